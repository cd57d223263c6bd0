"""Parity of the CUDA FRI folds / decompose / accumulation / quotients with the oracle.  Mirrors
packages/core/test/fri.test.ts:206-267, test/backend/cpu/fri.test.ts:44-183, accumulation.test.ts, quotients.test.ts."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
P = 2147483647


@pytest.fixture(scope="module")
def tb():
    import tstwo_b200
    return tstwo_b200


def rq(rng):
    return [int(v) for v in rng.integers(1, P, 4)]


@pytest.mark.parametrize("with_tree", [False, True])
@pytest.mark.parametrize("log", [1, 2, 3, 6, 11, 16])
def test_fold_line(tb, orc, log, with_tree):
    rng = np.random.default_rng(log)
    coset = tb.Coset.half_odds(log)
    src = [rng.integers(0, P, 1 << log, dtype=np.uint32) for _ in range(4)]
    alpha = rq(rng)
    tw = tb.CudaBackend.precomputeTwiddles(tb.Coset.half_odds(log + 2)) if with_tree else None
    ev = tb.LineEvaluation.new(tb.LineDomain.new(coset), tb.SecureColumnByCoords.from_cpu(src))
    out = tb.fold_line(ev, alpha, tw)
    exp = orc.fold_line(src, coset.initial_index.value, alpha)
    assert out.domain() == tb.LineDomain.new(coset.double()) if log > 0 else True
    for a, b in zip(out.values.to_cpu(), exp):
        assert (a == b).all()


def test_fold_line_too_small(tb):
    ev = tb.LineEvaluation.new(tb.LineDomain.new(tb.Coset.half_odds(0)), tb.SecureColumnByCoords.zeros(1))
    with pytest.raises(ValueError, match="Evaluation too small"):
        tb.fold_line(ev, [1, 0, 0, 0])


@pytest.mark.parametrize("with_tree", [False, True])
@pytest.mark.parametrize("log", [1, 2, 3, 4, 7, 12, 16])
def test_fold_circle_into_line(tb, orc, log, with_tree):
    rng = np.random.default_rng(50 + log)
    dom = tb.CanonicCoset.new(log).circleDomain()
    src = [rng.integers(0, P, 1 << log, dtype=np.uint32) for _ in range(4)]
    dst = [rng.integers(0, P, 1 << (log - 1), dtype=np.uint32) for _ in range(4)]
    alpha = rq(rng)
    tw = tb.CudaBackend.precomputeTwiddles(tb.Coset.half_odds(log + 1)) if with_tree else None
    d = tb.LineEvaluation.new(tb.LineDomain.new(dom.halfCoset), tb.SecureColumnByCoords.from_cpu(dst))
    s = tb.SecureEvaluation(dom, tb.SecureColumnByCoords.from_cpu(src))
    tb.fold_circle_into_line(d, s, alpha, tw)
    exp = orc.fold_circle_into_line(dst, src, dom.halfCoset.initial_index.value, alpha)
    for a, b in zip(d.values.to_cpu(), exp):
        assert (a == b).all()


def test_fold_circle_length_mismatch(tb):
    dom = tb.CanonicCoset.new(4).circleDomain()
    s = tb.SecureEvaluation(dom, tb.SecureColumnByCoords.zeros(16))
    d = tb.LineEvaluation.new(tb.LineDomain.new(tb.Coset.half_odds(2)), tb.SecureColumnByCoords.zeros(4))
    with pytest.raises(ValueError, match="Length mismatch"):
        tb.fold_circle_into_line(d, s, [1, 0, 0, 0])


def test_fri_layer_chain(tb, orc):
    """FriProver.commitInnerLayers data flow (fri.ts:676-716): circle fold then line folds down to 2 elements."""
    rng = np.random.default_rng(77)
    log = 12
    dom = tb.CanonicCoset.new(log).circleDomain()
    tw = tb.CudaBackend.precomputeTwiddles(dom.halfCoset)
    src = [rng.integers(0, P, 1 << log, dtype=np.uint32) for _ in range(4)]
    layer = tb.LineEvaluation.newZero(tb.LineDomain.new(dom.halfCoset))
    alpha = rq(rng)
    tb.fold_circle_into_line(layer, tb.SecureEvaluation(dom, tb.SecureColumnByCoords.from_cpu(src)), alpha, tw)
    cur = orc.fold_circle_into_line([np.zeros(1 << (log - 1), dtype=np.uint32)] * 4, src, dom.halfCoset.initial_index.value, alpha)
    coset = dom.halfCoset
    while layer.len() > 2:
        alpha = rq(rng)
        layer = tb.fold_line(layer, alpha, tw)
        cur = orc.fold_line(cur, coset.initial_index.value, alpha)
        coset = coset.double()
        for a, b in zip(layer.values.to_cpu(), cur):
            assert (a == b).all()


@pytest.mark.parametrize("log", [0, 1, 5, 11, 15])
def test_decompose(tb, orc, log):
    rng = np.random.default_rng(log)
    dom = tb.CircleDomain.new(tb.Coset.half_odds(log - 1)) if log > 0 else None
    src = [rng.integers(0, P, 1 << log, dtype=np.uint32) for _ in range(4)]
    if log == 0:
        class D:   # a size-1 "domain" as in the reference's MockCircleDomain (test/backend/cpu/fri.test.ts:23-42)
            def size(self): return 1
            def logSize(self): return 0
        dom = D()
    g, lam = tb.CudaBackend.friOps.decompose(tb.SecureEvaluation(dom, tb.SecureColumnByCoords.from_cpu(src)))
    eg, elam = orc.fri_decompose(src)
    assert lam == elam
    for a, b in zip(g.values.to_cpu(), eg):
        assert (a == b).all()


def test_accumulate(tb, orc):
    rng = np.random.default_rng(3)
    a = [rng.integers(0, P, 1000, dtype=np.uint32) for _ in range(4)]
    b = [rng.integers(0, P, 1000, dtype=np.uint32) for _ in range(4)]
    da, db = tb.SecureColumnByCoords.from_cpu(a), tb.SecureColumnByCoords.from_cpu(b)
    tb.accumulate(da, db)
    for x, y in zip(da.to_cpu(), orc.accumulate(a, b)):
        assert (x == y).all()
    with pytest.raises(ValueError, match="column length mismatch"):
        tb.accumulate(da, tb.SecureColumnByCoords.zeros(10))


@pytest.mark.parametrize("log", [3, 8, 13])
def test_accumulate_quotients(tb, orc, log):
    rng = np.random.default_rng(log)
    dom = tb.CanonicCoset.new(log).circleDomain()
    cols = [rng.integers(0, P, 1 << log, dtype=np.uint32) for _ in range(5)]
    rc = rq(rng)
    batches = [((rq(rng), rq(rng)), [(0, rq(rng)), (3, rq(rng)), (4, rq(rng))]), ((rq(rng), rq(rng)), [(1, rq(rng))]),
               ((rq(rng), rq(rng)), [(2, rq(rng)), (0, rq(rng))])]
    evs = [tb.CudaCircleEvaluation.new(dom, c) for c in cols]
    out = tb.accumulate_quotients(dom, evs, rc, batches)
    exp = orc.accumulate_quotients(log, dom.halfCoset.initial_index.value, cols, rc, batches)
    for a, b in zip(out.values.to_cpu(), exp):
        assert (a == b).all()


@pytest.mark.parametrize("mode", [1, 2])
@pytest.mark.parametrize("ts", [True, False])
def test_accumulate_quotients_both_kernels(tb, orc, mode, ts):
    """The integer-pipe kernel (mode 1) and the INT8 tensor-core kernel (mode 2) against the oracle: ragged batches (1, 7, 8, 9
    columns), a batch longer than one shared-memory chunk (300) and repeated columns."""
    from tstwo_b200 import _lib
    L = _lib.load()
    log = 10
    rng = np.random.default_rng(50 + mode)
    dom = tb.CanonicCoset.new(log).circleDomain()
    cols = [rng.integers(0, P, 1 << log, dtype=np.uint32) for _ in range(12)]
    cols[3][:] = P - 1          # extreme bytes: 0x7ffffffe in every row
    cols[4][:] = 0
    rc = rq(rng)
    batches = [((rq(rng), rq(rng)), [(int(c), rq(rng)) for c in rng.integers(0, 12, k)]) for k in (1, 7, 8, 9, 300)]
    batches[1][1][0] = (3, [P - 1, P - 1, P - 1, P - 1])
    evs = [tb.CudaCircleEvaluation.new(dom, c) for c in cols]
    L.tstwo_debug_set_quot_mode(mode)
    try:
        out = tb.accumulate_quotients(dom, evs, rc, batches, ts_denominators=ts)
        assert L.tstwo_debug_last_quot_mode() == mode
    finally:
        L.tstwo_debug_set_quot_mode(0)
    exp = orc.accumulate_quotients(log, dom.halfCoset.initial_index.value, cols, rc, batches, ts_denominators=ts)
    for a, b in zip(out.values.to_cpu(), exp):
        assert (a == b).all()


def test_accumulate_quotients_tensor_core_flush(tb, orc):
    """More columns in one batch than the s32 accumulators of the tensor-core kernel may hold (flush every 4096), all at the
    largest byte values."""
    from tstwo_b200 import _lib
    L = _lib.load()
    log = 9
    rng = np.random.default_rng(77)
    dom = tb.CanonicCoset.new(log).circleDomain()
    cols = [np.full(1 << log, 0x7f7fffff, dtype=np.uint32), rng.integers(0, P, 1 << log, dtype=np.uint32)]
    # c = alpha^k (conj(py) - py): not controllable per byte, so take many terms instead: 9000 > 2 * 4096
    batches = [((rq(rng), rq(rng)), [(int(k % 2 == 0 and 0 or rng.integers(0, 2)), rq(rng)) for k in range(9000)])]
    evs = [tb.CudaCircleEvaluation.new(dom, c) for c in cols]
    rc = rq(rng)
    out = tb.accumulate_quotients(dom, evs, rc, batches)
    assert L.tstwo_debug_last_quot_mode() == 2
    exp = orc.accumulate_quotients(log, dom.halfCoset.initial_index.value, cols, rc, batches)
    for a, b in zip(out.values.to_cpu(), exp):
        assert (a == b).all()
