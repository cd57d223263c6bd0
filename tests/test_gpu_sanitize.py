"""Small-size targets that touch every kernel family once -- what scripts/sanitize.sh runs under compute-sanitizer
(memcheck / racecheck / synccheck, ONE tool per call).  They are ordinary parity tests too (each result is compared with the
oracle), so they also run in the normal `-m gpu` suite."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
P = 2**31 - 1


@pytest.fixture(scope="module")
def tb():
    import tstwo_b200
    return tstwo_b200


@pytest.mark.parametrize("log,blow,n_cols", [(10, 1, 3), (14, 1, 2), (15, 1, 2)])
def test_lde_commit_small(tb, orc, log, blow, n_cols):
    """per-layer kernels (log 10), bulk-copy contiguous pass alone (log 14 -> 15), contiguous + partial strided pass (15 -> 16);
    Merkle leaf (bulk-copy fed) and node layers."""
    rng = np.random.default_rng(log)
    trace = rng.integers(0, P, (n_cols, 1 << log), dtype=np.uint32)
    root_coset = tb.CanonicCoset.new(log + blow).half_coset()
    tw = tb.CudaBackend.precomputeTwiddles(root_coset)
    tree = tb.CommitmentTreeProver.from_trace(tb.CudaColumn.from_cpu(trace.reshape(-1)), n_cols, log, blow, tw)
    otw, oitw = orc.precompute_twiddles(root_coset.initial_index.value, root_coset.log_size)
    coeffs, evals, layers = orc.lde_commit(trace, blow, otw, oitw, want_coeffs=True)
    for c in range(n_cols):
        assert (tree.polynomials[c].coeffs.toCpu() == coeffs[c]).all()
        assert (tree.evaluations[c].values.toCpu() == evals[c]).all()
    assert tree.root() == layers[0][0].tobytes()


def test_full_strided_pass_with_tensor_core_stage(tb, orc):
    """log 22, one column: contiguous pass + FULL strided pass; the forward one runs its top stage on the tensor cores
    (tcgen05 / TMEM).  Checked against the oracle in full."""
    from tstwo_b200 import _lib
    L = _lib.init(0)
    log = 22
    rng = np.random.default_rng(22)
    vals = rng.integers(0, P, 1 << log, dtype=np.uint32)
    dom = tb.CanonicCoset.new(log).circleDomain()
    tw = tb.CudaBackend.precomputeTwiddles(dom.halfCoset)
    otw, oitw = orc.precompute_twiddles(dom.halfCoset.initial_index.value, dom.halfCoset.log_size)
    orc.use_all_cores()
    L.tstwo_debug_set_fft_tc(7)                       # every pass that can: forward + inverse strided, contiguous
    try:
        poly = tb.CudaCircleEvaluation.new(dom, vals).interpolateWithTwiddles(tw)
        coeffs = poly.coeffs.toCpu()
        back = poly.evaluateWithTwiddles(dom, tw).values.toCpu()
    finally:
        L.tstwo_debug_set_fft_tc(1)
    want = orc.interpolate(vals, dom.halfCoset.initial_index.value, (dom.halfCoset.initial_index.value, dom.halfCoset.log_size), oitw)
    assert (coeffs == want).all()
    assert (back == vals).all()


def test_fri_and_elementwise_small(tb, orc):
    rng = np.random.default_rng(5)
    log = 10
    dom = tb.CanonicCoset.new(log).circleDomain()
    src = [rng.integers(0, P, 1 << log, dtype=np.uint32) for _ in range(4)]
    alpha = [int(v) for v in rng.integers(1, P, 4)]
    tw = tb.CudaBackend.precomputeTwiddles(dom.halfCoset)
    layer = tb.LineEvaluation.newZero(tb.LineDomain.new(dom.halfCoset))
    tb.fold_circle_into_line(layer, tb.SecureEvaluation(dom, tb.SecureColumnByCoords.from_cpu(src)), alpha, tw)
    cur = orc.fold_circle_into_line([np.zeros(1 << (log - 1), dtype=np.uint32)] * 4, src, dom.halfCoset.initial_index.value, alpha)
    coset = dom.halfCoset
    while layer.len() > 2:
        layer = tb.fold_line(layer, alpha, None)
        cur = orc.fold_line(cur, coset.initial_index.value, alpha)
        coset = coset.double()
    for a, b in zip(layer.values.to_cpu(), cur):
        assert (a == b).all()
    g, lam = tb.CudaBackend.friOps.decompose(tb.SecureEvaluation(dom, tb.SecureColumnByCoords.from_cpu(src)))
    eg, elam = orc.fri_decompose(src)
    assert lam == elam and all((a == b).all() for a, b in zip(g.values.to_cpu(), eg))
    col = tb.CudaColumn.from_cpu(src[0])
    tb.CudaBackend.bitReverseColumn(col)
    assert (col.toCpu() == orc.bit_reverse(src[0])).all()
