"""Parity of the CUDA LDE path (through the C ABI / CudaBackend mirror) with the oracle, bit-exact.
Mirrors packages/core/test/backend/cpu/circle.test.ts (evaluate / interpolate / round trip) plus differential
tests at every log size, ragged column counts, scattered columns, blow-ups and the log-3 quirk."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

P = 2147483647


@pytest.fixture(scope="module")
def tb():
    import tstwo_b200
    return tstwo_b200


def _tree(tb, orc, log_half, init=None):
    coset = tb.Coset.half_odds(log_half) if init is None else tb.Coset.new(tb.CirclePointIndex(init), log_half)
    tw = tb.CudaBackend.precomputeTwiddles(coset)
    otw, oitw = orc.precompute_twiddles(coset.initial_index.value, coset.log_size)
    return coset, tw, otw, oitw


@pytest.mark.parametrize("log", list(range(0, 14)) + [17])
def test_twiddles(tb, orc, log):
    coset, tw, otw, oitw = _tree(tb, orc, log)
    assert (tw.twiddles.toCpu() == otw).all()
    assert (tw.itwiddles.toCpu() == oitw).all()


def test_twiddles_non_canonic(tb, orc):
    coset, tw, otw, oitw = _tree(tb, orc, 9, init=12345 * 2 + 1)
    assert (tw.twiddles.toCpu() == otw).all() and (tw.itwiddles.toCpu() == oitw).all()


def test_bit_reverse(tb, orc):
    rng = np.random.default_rng(0)
    for log in [0, 1, 5, 11, 12, 13, 14, 17, 20]:      # 12 and up: the tiled kernel (odd and even numbers of middle bits, palindromic tiles)
        a = rng.integers(0, P, 1 << log, dtype=np.uint32)
        col = tb.CudaColumn.from_cpu(a)
        tb.CudaBackend.bitReverseColumn(col)
        assert (col.toCpu() == orc.bit_reverse(a)).all()
        tb.CudaBackend.bitReverseColumn(col)                 # an involution
        assert (col.toCpu() == a).all()
    with pytest.raises(ValueError):
        tb.CudaBackend.bitReverseColumn(tb.CudaColumn.from_cpu(np.zeros(6, dtype=np.uint32)))


@pytest.mark.parametrize("quirks", [True, False])
@pytest.mark.parametrize("log", list(range(1, 17)))
def test_evaluate_interpolate_vs_oracle(tb, orc, log, quirks):
    tb.CudaBackend.quirks = quirks
    try:
        dom = tb.CanonicCoset.new(log).circleDomain()
        coset, tw, otw, oitw = _tree(tb, orc, log - 1)
        rng = np.random.default_rng(log)
        n_cols = 3 if log < 14 else 9
        coeffs = rng.integers(0, P, (n_cols, 1 << log), dtype=np.uint32)
        polys = [tb.CudaCirclePoly.new(c) for c in coeffs]       # separately allocated -> scattered column table
        evs = tb.CudaBackend.evaluatePolynomialsOnDomain(polys, dom, tw)
        root = (coset.initial_index.value, coset.log_size)
        for c in range(n_cols):
            exp = orc.evaluate(coeffs[c], log, dom.halfCoset.initial_index.value, root, otw, quirks=quirks)
            assert (evs[c].values.toCpu() == exp).all(), (log, c)
        back = tb.CudaBackend.interpolateColumns(evs, tw)
        for c in range(n_cols):
            assert (back[c].coeffs.toCpu() == coeffs[c]).all(), (log, c)
        # interpolate of arbitrary values vs oracle
        vals = rng.integers(0, P, 1 << log, dtype=np.uint32)
        got = tb.CudaCircleEvaluation.new(dom, vals).interpolateWithTwiddles(tw).coeffs.toCpu()
        assert (got == orc.interpolate(vals, dom.halfCoset.initial_index.value, root, oitw, quirks=quirks)).all()
    finally:
        tb.CudaBackend.quirks = True


@pytest.mark.parametrize("log,blow,extra,n_cols", [(5, 2, 0, 3), (10, 1, 1, 2), (14, 1, 0, 8), (14, 2, 1, 5), (15, 1, 0, 17),
                                                   (16, 2, 0, 8), (18, 1, 0, 4)])
def test_blowup_contiguous_block(tb, orc, log, blow, extra, n_cols):
    """evaluatePolynomials (poly/circle/ops.ts:89-101) on a contiguous column block with a tree larger than needed."""
    ext = log + blow
    coset, tw, otw, oitw = _tree(tb, orc, ext - 1 + extra)
    root = (coset.initial_index.value, coset.log_size)
    rng = np.random.default_rng(100 + log)
    vals = rng.integers(0, P, (n_cols, 1 << log), dtype=np.uint32)
    block = tb.CudaColumn.from_cpu(vals.reshape(-1))
    dom = tb.CanonicCoset.new(log).circleDomain()
    evs = [tb.CudaCircleEvaluation(dom, block.slice(c << log, 1 << log)) for c in range(n_cols)]
    polys = tb.CudaBackend.interpolateColumns(evs, tw)
    big = tb.CanonicCoset.new(ext).circleDomain()
    out = tb.CudaBackend.evaluatePolynomials(polys, blow, tw)
    for c in range(n_cols):
        ec = orc.interpolate(vals[c], dom.halfCoset.initial_index.value, root, oitw)
        assert (polys[c].coeffs.toCpu() == ec).all()
        ee = orc.evaluate(ec, ext, big.halfCoset.initial_index.value, root, otw)
        assert (out[c].values.toCpu() == ee).all()


@pytest.mark.parametrize("radix", [8, 16])
@pytest.mark.parametrize("log,blow", [(14, 1), (15, 0), (17, 2), (21, 1)])
def test_both_tile_radixes_vs_oracle(tb, orc, radix, log, blow):
    """The tiled passes exist in two register blockings (8 / 16 rows per thread); both must match the oracle, including
    sizes where the strided window is only partially active (log < 22) and a 3-pass size is covered on device below."""
    from tstwo_b200 import _lib
    L = _lib.init(0)
    L.tstwo_debug_set_fft_radix(radix)
    try:
        ext = log + blow
        coset, tw, otw, oitw = _tree(tb, orc, ext - 1)
        root = (coset.initial_index.value, coset.log_size)
        rng = np.random.default_rng(radix * 100 + log)
        n_cols = 5 if log < 20 else 2
        vals = rng.integers(0, P, (n_cols, 1 << log), dtype=np.uint32)
        dom = tb.CanonicCoset.new(log).circleDomain()
        polys = tb.CudaBackend.interpolateColumns([tb.CudaCircleEvaluation.new(dom, v) for v in vals], tw)
        out = tb.CudaBackend.evaluatePolynomials(polys, blow, tw)
        big = tb.CanonicCoset.new(ext).circleDomain()
        for c in (0, n_cols - 1):
            ec = orc.interpolate(vals[c], dom.halfCoset.initial_index.value, root, oitw)
            assert (polys[c].coeffs.toCpu() == ec).all()
            assert (out[c].values.toCpu() == orc.evaluate(ec, ext, big.halfCoset.initial_index.value, root, otw)).all()
    finally:
        L.tstwo_debug_set_fft_radix(16)


@pytest.mark.parametrize("radix", [8, 16])
def test_three_pass_transform_on_device(tb, orc, radix):
    """log 23 needs three passes (12 + 10 + partial window); cross-checked against the per-layer kernels."""
    from tstwo_b200 import _lib
    L = _lib.init(0)
    L.tstwo_debug_set_fft_radix(radix)
    log = 23
    tw = tb.CudaBackend.precomputeTwiddles(tb.Coset.half_odds(log - 1))
    vals = np.random.default_rng(23).integers(0, P, 1 << log, dtype=np.uint32)
    dom = tb.CanonicCoset.new(log).circleDomain()
    res = []
    try:
        for naive in (0, 1):
            L.tstwo_debug_force_naive_fft(naive)
            poly = tb.CudaCircleEvaluation.new(dom, vals).interpolateWithTwiddles(tw)
            back = poly.evaluateWithTwiddles(dom, tw)
            res.append((poly.coeffs.toCpu(), back.values.toCpu()))
    finally:
        L.tstwo_debug_force_naive_fft(0)
        L.tstwo_debug_set_fft_radix(16)
    assert (res[0][0] == res[1][0]).all() and (res[0][1] == res[1][1]).all() and (res[0][1] == vals).all()


@pytest.mark.parametrize("log", [14, 19, 20])
def test_tiled_passes_match_layer_kernels_on_device(tb, orc, log):
    """At sizes where the oracle is slow the tiled radix-2^11 passes are cross-checked against the per-layer kernels."""
    from tstwo_b200 import _lib
    L = _lib.init(0)
    coset, tw, _, _ = _tree(tb, orc, log) if log < 18 else (None, tb.CudaBackend.precomputeTwiddles(tb.Coset.half_odds(log)), None, None)
    rng = np.random.default_rng(log)
    n_cols = 3
    vals = rng.integers(0, P, (n_cols, 1 << log), dtype=np.uint32)
    dom = tb.CanonicCoset.new(log).circleDomain()
    res = []
    for naive in (0, 1):
        L.tstwo_debug_force_naive_fft(naive)
        try:
            evs = [tb.CudaCircleEvaluation.new(dom, v) for v in vals]
            polys = tb.CudaBackend.interpolateColumns(evs, tw)
            out = tb.CudaBackend.evaluatePolynomials(polys, 1, tw)
            res.append(([p.coeffs.toCpu() for p in polys], [o.values.toCpu() for o in out]))
        finally:
            L.tstwo_debug_force_naive_fft(0)
    for c in range(n_cols):
        assert (res[0][0][c] == res[1][0][c]).all()
        assert (res[0][1][c] == res[1][1][c]).all()
    if log <= 19:   # and one column against the oracle
        otw, oitw = orc.precompute_twiddles(tw.rootCoset.initial_index.value, tw.rootCoset.log_size)
        root = (tw.rootCoset.initial_index.value, tw.rootCoset.log_size)
        assert (res[0][0][0] == orc.interpolate(vals[0], dom.halfCoset.initial_index.value, root, oitw)).all()


def test_twiddle_tree_mismatch(tb, orc):
    # backend/cpu/circle.ts:89-91,140-142
    tw = tb.CudaBackend.precomputeTwiddles(tb.Coset.half_odds(5))
    dom = tb.CanonicCoset.new(8).circleDomain()
    with pytest.raises(tb.TwiddleMismatch, match="twiddle tree mismatch"):
        tb.CudaCirclePoly.new(np.ones(256, dtype=np.uint32)).evaluateWithTwiddles(dom, tw)
    bad = tb.CircleDomain.new(tb.Coset.new(tb.CirclePointIndex(7), 4))
    with pytest.raises(tb.TwiddleMismatch):
        tb.CudaCircleEvaluation.new(bad, np.ones(32, dtype=np.uint32)).interpolateWithTwiddles(tw)
    with pytest.raises(ValueError):
        tb.CudaCircleEvaluation.new(dom, np.ones(8, dtype=np.uint32))      # evaluation.ts:106-108


@pytest.mark.parametrize("log", [0, 1, 2, 3, 4, 5, 9, 13])
def test_eval_at_point(tb, orc, log):
    rng = np.random.default_rng(log)
    coeffs = rng.integers(0, P, 1 << log, dtype=np.uint32)
    px = [int(v) for v in rng.integers(0, P, 4)]; py = [int(v) for v in rng.integers(0, P, 4)]
    got = tb.CudaCirclePoly.eval_at_point(tb.CudaCirclePoly.new(coeffs), (px, py))
    assert got == orc.eval_at_point(coeffs, px, py)


def test_eval_many_at_point(tb, orc):
    """Batched form: mixed sizes (grouped per size), more polynomials than one pass holds at the largest size."""
    rng = np.random.default_rng(77)
    logs = [0, 3, 3, 7, 12, 12, 12, 5, 16, 16, 1]
    coeffs = [rng.integers(0, P, 1 << l, dtype=np.uint32) for l in logs]
    px = [int(v) for v in rng.integers(0, P, 4)]; py = [int(v) for v in rng.integers(0, P, 4)]
    got = tb.CudaCirclePoly.eval_many_at_point([tb.CudaCirclePoly.new(c) for c in coeffs], (px, py))
    assert got == [orc.eval_at_point(c, px, py) for c in coeffs]
    assert tb.CudaCirclePoly.eval_many_at_point([], (px, py)) == []
    big = [tb.CudaCirclePoly.new(rng.integers(0, P, 1 << 22, dtype=np.uint32)) for _ in range(50)]     # 6 MiB of scratch each, 256 MiB per pass: 2 passes
    got = tb.CudaCirclePoly.eval_many_at_point(big, (px, py))
    assert got == [tb.CudaCirclePoly.eval_at_point(p, (px, py)) for p in big]
    assert got[7] == orc.eval_at_point(big[7].coeffs.toCpu(), px, py)


def test_column_interface(tb):
    col = tb.CudaBackend.createBaseFieldColumn([1, 2, 3, 4])
    assert col.len() == 4 and not col.isEmpty() and col.at(2) == 3
    col.set(2, 77)
    assert col.toCpu().tolist() == [1, 2, 77, 4]
    with pytest.raises(IndexError):
        col.at(4)
    with pytest.raises(ValueError):
        tb.CudaBackend.createBaseFieldColumn([P])
    z = tb.CudaColumn.zeros(5)
    assert z.toCpu().tolist() == [0] * 5
