"""The three families of tiled FFT passes -- bulk-copy/TMA (default), cp.async persistent, one CTA per tile -- must
produce bit-identical coefficients, evaluations and Merkle roots, equal to the oracle where the oracle is fast.
Mirrors core/test/backend/cpu/circle.test.ts:52-141 (interpolate/evaluate round trips at many sizes) through the
fused C-ABI entry point and through the PolyOps batch entry points."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
P = 2**31 - 1


@pytest.fixture(scope="module")
def tb():
    import tstwo_b200 as t
    return t


@pytest.fixture(scope="module")
def orc():
    import oracle
    return oracle


def _lde(tb, L, mode, trace, n_cols, log, blow, tw):
    L.tstwo_debug_set_fft_stream(mode)
    try:
        tree = tb.CommitmentTreeProver.from_trace(tb.CudaColumn.from_cpu(trace.reshape(-1)), n_cols, log, blow, tw)
        used = L.tstwo_debug_last_fft_mode()
        return ([p.coeffs.toCpu() for p in tree.polynomials], [e.values.toCpu() for e in tree.evaluations], tree.root(), used)
    finally:
        L.tstwo_debug_set_fft_stream(2)


@pytest.mark.parametrize("log,blow,n_cols", [(14, 0, 1), (14, 1, 3), (14, 2, 5), (15, 1, 2), (15, 0, 7), (16, 3, 3), (17, 2, 4),
                                             (18, 4, 2), (20, 1, 6), (21, 0, 3)])
def test_pass_families_agree(tb, orc, log, blow, n_cols):
    from tstwo_b200 import _lib
    L = _lib.init(0)
    ext = log + blow
    root_coset = tb.CanonicCoset.new(ext).half_coset()
    tw = tb.CudaBackend.precomputeTwiddles(root_coset)
    trace = np.random.default_rng(1000 * log + 10 * blow + n_cols).integers(0, P, (n_cols, 1 << log), dtype=np.uint32)
    res = {m: _lde(tb, L, m, trace, n_cols, log, blow, tw) for m in (2, 1, 0)}
    assert res[2][3] == 2 and res[1][3] == 1 and res[0][3] == 0, "the requested pass family did not run"
    for m in (1, 0):
        assert res[m][2] == res[2][2], f"Merkle root of family {m} differs from the bulk-copy passes"
        for c in range(n_cols):
            assert (res[m][0][c] == res[2][0][c]).all(), f"coefficients differ (family {m}, column {c})"
            assert (res[m][1][c] == res[2][1][c]).all(), f"evaluations differ (family {m}, column {c})"
    if log <= 17:   # and against the oracle
        otw, oitw = orc.precompute_twiddles(root_coset.initial_index.value, root_coset.log_size)
        coeffs, evals, layers = orc.lde_commit(trace, blow, otw, oitw, want_coeffs=True)
        for c in range(n_cols):
            assert (res[2][0][c] == coeffs[c]).all() and (res[2][1][c] == evals[c]).all()
        assert res[2][2] == layers[0][0].tobytes()


@pytest.mark.parametrize("log", [22, 23, 24])
def test_large_transforms_round_trip(tb, log):
    """Sizes with a full strided window (22), a partial third window (23, 24): bulk-copy passes vs cp.async passes on the
    device, and evaluate(interpolate(v)) == v."""
    from tstwo_b200 import _lib
    L = _lib.init(0)
    tw = tb.CudaBackend.precomputeTwiddles(tb.Coset.half_odds(log - 1))
    vals = np.random.default_rng(log).integers(0, P, 1 << log, dtype=np.uint32)
    dom = tb.CanonicCoset.new(log).circleDomain()
    res = []
    try:
        for mode in (2, 1):
            L.tstwo_debug_set_fft_stream(mode)
            poly = tb.CudaCircleEvaluation.new(dom, vals).interpolateWithTwiddles(tw)
            assert L.tstwo_debug_last_fft_mode() == mode
            back = poly.evaluateWithTwiddles(dom, tw)
            res.append((poly.coeffs.toCpu(), back.values.toCpu()))
    finally:
        L.tstwo_debug_set_fft_stream(2)
    assert (res[0][0] == res[1][0]).all() and (res[0][1] == res[1][1]).all() and (res[0][1] == vals).all()


def test_batch_ops_with_scattered_inputs_use_bulk_copies(tb, orc):
    """PolyOps.interpolateColumns / evaluatePolynomials on separately allocated columns: the contiguous pass copies column
    by column (any table), the outputs are one block, so the bulk-copy family still runs; results match the oracle."""
    from tstwo_b200 import _lib
    L = _lib.init(0)
    log, blow, n_cols = 15, 1, 5
    ext = log + blow
    root_coset = tb.CanonicCoset.new(ext).half_coset()
    tw = tb.CudaBackend.precomputeTwiddles(root_coset)
    otw, oitw = orc.precompute_twiddles(root_coset.initial_index.value, root_coset.log_size)
    root = (root_coset.initial_index.value, root_coset.log_size)
    vals = np.random.default_rng(7).integers(0, P, (n_cols, 1 << log), dtype=np.uint32)
    dom, big = tb.CanonicCoset.new(log).circleDomain(), tb.CanonicCoset.new(ext).circleDomain()
    pad = [tb.CudaColumn.uninitialized(4 * (k + 1)) for k in range(n_cols)]   # keeps the inputs unevenly spaced
    polys = tb.CudaBackend.interpolateColumns([tb.CudaCircleEvaluation.new(dom, v) for v in vals], tw)
    assert L.tstwo_debug_last_fft_mode() == 2
    out = tb.CudaBackend.evaluatePolynomials(polys, blow, tw)
    assert L.tstwo_debug_last_fft_mode() == 2
    del pad
    for c in range(n_cols):
        ec = orc.interpolate(vals[c], dom.halfCoset.initial_index.value, root, oitw)
        assert (polys[c].coeffs.toCpu() == ec).all()
        assert (out[c].values.toCpu() == orc.evaluate(ec, ext, big.halfCoset.initial_index.value, root, otw)).all()


def test_partial_top_pass_repeats_exactly(tb):
    """Regression: the 2-layer partial pass of a 2^24 transform finishes a tile faster than a bulk copy lands, so a group
    could reach its next wait on a buffer one phase early and fall through on the barrier's parity (stale tile -> every
    output wrong, about one run in eight).  Buffers now alternate between two barriers; repeat the transform and demand
    the identical, correct result every time in both persistent families."""
    from tstwo_b200 import _lib
    L = _lib.init(0)
    log = 24
    tw = tb.CudaBackend.precomputeTwiddles(tb.Coset.half_odds(log - 1))
    vals = np.random.default_rng(99).integers(0, P, 1 << log, dtype=np.uint32)
    dom = tb.CanonicCoset.new(log).circleDomain()
    try:
        for mode in (2, 1):
            L.tstwo_debug_set_fft_stream(mode)
            poly = tb.CudaCircleEvaluation.new(dom, vals).interpolateWithTwiddles(tw)
            for _ in range(12):
                back = poly.evaluateWithTwiddles(dom, tw)
                assert L.tstwo_debug_last_fft_mode() == mode
                assert (back.values.toCpu() == vals).all()
    finally:
        L.tstwo_debug_set_fft_stream(2)


@pytest.mark.parametrize("blow,n_cols", [(0, 1), (1, 3), (2, 2)])
def test_nine_layer_window_log23(tb, orc, blow, n_cols):
    """Log size 23 leaves exactly nine layers above the 14-layer contiguous pass: the bulk-copy family then runs ONE 512-row
    strided pass (fft_tma_kernel<.., R9>) instead of an 8-layer window plus a one-layer pass.  Coefficients, evaluations (also with
    blow-up, where the pass reads the zero-extended coefficients coset by coset) and the Merkle root must equal what the planner
    without that pass gives (tstwo_debug_set_fft_rows9(0)), and column 0 must equal the oracle's transform."""
    from tstwo_b200 import _lib
    L = _lib.init(0)
    log = 23
    ext = log + blow
    root_coset = tb.CanonicCoset.new(ext).half_coset()
    tw = tb.CudaBackend.precomputeTwiddles(root_coset)
    trace = np.random.default_rng(2300 + 10 * blow + n_cols).integers(0, P, (n_cols, 1 << log), dtype=np.uint32)
    res = {}
    try:
        for on in (1, 0):
            L.tstwo_debug_set_fft_rows9(on)
            res[on] = _lde(tb, L, 2, trace, n_cols, log, blow, tw)
            assert res[on][3] == 2
    finally:
        L.tstwo_debug_set_fft_rows9(1)
    assert res[1][2] == res[0][2], "Merkle root differs with the 512-row pass"
    for c in range(n_cols):
        assert (res[1][0][c] == res[0][0][c]).all(), f"coefficients differ (column {c})"
        assert (res[1][1][c] == res[0][1][c]).all(), f"evaluations differ (column {c})"
    otw, oitw = orc.precompute_twiddles(root_coset.initial_index.value, root_coset.log_size)
    orc.use_all_cores()
    coeffs, evals, _ = orc.lde_commit(trace[:1], blow, otw, oitw, want_coeffs=True, want_layers=False)
    assert (res[1][0][0] == coeffs[0]).all() and (res[1][1][0] == evals[0]).all()
