"""GPU: tstwo_widefib_constraints (both kernel forms) against the oracle, and the accumulator classes of tstwo_b200/air.py on
the device: DomainEvaluationAccumulator.finalize against the same steps through the oracle, the wide-Fibonacci component end to
end (trace -> interpolate -> constraint quotients -> finalize) with the degree property of a valid trace."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
P = 2147483647


def rq(rng): return [int(x) for x in rng.integers(0, P, 4)]


@pytest.fixture(scope="module")
def tb():
    import tstwo_b200
    return tstwo_b200


@pytest.mark.parametrize("tlog,elog,n_cols,mode", [(3, 4, 3, 0), (4, 5, 10, 0), (8, 9, 100, 1), (8, 9, 100, 2), (9, 11, 257, 2),
                                                   (9, 11, 257, 1), (10, 10, 8, 2), (9, 10, 2, 0)])
def test_widefib_constraints_kernels(tb, orc, tlog, elog, n_cols, mode):
    from tstwo_b200 import _lib, air
    L = _lib.load()
    rng = np.random.default_rng(tlog * 100 + n_cols)
    cols = [rng.integers(0, P, 1 << elog, dtype=np.uint32) for _ in range(n_cols)]
    cols[0][:] = P - 1                                   # extreme bytes / squares
    coeffs = [rq(rng) for _ in range(n_cols - 2)]
    den = [int(x) for x in rng.integers(0, P, 1 << (elog - tlog))]
    start = [rng.integers(0, P, 1 << elog, dtype=np.uint32) for _ in range(4)]
    out = tb.SecureColumnByCoords.from_cpu(start)
    L.tstwo_debug_set_quot_mode(mode)
    try:
        air.widefib_constraints(tlog, elog, [tb.CudaColumn.from_cpu(c) for c in cols], coeffs, den, out)
        if mode and n_cols > 2:
            assert L.tstwo_debug_last_quot_mode() == mode
    finally:
        L.tstwo_debug_set_quot_mode(0)
    exp = orc.widefib_constraints(tlog, elog, cols, coeffs, den, start)
    for a, b in zip(out.to_cpu(), exp):
        assert (a == b).all()


def test_widefib_constraints_many_columns_flush(tb, orc):
    """More columns than the s32 accumulators may take between flushes (2048), all at the largest values."""
    from tstwo_b200 import _lib, air
    rng = np.random.default_rng(1)
    tlog, elog, n_cols = 8, 9, 2300
    base = [np.full(1 << elog, P - 1, dtype=np.uint32), rng.integers(0, P, 1 << elog, dtype=np.uint32)]
    cols = [base[k % 2] for k in range(n_cols)]
    dcols = [tb.CudaColumn.from_cpu(c) for c in base]
    coeffs = [[P - 1, P - 1, P - 1, P - 1] for _ in range(n_cols - 2)]
    den = [5, 7]
    out = tb.SecureColumnByCoords.zeros(1 << elog)
    air.widefib_constraints(tlog, elog, [dcols[k % 2] for k in range(n_cols)], coeffs, den, out)
    assert _lib.load().tstwo_debug_last_quot_mode() == 2
    exp = orc.widefib_constraints(tlog, elog, cols, coeffs, den, [np.zeros(1 << elog, dtype=np.uint32)] * 4)
    for a, b in zip(out.to_cpu(), exp):
        assert (a == b).all()


def test_widefib_bad_arguments(tb):
    from tstwo_b200 import air
    out = tb.SecureColumnByCoords.zeros(16)
    cols = [tb.CudaColumn.zeros(16) for _ in range(3)]
    with pytest.raises(tb.TstwoError, match="BAD_ARG"):
        air.widefib_constraints(5, 4, cols, [[1, 0, 0, 0]], [1], out)            # eval domain smaller than the trace
    with pytest.raises(tb.TstwoError, match="out of range"):
        air.widefib_constraints(3, 4, cols, [[P, 0, 0, 0]], [1, 1], out)         # not a field element


def test_domain_evaluation_accumulator_finalize(tb, orc):
    """Two sizes: the size-5 sub-accumulation is interpolated, lifted to size 7, added, interpolated (accumulator.ts:189-238)."""
    from tstwo_b200 import air
    rng = np.random.default_rng(3)
    alpha = rq(rng)
    acc = air.DomainEvaluationAccumulator.new(alpha, 7, 5)
    assert acc.logSize() == 7
    with pytest.raises(ValueError):
        acc.columns([(5, 1), (5, 1)])
    with pytest.raises(ValueError):
        acc.columns([(9, 1)])
    c7, c5 = acc.columns([(7, 3), (5, 2)])
    pows = air.generate_secure_powers(alpha, 5)
    assert c7.randomCoeffPowers == pows[2:] and c5.randomCoeffPowers == pows[:2]
    with pytest.raises(ValueError):
        acc.columns([(6, 1)])                                   # no coefficients left
    v7 = [rng.integers(0, P, 1 << 7, dtype=np.uint32) for _ in range(4)]
    v5 = [rng.integers(0, P, 1 << 5, dtype=np.uint32) for _ in range(4)]
    c7.accumulate_column(tb.SecureColumnByCoords.from_cpu(v7))
    c5.accumulate_column(tb.SecureColumnByCoords.from_cpu(v5))
    extra = rq(rng)
    c5.accumulate(3, extra)                                     # the per-row form of the reference
    with pytest.raises(IndexError):
        c5.accumulate(32, extra)
    for k in range(4):
        v5[k][3] = (int(v5[k][3]) + extra[k]) % P
    poly = acc.finalize()
    assert poly.log_size() == 7
    # the same through the oracle
    h7, l7 = orc.canonic_half_coset(7)
    h5, l5 = orc.canonic_half_coset(5)
    tw, itw = orc.precompute_twiddles(h7, l7)
    root = (h7, l7)
    for k in range(4):
        p5 = orc.interpolate(v5[k], h5, root, itw, quirks=False)
        lifted = orc.evaluate(p5, 7, h7, root, tw, quirks=False)
        exp = orc.interpolate(((v7[k].astype(np.uint64) + lifted) % P).astype(np.uint32), h7, root, itw, quirks=False)
        assert (poly.polys[k].coeffs.toCpu() == exp).all()
    # eval_at_point combines the 4 coordinate evaluations
    pt = ((1, 0, 478637715, 513582971), (992285211, 649143431, 740191619, 1186584352))
    e = [orc.eval_at_point(poly.polys[k].coeffs.toCpu(), pt[0], pt[1]) for k in range(4)]
    got = poly.eval_at_point(pt)
    units = ([1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, 0], [0, 0, 0, 1])
    exp_pt = [0, 0, 0, 0]
    for ek, u in zip(e, units):
        exp_pt = air.qm31_add(exp_pt, air.qm31_mul(ek, u))
    assert got == exp_pt
    # not all coefficients used -> finalize throws
    acc2 = air.DomainEvaluationAccumulator.new(alpha, 4, 2)
    with pytest.raises(ValueError):
        acc2.finalize()
    # nothing accumulated -> zero polynomial
    z = air.DomainEvaluationAccumulator.new(alpha, 4, 0).finalize()
    assert z.log_size() == 4 and not any(p.coeffs.toCpu().any() for p in z.polys)


@pytest.mark.parametrize("log,n_cols", [(6, 5), (10, 64)])
def test_wide_fibonacci_component(tb, orc, log, n_cols):
    from tstwo_b200 import air
    rng = np.random.default_rng(log)
    trace = air.generate_wide_fibonacci_trace(log, n_cols)
    dom = tb.CanonicCoset.new(log).circleDomain()
    elog = log + 1
    edom = tb.CanonicCoset.new(elog).circleDomain()
    tw = tb.CudaBackend.precomputeTwiddles(edom.halfCoset)
    polys = tb.CudaBackend.interpolateColumns([tb.CudaCircleEvaluation.new(dom, c) for c in trace], tw)
    comp = air.WideFibonacciComponent(air.WideFibonacciEval.create(log, n_cols))
    alpha = rq(rng)
    acc = air.DomainEvaluationAccumulator.new(alpha, elog, comp.nConstraints())
    comp.evaluateConstraintQuotientsOnDomain(polys, acc, tw)
    quot = [c.toCpu() for c in acc.subAccumulations[elog].columns]
    # oracle: same evaluations, coefficient of constraint i = alpha^(N-1-i)
    evals = [e.values.toCpu() for e in tb.CudaBackend.evaluatePolynomialsOnDomain(polys, edom, tw)]
    pows = orc.generate_secure_powers(alpha, n_cols - 2)
    coeffs = [list(map(int, pows[n_cols - 3 - i])) for i in range(n_cols - 2)]
    exp = orc.widefib_constraints(log, elog, evals, coeffs, orc.widefib_denom_inv(log, elog), [np.zeros(1 << elog, dtype=np.uint32)] * 4)
    for a, b in zip(quot, exp):
        assert (a == b).all()
    # composition polynomial of a valid trace: only the first 2^log + 1 basis functions (tests/test_oracle_air.py)
    poly = acc.finalize()
    for p in poly.polys:
        co = p.coeffs.toCpu()
        assert co[:1 << log].any() and not co[(1 << log) + 1:].any()
    # an invalid trace does not give such a polynomial
    bad = trace.copy()
    bad[3][5] = (int(bad[3][5]) + 1) % P
    polys = tb.CudaBackend.interpolateColumns([tb.CudaCircleEvaluation.new(dom, c) for c in bad], tw)
    acc = air.DomainEvaluationAccumulator.new(alpha, elog, comp.nConstraints())
    comp.evaluateConstraintQuotientsOnDomain(polys, acc, tw)
    assert acc.finalize().polys[0].coeffs.toCpu()[(1 << log) + 1:].any()
