"""CPU: the oracle's wide-Fibonacci constraint quotients (oracle.widefib_constraints) against python big-int arithmetic and
against the property that defines them (a valid trace gives a quotient of degree < 2^log), and the host-side accumulator
classes of tstwo_b200/air.py against the oracle / direct sums.  No reference golden exists for these (parity unpinned)."""
import numpy as np
import pytest

import pyfield as pf

P = 2147483647


def rq(rng): return [int(x) for x in rng.integers(0, P, 4)]


def fib_trace(log, n_cols):
    n = 1 << log
    cols = [np.ones(n, dtype=np.uint64), np.arange(n, dtype=np.uint64)]
    for k in range(2, n_cols):
        cols.append((cols[k - 2] ** 2 % P + cols[k - 1] ** 2 % P) % P)
    return [c.astype(np.uint32) for c in cols]


def test_widefib_constraints_bigint(orc):
    rng = np.random.default_rng(5)
    tlog, elog, n_cols = 3, 4, 6
    cols = [rng.integers(0, P, 1 << elog, dtype=np.uint32) for _ in range(n_cols)]
    coeffs = [rq(rng) for _ in range(n_cols - 2)]
    den = [int(x) for x in rng.integers(1, P, 1 << (elog - tlog))]
    start = [rng.integers(0, P, 1 << elog, dtype=np.uint32) for _ in range(4)]
    out = orc.widefib_constraints(tlog, elog, cols, coeffs, den, start)
    for row in range(1 << elog):
        acc = (0, 0, 0, 0)
        for i in range(n_cols - 2):
            a, b, c = int(cols[i][row]), int(cols[i + 1][row]), int(cols[i + 2][row])
            acc = pf.qm_add(acc, pf.qm_scale(coeffs[i], (c - a * a - b * b) % P))
        acc = pf.qm_scale(acc, den[row >> tlog])
        assert [int(o[row]) for o in out] == [(int(s[row]) + v) % P for s, v in zip(start, acc)]


def test_coset_vanishing_vanishes_on_the_coset(orc):
    log = 4
    init = orc.subgroup_gen(log + 1)                     # Coset.odds(log), the trace coset of CanonicCoset(log)
    step = orc.subgroup_gen(log)
    for i in range(1 << log):
        assert orc.coset_vanishing(init, log, orc.index_to_point(init + i * step)) == 0
    hi, hl = orc.canonic_half_coset(log + 1)
    assert all(orc.coset_vanishing(init, log, orc.domain_point(hi, hl, i)) != 0 for i in range(1 << (log + 1)))


@pytest.mark.parametrize("log,n_cols", [(4, 3), (5, 7)])
def test_widefib_quotient_is_low_degree(orc, log, n_cols):
    rng = np.random.default_rng(log)
    elog = log + 1
    hi, hl = orc.canonic_half_coset(elog)
    tw, itw = orc.precompute_twiddles(hi, hl)
    root = (hi, hl)
    thi, _ = orc.canonic_half_coset(log)
    trace = fib_trace(log, n_cols)
    evals = [orc.evaluate(orc.interpolate(c, thi, root, itw, quirks=False), elog, hi, root, tw, quirks=False) for c in trace]
    coeffs = [rq(rng) for _ in range(n_cols - 2)]
    den = orc.widefib_denom_inv(log, elog)
    assert len(den) == 2
    zero = [np.zeros(1 << elog, dtype=np.uint32) for _ in range(4)]
    out = orc.widefib_constraints(log, elog, evals, coeffs, den, zero)
    for c in out:
        # degree(constraint) <= 2^log and the vanishing polynomial has degree 2^(log-1): the quotient lies in the span of the first
        # 2^log + 1 basis functions of the 2^(log+1) domain (index 2^log is the x^(2^(log-1)) one)
        co = orc.interpolate(c, hi, root, itw, quirks=False)
        assert co[:1 << log].any() and not co[(1 << log) + 1:].any()
    # break one trace cell: the quotient is no longer a polynomial of that degree
    trace[2][3] = (int(trace[2][3]) + 1) % P
    evals = [orc.evaluate(orc.interpolate(c, thi, root, itw, quirks=False), elog, hi, root, tw, quirks=False) for c in trace]
    out = orc.widefib_constraints(log, elog, evals, coeffs, den, zero)
    assert orc.interpolate(out[0], hi, root, itw, quirks=False)[(1 << log) + 1:].any()


def test_host_accumulators(orc):
    from tstwo_b200 import air
    rng = np.random.default_rng(9)
    alpha = rq(rng)
    pows = air.generate_secure_powers(alpha, 7)
    assert (np.array(pows, dtype=np.uint32) == orc.generate_secure_powers(alpha, 7)).all()
    assert air.generate_secure_powers(alpha, 0) == []
    x, y = rq(rng), rq(rng)
    assert air.qm31_mul(x, y) == list(pf.qm_mul(x, y))
    # PointEvaluationAccumulator: sum_i alpha^(N-1-i) u_i
    us = [rq(rng) for _ in range(5)]
    acc = air.PointEvaluationAccumulator.new(alpha)
    for u in us:
        acc.accumulate(u)
    exp = (0, 0, 0, 0)
    for i, u in enumerate(us):
        exp = pf.qm_add(exp, pf.qm_mul(pows[len(us) - 1 - i], u))
    assert acc.finalize() == list(exp)


def test_coset_vanishing_host_matches_oracle(orc):
    from tstwo_b200 import air
    from tstwo_b200.circle import CanonicCoset
    for log in (1, 3, 6):
        comp = air.WideFibonacciComponent(air.WideFibonacciEval(log, 4))
        assert comp.denominator_inverses() == orc.widefib_denom_inv(log, log + 1)
        dom = CanonicCoset.new(log + 1).circle_domain()
        hi, hl = orc.canonic_half_coset(log + 1)
        for i in range(1 << (log + 1)):
            assert air.coset_vanishing(CanonicCoset.new(log).coset, dom.at(i)) == orc.coset_vanishing(orc.subgroup_gen(log + 1), log, orc.domain_point(hi, hl, i))


def test_wide_fibonacci_eval_walk():
    from tstwo_b200 import air
    with pytest.raises(ValueError):
        air.WideFibonacciEval(-1, 5)
    with pytest.raises(ValueError):
        air.WideFibonacciEval(3, 1)

    class RowEval:   # one row of a valid trace: every constraint must be 0
        def __init__(self, row): self.row, self.k, self.cons = row, 0, []
        def nextTraceMask(self):
            self.k += 1
            return self.row[self.k - 1]
        def addConstraint(self, c): self.cons.append(c % P)

    tr = air.generate_wide_fibonacci_trace(3, 6)
    ev = air.WideFibonacciEval.create(3, 6)
    assert ev.maxConstraintLogDegreeBound() == 4 and ev.nConstraints() == 4
    for r in range(8):
        assert ev.evaluate(RowEval([int(tr[k][r]) for k in range(6)])).cons == [0, 0, 0, 0]
    assert (tr[0] == 1).all() and (tr[1] == np.arange(8)).all()
