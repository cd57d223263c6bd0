"""Random-linear-combination accumulators and the wide-Fibonacci component over the device kernels.

Mirrors air/accumulator.ts (PointEvaluationAccumulator :33-76, DomainEvaluationAccumulator :91-262, ColumnAccumulator
:275-298), secure_poly.ts (SecureCirclePoly), constraints.ts:32-47 (cosetVanishing) and examples/fibonacci.ts
(WideFibonacciEval :36-92, generateTestTrace :245-258).  The loop over the evaluation domain that the TS code leaves as a
placeholder (constraint_framework/index.ts:131-152) is Stwo's evaluate_constraint_quotients_on_domain, done here by one
kernel (tstwo_widefib_constraints).  Field elements are python ints; QM31 = [a, b, c, d]."""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence, Tuple

import numpy as np

from . import backend as _b
from ._lib import check, dptr_array, u32_array
from .circle import CanonicCoset, Coset

P = (1 << 31) - 1


# ---- QM31 on the host (fields/qm31.ts:183-233), only what the accumulators need --------------------------------------------
def _cm_mul(a, b): return ((a[0] * b[0] - a[1] * b[1]) % P, (a[0] * b[1] + a[1] * b[0]) % P)


def qm31_add(x, y): return [(a + b) % P for a, b in zip(x, y)]


def qm31_mul(x, y):
    """(a + bu)(c + du) = (ac + R bd) + (ad + bc)u with R = 2 + i (qm31.ts:9, 223-233)."""
    a, b, c, d = (x[0], x[1]), (x[2], x[3]), (y[0], y[1]), (y[2], y[3])
    ac, bd, ad, bc = _cm_mul(a, c), _cm_mul(b, d), _cm_mul(a, d), _cm_mul(b, c)
    rbd = _cm_mul((2, 1), bd)
    return [(ac[0] + rbd[0]) % P, (ac[1] + rbd[1]) % P, (ad[0] + bc[0]) % P, (ad[1] + bc[1]) % P]


def qm31_sub(x, y): return [(a - b) % P for a, b in zip(x, y)]


def qm31_scale(x, k: int): return [a * k % P for a in x]


def qm31_inv(x):
    """qm31.ts:282-305: (a + bu)^-1 = (a - bu) / (a^2 - (2 + i) b^2), the denominator a CM31 (cm31.ts inverse)."""
    a, b = (x[0], x[1]), (x[2], x[3])
    b2 = _cm_mul(b, b)
    rb2 = _cm_mul((2, 1), b2)
    a2 = _cm_mul(a, a)
    den = ((a2[0] - rb2[0]) % P, (a2[1] - rb2[1]) % P)
    norm = (den[0] * den[0] + den[1] * den[1]) % P
    if norm == 0:
        raise ZeroDivisionError("0 has no inverse")
    ninv = pow(norm, P - 2, P)
    dinv = (den[0] * ninv % P, (-den[1]) * ninv % P)
    ra, rb = _cm_mul(a, dinv), _cm_mul(((-b[0]) % P, (-b[1]) % P), dinv)
    return [ra[0], ra[1], rb[0], rb[1]]


def from_partial_evals(evals: Sequence[Sequence[int]]) -> List[int]:
    """qm31.ts:411-428: e0 + e1 i + e2 u + e3 iu."""
    res = list(evals[0])
    for k, unit in ((1, [0, 1, 0, 0]), (2, [0, 0, 1, 0]), (3, [0, 0, 0, 1])):
        res = qm31_add(res, qm31_mul(list(evals[k]), unit))
    return res


def generate_secure_powers(felt: Sequence[int], n_powers: int) -> List[List[int]]:
    """backend/cpu/accumulation.ts:52-63 / DomainEvaluationAccumulator.generateSecurePowers (accumulator.ts:247-257)."""
    out, cur = [], [1, 0, 0, 0]
    for _ in range(n_powers):
        out.append(cur)
        cur = qm31_mul(cur, list(felt))
    return out


class PointEvaluationAccumulator:
    """accumulator.ts:33-76: f(P0) = sum_i alpha^(N-1-i) u_i(P0), evaluations fed in order (Horner)."""

    def __init__(self, random_coeff, accumulation):
        self.randomCoeff, self.accumulation = list(random_coeff), list(accumulation)

    @staticmethod
    def new(random_coeff) -> "PointEvaluationAccumulator":
        return PointEvaluationAccumulator(random_coeff, [0, 0, 0, 0])

    def accumulate(self, evaluation) -> None:
        self.accumulation = qm31_add(qm31_mul(self.accumulation, self.randomCoeff), list(evaluation))

    def finalize(self): return list(self.accumulation)
    getCurrentAccumulation = finalize


class SecureCirclePoly:
    """poly/circle/secure_poly.ts: 4 coordinate polynomials of a QM31-valued circle polynomial, on the device."""

    def __init__(self, polys: Sequence[_b.CudaCirclePoly]):
        if len(polys) != 4:
            raise ValueError("SecureCirclePoly needs 4 coordinate polynomials")
        self.polys = list(polys)

    def log_size(self) -> int: return self.polys[0].logSize()
    logSize = log_size

    def into_coordinate_polys(self): return list(self.polys)

    def eval_columns_at_point(self, point) -> List[List[int]]:
        return [_b.CudaCirclePoly.eval_at_point(p, point) for p in self.polys]

    def eval_at_point(self, point) -> List[int]:
        """from_partial_evals (qm31.ts:411-428): e0 + e1 i + e2 u + e3 iu."""
        return from_partial_evals(self.eval_columns_at_point(point))

    def evaluateWithTwiddles(self, domain, twiddles) -> "_b.SecureEvaluation":
        evs = _b.CudaBackend.evaluatePolynomialsOnDomain(self.polys, domain, twiddles)
        return _b.SecureEvaluation(domain, _b.SecureColumnByCoords([e.values for e in evs]))


class ColumnAccumulator:
    """accumulator.ts:275-298."""

    def __init__(self, random_coeff_powers, col: _b.SecureColumnByCoords):
        if col is None:
            raise ValueError("ColumnAccumulator: invalid inputs")
        self.randomCoeffPowers = [list(p) for p in random_coeff_powers]
        self.col = col

    def accumulate(self, index: int, evaluation) -> None:
        if index < 0 or index >= self.col.len():
            raise IndexError(f"ColumnAccumulator.accumulate: index {index} out of bounds (length: {self.col.len()})")
        self.col.set(index, qm31_add(self.col.at(index), list(evaluation)))

    def accumulate_column(self, other: _b.SecureColumnByCoords) -> None:
        """The whole column at once (what a device component does instead of one accumulate per row)."""
        _b.accumulate(self.col, other)


class DomainEvaluationAccumulator:
    """accumulator.ts:91-262.  sub-accumulations are device columns; finalize() runs on the device FFTs."""

    def __init__(self, random_coeff_powers, sub_accumulations):
        self.randomCoeffPowers = random_coeff_powers
        self.subAccumulations: List[Optional[_b.SecureColumnByCoords]] = sub_accumulations

    @staticmethod
    def new(random_coeff, max_log_size: int, total_columns: int, backend=None) -> "DomainEvaluationAccumulator":
        return DomainEvaluationAccumulator(generate_secure_powers(random_coeff, total_columns), [None] * (max_log_size + 1))

    generateSecurePowers = staticmethod(generate_secure_powers)

    def columns(self, n_cols_per_size: Sequence[Tuple[int, int]]) -> List[ColumnAccumulator]:
        sizes = [s for s, _ in n_cols_per_size]
        if len(set(sizes)) != len(sizes):
            raise ValueError("DomainEvaluationAccumulator.columns: duplicate log sizes not allowed")
        out = []
        for log_size, n_cols in n_cols_per_size:
            if log_size < 0 or log_size >= len(self.subAccumulations):
                raise ValueError(f"DomainEvaluationAccumulator.columns: invalid log_size {log_size}")
            if len(self.randomCoeffPowers) < n_cols:
                raise ValueError("DomainEvaluationAccumulator.columns: not enough random coefficients")
            cut = len(self.randomCoeffPowers) - n_cols           # splice(-nCols, nCols)
            coeffs, self.randomCoeffPowers = self.randomCoeffPowers[cut:], self.randomCoeffPowers[:cut]
            if self.subAccumulations[log_size] is None:
                self.subAccumulations[log_size] = _b.SecureColumnByCoords.zeros(1 << log_size)
            out.append(ColumnAccumulator(coeffs, self.subAccumulations[log_size]))
        return out

    def logSize(self) -> int: return len(self.subAccumulations) - 1
    log_size = logSize

    def finalize(self) -> SecureCirclePoly:
        """accumulator.ts:189-238: lift the running polynomial to each larger domain, add, interpolate."""
        if self.randomCoeffPowers:
            raise ValueError("DomainEvaluationAccumulator.finalize: not all random coefficients were used")
        cur: Optional[SecureCirclePoly] = None
        for log in range(1, self.logSize() + 1):
            values = self.subAccumulations[log]
            if values is None:
                continue
            domain = CanonicCoset.new(log).circle_domain()
            tw = _b.CudaBackend.precomputeTwiddles(domain.halfCoset)
            if cur is not None:
                _b.accumulate(values, cur.evaluateWithTwiddles(domain, tw).values)
            evs = [_b.CudaCircleEvaluation(domain, c) for c in values.columns]
            cur = SecureCirclePoly(_b.CudaBackend.interpolateColumns(evs, tw))
        if cur is None:
            n = 1 << self.logSize()
            return SecureCirclePoly([_b.CudaCirclePoly(_b.CudaColumn.zeros(n)) for _ in range(4)])
        return cur


# ---- constraints.ts:32-47 ------------------------------------------------------------------------------------------------
def coset_vanishing(coset: Coset, p: Tuple[int, int]) -> int:
    """Vanishing polynomial of `coset` at the M31 circle point p."""
    hx, hy = coset.step_size.half().to_point()
    ix, iy = coset.initial_index.to_point()
    # p - initial + half
    sx, sy = (p[0] * ix + p[1] * iy) % P, (p[1] * ix - p[0] * iy) % P
    x = (sx * hx - sy * hy) % P
    for _ in range(1, coset.log_size):
        x = (2 * x * x - 1) % P
    return x


def coset_vanishing_secure(coset: Coset, p) -> List[int]:
    """constraints.ts:32-47 at a CirclePoint<SecureField> p = (x[4], y[4])."""
    hx, hy = coset.step_size.half().to_point()
    ix, iy = coset.initial_index.to_point()
    px, py = list(p[0]), list(p[1])
    sx = qm31_add(qm31_scale(px, ix), qm31_scale(py, iy))            # p - initial
    sy = qm31_sub(qm31_scale(py, ix), qm31_scale(px, iy))
    x = qm31_sub(qm31_scale(sx, hx), qm31_scale(sy, hy))             # + half, x only
    for _ in range(1, coset.log_size):
        x = qm31_sub(qm31_scale(qm31_mul(x, x), 2), [1, 0, 0, 0])
    return x


def get_random_point(channel):
    """CirclePoint.get_random_point (circle.ts:126-133): t = draw_felt(); ((1 - t^2) / (1 + t^2), 2t / (1 + t^2))."""
    t = channel.draw_felt()
    t2 = qm31_mul(t, t)
    inv = qm31_inv(qm31_add(t2, [1, 0, 0, 0]))
    return (tuple(qm31_mul(qm31_sub([1, 0, 0, 0], t2), inv)), tuple(qm31_mul(qm31_scale(t, 2), inv)))


def bit_reverse_list(items):
    n = len(items)
    bits = n.bit_length() - 1
    return [items[int(format(i, f"0{bits}b")[::-1], 2) if bits else 0] for i in range(n)]


# ---- examples/fibonacci.ts -------------------------------------------------------------------------------------------------
def generate_wide_fibonacci_trace(log_n_rows: int, sequence_length: int) -> np.ndarray:
    """generateTestTrace (fibonacci.ts:245-258): row r starts (1, r), then c = a^2 + b^2; (n_cols, 2^log) in natural row order."""
    n = 1 << log_n_rows
    cols = np.zeros((sequence_length, n), dtype=np.uint64)
    cols[0], cols[1] = 1, np.arange(n, dtype=np.uint64)
    for k in range(2, sequence_length):
        cols[k] = (cols[k - 2] * cols[k - 2] % P + cols[k - 1] * cols[k - 1] % P) % P
    return cols.astype(np.uint32)


class WideFibonacciEval:
    """fibonacci.ts:36-135."""

    def __init__(self, log_n_rows: int, sequence_length: int = 100):
        if log_n_rows < 0:
            raise ValueError("WideFibonacciEval: logNRows must be non-negative")
        if sequence_length < 2:
            raise ValueError("WideFibonacciEval: sequenceLength must be at least 2")
        self.logNRows, self.sequenceLength = log_n_rows, sequence_length

    create = staticmethod(lambda log_n_rows, sequence_length=100: WideFibonacciEval(log_n_rows, sequence_length))

    def logSize(self) -> int: return self.logNRows
    def maxConstraintLogDegreeBound(self) -> int: return self.logNRows + 1
    def getSequenceLength(self) -> int: return self.sequenceLength
    def nConstraints(self) -> int: return self.sequenceLength - 2

    def evaluate(self, evaluator):
        """The symbolic walk of the TS code: nextTraceMask() per column, addConstraint(c - (a^2 + b^2))."""
        a, b = evaluator.nextTraceMask(), evaluator.nextTraceMask()
        for _ in range(2, self.sequenceLength):
            c = evaluator.nextTraceMask()
            evaluator.addConstraint(c - (a * a + b * b))
            a, b = b, c
        return evaluator


class WideFibonacciComponent:
    """FrameworkComponent<WideFibonacciEval> (constraint_framework/index.ts:38-170) for the prover side."""

    def __init__(self, evaluator: WideFibonacciEval, trace_tree: int = 1, col_offset: int = 0):
        self.evaluator, self.trace_tree, self.col_offset = evaluator, trace_tree, col_offset

    def nConstraints(self) -> int: return self.evaluator.nConstraints()
    def maxConstraintLogDegreeBound(self) -> int: return self.evaluator.maxConstraintLogDegreeBound()
    def getEvaluator(self): return self.evaluator

    def traceLogDegreeBounds(self):
        """TreeVec [preprocessed, main trace] of column log degree bounds (constraint_framework/index.ts:97-106)."""
        return [[], [self.evaluator.logSize()] * self.evaluator.sequenceLength]

    def maskPoints(self, point):
        """Every column is read at offset 0 only (constraint_framework/index.ts:108-114)."""
        return [[], [[point] for _ in range(self.evaluator.sequenceLength)]]

    def evaluateConstraintQuotientsAtPoint(self, point, mask, accumulator: PointEvaluationAccumulator) -> None:
        """constraint_framework/index.ts:120-129 with Stwo's PointEvaluator: accumulate(denom_inverse * constraint) per constraint,
        denom = coset_vanishing(CanonicCoset(log).coset, point).  mask[tree][column] = sampled values of that column."""
        ev = self.evaluator
        cols = mask[self.trace_tree][self.col_offset:self.col_offset + ev.sequenceLength]
        denom_inv = qm31_inv(coset_vanishing_secure(CanonicCoset.new(ev.logSize()).coset, point))
        a, b = list(cols[0][0]), list(cols[1][0])
        for k in range(2, ev.sequenceLength):
            c = list(cols[k][0])
            constraint = qm31_sub(c, qm31_add(qm31_mul(a, a), qm31_mul(b, b)))
            accumulator.accumulate(qm31_mul(denom_inv, constraint))
            a, b = b, c

    def denominator_inverses(self) -> List[int]:
        """1 / coset_vanishing(trace coset, eval_domain.at(i)), i < 2^log_expand, bit-reversed (Stwo component.rs)."""
        ev = self.evaluator
        eval_domain = CanonicCoset.new(ev.maxConstraintLogDegreeBound()).circle_domain()
        trace_coset = CanonicCoset.new(ev.logSize()).coset
        n = 1 << (ev.maxConstraintLogDegreeBound() - ev.logSize())
        return bit_reverse_list([pow(coset_vanishing(trace_coset, eval_domain.at(i)), P - 2, P) for i in range(n)])

    def evaluateConstraintQuotientsOnDomain(self, trace, accumulator: DomainEvaluationAccumulator, twiddles=None) -> None:
        """trace: the interpolated trace columns, or a Trace (polys / evals per tree, prover.py) of which this component reads
        its span of the main-trace tree.  Committed evaluations on CanonicCoset(log + 1) are used as they are (blow-up 2);
        otherwise the polynomials are evaluated there.  One kernel then adds denom_inv * sum_i alpha^(N-1-i) constraint_i into
        the accumulator's column of that size."""
        ev = self.evaluator
        eval_log = ev.maxConstraintLogDegreeBound()
        span = slice(self.col_offset, self.col_offset + ev.sequenceLength)
        if hasattr(trace, "polys"):
            evals = trace.evals[self.trace_tree][span]
            if len(evals) == ev.sequenceLength and all(e.values.len() == 1 << eval_log for e in evals):
                return self.evaluate_on_evals(evals, accumulator)
            trace_polys = trace.polys[self.trace_tree][span]
        else:
            trace_polys = list(trace)
        if len(trace_polys) != ev.sequenceLength:
            raise ValueError("trace width does not match the sequence length")
        domain = CanonicCoset.new(eval_log).circle_domain()
        tw = twiddles or _b.CudaBackend.precomputeTwiddles(domain.halfCoset)
        evals = _b.CudaBackend.evaluatePolynomialsOnDomain(list(trace_polys), domain, tw)
        self.evaluate_on_evals(evals, accumulator)

    def evaluate_on_evals(self, evals: Sequence[_b.CudaCircleEvaluation], accumulator: DomainEvaluationAccumulator) -> None:
        ev = self.evaluator
        eval_log = ev.maxConstraintLogDegreeBound()
        if any(e.values.len() != 1 << eval_log for e in evals):
            raise ValueError("evaluations must be on CanonicCoset(log_size + 1)")
        [accum] = accumulator.columns([(eval_log, self.nConstraints())])
        coeffs = list(reversed(accum.randomCoeffPowers))          # constraint i gets alpha^(N-1-i)
        widefib_constraints(ev.logSize(), eval_log, [e.values for e in evals], coeffs, self.denominator_inverses(), accum.col)


def widefib_constraints(trace_log: int, eval_log: int, cols: Sequence[_b.CudaColumn], coeffs, denom_inv, out: _b.SecureColumnByCoords) -> None:
    flat = [int(x) for q in coeffs for x in q]
    check(_b._L().tstwo_widefib_constraints(trace_log, eval_log, dptr_array([c.ptr for c in cols]), len(cols),
                                            u32_array(flat if flat else [0]), u32_array([int(d) for d in denom_inv]), out.ptrs()))
