"""FRI folds / decompose / accumulation / quotients of the oracle vs the reference's property tests
(packages/core/test/fri.test.ts:206-267, test/backend/cpu/fri.test.ts:44-183, accumulation.test.ts)
and an independent big-int model."""
import numpy as np

import pyfield as F


def line_poly_eval(coeffs, x):
    # poly/line.ts LinePoly.evalAtPoint: fold(coeffs, [x, pi(x), ...]) (poly/utils.ts:36-59)
    n = len(coeffs).bit_length() - 1
    d = []
    cur = x
    for _ in range(n):
        d.append(cur); cur = (2 * cur * cur - 1) % F.P
    def fold(vals, fac):
        if len(vals) == 1:
            return vals[0]
        hlf = len(vals) // 2
        return (fold(vals[:hlf], fac[1:]) + fold(vals[hlf:], fac[1:]) * fac[0]) % F.P
    return fold(coeffs, d)


def test_fold_line_works(orc):
    # fri.test.ts:206-254
    even, odd = [1, 2, 1, 3], [3, 5, 4, 1]
    coeffs = even + odd
    alpha = 19283
    log = 3
    init = orc.half_odds_index(log)
    step = 1 << (31 - log)
    xs = [F.to_point(init + k * step)[0] for k in range(8)]
    vals = [line_poly_eval(coeffs, x) for x in xs]
    br = [vals[F.bitrev(i, log)] for i in range(8)]
    src = [np.array(br, dtype=np.uint32)] + [np.zeros(8, dtype=np.uint32)] * 3
    out = orc.fold_line(src, init, [alpha, 0, 0, 0])
    got = [int(out[0][F.bitrev(i, 2)]) for i in range(4)]
    dinit, dstep = (2 * init) % 2**31, 1 << (31 - 2)
    for i in range(4):
        x = F.to_point(dinit + i * dstep)[0]
        exp = 2 * (line_poly_eval(even, x) + alpha * line_poly_eval(odd, x)) % F.P
        assert got[i] == exp
    assert all((o == 0).all() for o in out[1:])


def test_fold_line_bigint(orc):
    rng = np.random.default_rng(3)
    log = 6
    init = orc.half_odds_index(log)
    src = [rng.integers(0, F.P, 1 << log, dtype=np.uint32) for _ in range(4)]
    alpha = tuple(int(v) for v in rng.integers(0, F.P, 4))
    out = orc.fold_line(src, init, alpha)
    step = 1 << (31 - log)
    for i in range(1 << (log - 1)):
        f0 = tuple(int(c[2 * i]) for c in src); f1 = tuple(int(c[2 * i + 1]) for c in src)
        x = F.to_point(init + F.bitrev(2 * i, log) * step)[0]
        a = F.qm_add(f0, f1); b = F.qm_scale(F.qm_sub(f0, f1), F.inv(x))
        exp = F.qm_add(a, F.qm_mul(alpha, b))
        assert tuple(int(c[i]) for c in out) == exp


def test_fold_circle_into_line_bigint(orc):
    rng = np.random.default_rng(4)
    log = 6
    hinit, hlog = orc.canonic_half_coset(log)
    src = [rng.integers(0, F.P, 1 << log, dtype=np.uint32) for _ in range(4)]
    dst = [rng.integers(0, F.P, 1 << (log - 1), dtype=np.uint32) for _ in range(4)]
    alpha = tuple(int(v) for v in rng.integers(0, F.P, 4))
    out = orc.fold_circle_into_line(dst, src, hinit, alpha)
    a2 = F.qm_mul(alpha, alpha)
    for i in range(1 << (log - 1)):
        f0 = tuple(int(c[2 * i]) for c in src); f1 = tuple(int(c[2 * i + 1]) for c in src)
        y = F.to_point(F.domain_index_at(hinit, hlog, F.bitrev(2 * i, log)))[1]
        a = F.qm_add(f0, f1); b = F.qm_scale(F.qm_sub(f0, f1), F.inv(y))
        fp = F.qm_add(F.qm_mul(alpha, b), a)
        exp = F.qm_add(F.qm_mul(tuple(int(c[i]) for c in dst), a2), fp)
        assert tuple(int(c[i]) for c in out) == exp


def test_fold_circle_degree_drops(orc):
    # fri.test.ts:256-267: a degree < 2^4 circle poly on a blowup-2 domain folds into a line eval of degree < 2^3
    log_deg, blow = 4, 1
    log = log_deg + blow
    hinit, hlog = orc.canonic_half_coset(log)
    tw, itw = orc.precompute_twiddles(hinit, hlog)
    coeffs = np.arange(1, 17, dtype=np.uint32)
    ev = orc.evaluate(coeffs, log, hinit, (hinit, hlog), tw)
    src = [ev] + [np.zeros_like(ev)] * 3
    dst = [np.zeros(1 << (log - 1), dtype=np.uint32)] * 4
    out = orc.fold_circle_into_line(dst, src, hinit, [1, 0, 0, 0])
    # interpolate the line evaluation (poly/line.ts:312-390) with big ints: values bit-reversed -> natural
    n = 1 << (log - 1)
    v = [int(out[0][F.bitrev(i, log - 1)]) for i in range(n)]
    ci, cl = hinit, hlog
    size = n
    while size > 1:
        step = 1 << (31 - cl)
        for s in range(0, n, size):
            for i in range(size // 2):
                x = F.inv(F.to_point(ci + i * step)[0])
                a, b = v[s + i], v[s + i + size // 2]
                v[s + i], v[s + i + size // 2] = (a + b) % F.P, (a - b) * x % F.P
        ci, cl, size = (2 * ci) % 2**31, cl - 1, size // 2
    ordered = [v[F.bitrev(i, log - 1)] for i in range(n)]   # LinePoly.intoOrderedCoefficients
    assert any(ordered[: 1 << (log_deg - 1)]) and not any(ordered[1 << (log_deg - 1):])


def test_decompose(orc):
    # backend/cpu/fri.test.ts:44-124
    rng = np.random.default_rng(5)
    for log in [0, 1, 5, 8]:
        src = [rng.integers(0, F.P, 1 << log, dtype=np.uint32) for _ in range(4)]
        g, lam = orc.fri_decompose(src)
        n = 1 << log
        for k in range(4):
            s = src[k].astype(object)
            if n == 1:
                exp = (-int(s[0])) % F.P
            else:
                exp = (int(s[: n // 2].sum()) - int(s[n // 2:].sum())) * F.inv(n) % F.P
            assert lam[k] == exp
            sign = np.array([-1] * max(n // 2, 1) + [1] * (n // 2), dtype=object)[:n]
            assert ((s + sign * exp) % F.P == g[k].astype(object)).all()


def test_accumulate_and_powers(orc):
    rng = np.random.default_rng(6)
    a = [rng.integers(0, F.P, 100, dtype=np.uint32) for _ in range(4)]
    b = [rng.integers(0, F.P, 100, dtype=np.uint32) for _ in range(4)]
    out = orc.accumulate(a, b)
    for k in range(4):
        assert ((a[k].astype(np.uint64) + b[k]) % F.P == out[k]).all()
    f = (3, 5, 7, 11)
    pw = orc.generate_secure_powers(f, 5)
    acc = (1, 0, 0, 0)
    for i in range(5):
        assert tuple(int(v) for v in pw[i]) == acc
        acc = F.qm_mul(acc, f)


def test_accumulate_quotients_bigint(orc):
    # follows backend/cpu/quotients.ts exactly (incl. the M31-part denominators)
    rng = np.random.default_rng(8)
    log = 5
    hinit, hlog = orc.canonic_half_coset(log)
    cols = [rng.integers(0, F.P, 1 << log, dtype=np.uint32) for _ in range(3)]
    rc = tuple(int(v) for v in rng.integers(1, F.P, 4))
    def rq(): return tuple(int(v) for v in rng.integers(1, F.P, 4))
    batches = [((rq(), rq()), [(0, rq()), (2, rq())]), ((rq(), rq()), [(1, rq())])]
    out = orc.accumulate_quotients(log, hinit, cols, rc, batches)
    conj = lambda q: (q[0], (-q[1]) % F.P, q[2], (-q[3]) % F.P)
    for row in [0, 1, 7, 31]:
        dx, dy = F.to_point(F.domain_index_at(hinit, hlog, F.bitrev(row, log)))
        acc = (0, 0, 0, 0)
        for (px, py), entries in batches:
            alpha = (1, 0, 0, 0); num = (0, 0, 0, 0)
            for (ci, v) in entries:
                alpha = F.qm_mul(alpha, rc)
                a = F.qm_sub(conj(v), v); c = F.qm_sub(conj(py), py)
                b = F.qm_sub(F.qm_mul(v, c), F.qm_mul(a, py))
                a, b, c = F.qm_mul(alpha, a), F.qm_mul(alpha, b), F.qm_mul(alpha, c)
                val = F.qm_scale(c, int(cols[ci][row]))
                lin = F.qm_add(F.qm_scale(a, dy), b)
                num = F.qm_add(num, F.qm_sub(val, lin))
            den = ((px[0] - dx) * py[1] - (py[0] - dy) * px[1]) % F.P
            acc = F.qm_add(F.qm_mul(acc, alpha), F.qm_scale(num, F.inv(den)))
        assert tuple(int(o[row]) for o in out) == acc
