"""Independent big-int model of the fields / circle group, used only to cross-check the C oracle."""
P = 2**31 - 1
GEN = (2, 1268011823)  # circle.ts:137


def inv(a):
    return pow(a, P - 2, P)


def cadd(p, q):
    return ((p[0] * q[0] - p[1] * q[1]) % P, (p[0] * q[1] + p[1] * q[0]) % P)


def cmul(p, k):
    r = (1, 0)
    while k:
        if k & 1:
            r = cadd(r, p)
        p = cadd(p, p)
        k >>= 1
    return r


def to_point(index):
    return cmul(GEN, index % (1 << 31))


def bitrev(i, n):
    r = 0
    for _ in range(n):
        r = (r << 1) | (i & 1)
        i >>= 1
    return r


def domain_index_at(half_init, half_log, i):
    half = 1 << half_log
    step = (1 << (31 - half_log)) if half_log else 0
    if i < half:
        return (half_init + step * i) % (1 << 31)
    return (-(half_init + step * (i - half))) % (1 << 31)


def circle_basis_eval(coeffs, x, y):
    """sum_j coeffs[j] * y^{j0} x^{j1} pi(x)^{j2} pi^2(x)^{j3} ...  (bit0<->y, bit1<->x, ...)."""
    n = len(coeffs).bit_length() - 1
    maps = [y]
    cur = x
    for _ in range(1, n):
        maps.append(cur)
        cur = (2 * cur * cur - 1) % P
    tot = 0
    for j, c in enumerate(coeffs):
        t = c
        for b in range(n):
            if (j >> b) & 1:
                t = t * maps[b] % P
        tot += t
    return tot % P


# QM31 as 4-tuples (a,b,c,d) = (a+bi) + (c+di)u, u^2 = 2+i
def cm_mul(a, b):
    return ((a[0] * b[0] - a[1] * b[1]) % P, (a[0] * b[1] + a[1] * b[0]) % P)


def qm_mul(a, b):
    a0, a1, b0, b1 = a[:2], a[2:], b[:2], b[2:]
    r = (2, 1)
    x = cm_mul(a0, b0); y = cm_mul(cm_mul(r, a1), b1)
    z = cm_mul(a0, b1); w = cm_mul(a1, b0)
    return ((x[0] + y[0]) % P, (x[1] + y[1]) % P, (z[0] + w[0]) % P, (z[1] + w[1]) % P)


def qm_add(a, b):
    return tuple((x + y) % P for x, y in zip(a, b))


def qm_sub(a, b):
    return tuple((x - y) % P for x, y in zip(a, b))


def qm_scale(a, k):
    return tuple(x * k % P for x in a)
