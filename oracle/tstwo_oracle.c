/*
 * tstwo_oracle.c -- CPU restatement of TStwo's CpuBackend for the LDE + Blake2s-commit hot path.
 *
 * THIS IS TEST INFRASTRUCTURE, NOT THE PRODUCT.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load this library.  The product path
 * (tstwo_b200/) never links, imports or calls anything here.
 *
 * The reference (teddyjfpender/tstwo, TypeScript) cannot be executed in the build container
 * (no bun/node), so this file restates its algorithms in plain C, function by function, each
 * citing the reference file:line (paths relative to /root/reference/packages/core/src).
 *
 * Pinning status:
 *   - field arithmetic (M31/CM31/QM31): pinned by the reference's Rust-generated golden vectors
 *     test-vectors/{m31,cm31,qm31}-test-vectors.json (tests/golden/field_vectors.json);
 *   - BLAKE2s: pinned by RFC 7693 / the literals in test/vcs/blake2_hash.test.ts:7-8,127-128 and
 *     by Python hashlib.blake2s (an independent implementation);
 *   - twiddles / interpolate / evaluate / FRI folds / quotients: "parity unpinned" by literals --
 *     the reference holds no golden outputs for them.  They are pinned only by the
 *     restated algorithm plus the reference's own property tests re-run on this code
 *     (evaluate == eval_at_point on domain points, interpolate o evaluate = id, fold identities).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define P 2147483647u
typedef uint32_t m31;
typedef struct { m31 re, im; } cm31;
typedef struct { cm31 c0, c1; } qm31;          /* (a + bi) + (c + di)u, u^2 = 2 + i */
typedef struct { m31 x, y; } cpoint;
typedef struct { qm31 x, y; } qpoint;

#define EXPORT __attribute__((visibility("default")))

/* ------------------------------------------------------------------ M31  (fields/m31.ts) */
/* fields/m31.ts:60-75 `_reduceBig`: ((((v >> 31) + v + 1) >> 31) + v) & P */
static inline m31 m31_reduce(uint64_t v) { return (m31)(((((v >> 31) + v + 1) >> 31) + v) & P); }
/* fields/m31.ts:39-44 partialReduce */
static inline m31 m31_partial(uint32_t v) { return v >= P ? v - P : v; }
static inline m31 m31_add(m31 a, m31 b) { return m31_partial(a + b); }            /* m31.ts:126-128 */
static inline m31 m31_sub(m31 a, m31 b) { return m31_partial(a + P - b); }        /* m31.ts:134-136 */
static inline m31 m31_neg(m31 a) { return a == 0 ? 0 : P - a; }                   /* m31.ts:104-109 */
static inline m31 m31_mul(m31 a, m31 b) { return m31_reduce((uint64_t)a * b); }   /* m31.ts:147-150 */
static inline m31 m31_sqn(m31 v, int n) { while (n-- > 0) v = m31_mul(v, v); return v; }
/* fields/m31.ts:305-326 pow2147483645 (37-multiplication addition chain) */
static m31 m31_inv(m31 v)
{
    m31 t0 = m31_mul(m31_sqn(v, 2), v);
    m31 t1 = m31_mul(m31_sqn(t0, 1), t0);
    m31 t2 = m31_mul(m31_sqn(t1, 3), t0);
    m31 t3 = m31_mul(m31_sqn(t2, 1), t0);
    m31 t4 = m31_mul(m31_sqn(t3, 8), t3);
    m31 t5 = m31_mul(m31_sqn(t4, 8), t3);
    return m31_mul(m31_sqn(t5, 7), t2);
}

/* ------------------------------------------------------------------ CM31 (fields/cm31.ts) */
static inline cm31 cm(m31 a, m31 b) { cm31 r = { a, b }; return r; }
static inline cm31 cm_add(cm31 a, cm31 b) { return cm(m31_add(a.re, b.re), m31_add(a.im, b.im)); }
static inline cm31 cm_sub(cm31 a, cm31 b) { return cm(m31_sub(a.re, b.re), m31_sub(a.im, b.im)); }
static inline cm31 cm_neg(cm31 a) { return cm(m31_neg(a.re), m31_neg(a.im)); }
/* cm31.ts `mul`: (ac - bd) + (ad + bc) i */
static inline cm31 cm_mul(cm31 a, cm31 b)
{
    return cm(m31_sub(m31_mul(a.re, b.re), m31_mul(a.im, b.im)),
              m31_add(m31_mul(a.re, b.im), m31_mul(a.im, b.re)));
}
static inline cm31 cm_mul_m31(cm31 a, m31 b) { return cm(m31_mul(a.re, b), m31_mul(a.im, b)); }
static inline cm31 cm_sqr(cm31 a) { return cm_mul(a, a); }
/* cm31.ts `inverse`: conj / (re^2 + im^2) */
static cm31 cm_inv(cm31 a)
{
    m31 norm = m31_add(m31_mul(a.re, a.re), m31_mul(a.im, a.im));
    return cm_mul_m31(cm(a.re, m31_neg(a.im)), m31_inv(norm));
}

/* ------------------------------------------------------------------ QM31 (fields/qm31.ts) */
static inline qm31 qm(m31 a, m31 b, m31 c, m31 d) { qm31 r = { { a, b }, { c, d } }; return r; }
static inline qm31 qm_from_m31(m31 a) { return qm(a, 0, 0, 0); }
static inline qm31 qm_add(qm31 a, qm31 b) { qm31 r = { cm_add(a.c0, b.c0), cm_add(a.c1, b.c1) }; return r; }
static inline qm31 qm_sub(qm31 a, qm31 b) { qm31 r = { cm_sub(a.c0, b.c0), cm_sub(a.c1, b.c1) }; return r; }
static inline qm31 qm_neg(qm31 a) { qm31 r = { cm_neg(a.c0), cm_neg(a.c1) }; return r; }
/* qm31.ts:183-193 `mul`: (a + bu)(c + du) = (ac + R bd) + (ad + bc)u, R = 2 + i (qm31.ts:9) */
static inline qm31 qm_mul(qm31 a, qm31 b)
{
    const cm31 R = { 2, 1 };
    qm31 r;
    r.c0 = cm_add(cm_mul(a.c0, b.c0), cm_mul(cm_mul(R, a.c1), b.c1));
    r.c1 = cm_add(cm_mul(a.c0, b.c1), cm_mul(a.c1, b.c0));
    return r;
}
static inline qm31 qm_mul_m31(qm31 a, m31 b) { qm31 r = { cm_mul_m31(a.c0, b), cm_mul_m31(a.c1, b) }; return r; }
static inline qm31 qm_mul_cm31(qm31 a, cm31 b) { qm31 r = { cm_mul(a.c0, b), cm_mul(a.c1, b) }; return r; }
/* qm31.ts `inverse`: b2 = c1^2; ib2 = i*b2; denom = c0^2 - (2 b2 + ib2); (c0 - c1 u)/denom */
static qm31 qm_inv(qm31 a)
{
    cm31 b2 = cm_sqr(a.c1);
    cm31 ib2 = cm(m31_neg(b2.im), b2.re);
    cm31 denom = cm_sub(cm_sqr(a.c0), cm_add(cm_add(b2, b2), ib2));
    cm31 di = cm_inv(denom);
    qm31 r = { cm_mul(a.c0, di), cm_mul(cm_neg(a.c1), di) };
    return r;
}
static inline qm31 qm_conj(qm31 a) { return qm(a.c0.re, m31_neg(a.c0.im), a.c1.re, m31_neg(a.c1.im)); }
static inline int qm_eq(qm31 a, qm31 b)
{
    return a.c0.re == b.c0.re && a.c0.im == b.c0.im && a.c1.re == b.c1.re && a.c1.im == b.c1.im;
}
static inline qm31 qm_ld(const uint32_t *v) { return qm(v[0], v[1], v[2], v[3]); }
static inline void qm_st(uint32_t *v, qm31 a) { v[0] = a.c0.re; v[1] = a.c0.im; v[2] = a.c1.re; v[3] = a.c1.im; }

/* Exported scalar entry points so the golden field vectors can be replayed (tests/test_oracle_fields.py). */
EXPORT uint32_t orc_m31_add(uint32_t a, uint32_t b) { return m31_add(a, b); }
EXPORT uint32_t orc_m31_sub(uint32_t a, uint32_t b) { return m31_sub(a, b); }
EXPORT uint32_t orc_m31_mul(uint32_t a, uint32_t b) { return m31_mul(a, b); }
EXPORT uint32_t orc_m31_neg(uint32_t a) { return m31_neg(a); }
EXPORT uint32_t orc_m31_inv(uint32_t a) { return m31_inv(a); }
EXPORT uint32_t orc_m31_reduce(uint64_t a) { return m31_reduce(a); }
EXPORT uint32_t orc_m31_partial_reduce(uint32_t a) { return m31_partial(a); }
EXPORT void orc_cm31_op(int op, const uint32_t *a, const uint32_t *b, uint32_t *out)
{
    cm31 x = cm(a[0], a[1]), y = b ? cm(b[0], b[1]) : cm(0, 0), r;
    switch (op) {
    case 0: r = cm_add(x, y); break;
    case 1: r = cm_sub(x, y); break;
    case 2: r = cm_mul(x, y); break;
    case 3: r = cm_neg(x); break;
    case 4: r = cm_inv(x); break;
    default: r = cm(0, 0);
    }
    out[0] = r.re; out[1] = r.im;
}
EXPORT void orc_qm31_op(int op, const uint32_t *a, const uint32_t *b, uint32_t *out)
{
    qm31 x = qm_ld(a), y = b ? qm_ld(b) : qm(0, 0, 0, 0), r;
    switch (op) {
    case 0: r = qm_add(x, y); break;
    case 1: r = qm_sub(x, y); break;
    case 2: r = qm_mul(x, y); break;
    case 3: r = qm_neg(x); break;
    case 4: r = qm_inv(x); break;
    case 5: r = qm_mul_cm31(x, cm(b[0], b[1])); break;
    case 6: r = qm_mul_m31(x, b[0]); break;
    default: r = qm(0, 0, 0, 0);
    }
    qm_st(out, r);
}

/* ------------------------------------------------------------------ circle group (circle.ts) */
static const cpoint M31_CIRCLE_GEN = { 2u, 1268011823u };       /* circle.ts:137 */
#define LOG_ORDER 31
#define IDX_MASK 0x7fffffffu

/* circle.ts:102-106 `add` */
static inline cpoint cp_add(cpoint a, cpoint b)
{
    cpoint r;
    r.x = m31_sub(m31_mul(a.x, b.x), m31_mul(a.y, b.y));
    r.y = m31_add(m31_mul(a.x, b.y), m31_mul(a.y, b.x));
    return r;
}
/* circle.ts:56-68 `mul` (double and add), circle.ts:172-174 CirclePointIndex.to_point */
static cpoint idx_to_point(uint32_t index)
{
    cpoint res = { 1, 0 }, cur = M31_CIRCLE_GEN;
    uint32_t s = index & IDX_MASK;
    while (s) {
        if (s & 1) res = cp_add(res, cur);
        cur = cp_add(cur, cur);
        s >>= 1;
    }
    return res;
}
/* circle.ts:167-170 subgroup_gen; for log_size 0 the index 2^31 == 0 (mod 2^31). */
static inline uint32_t subgroup_gen(int log_size) { return (log_size == 0) ? 0u : (1u << (LOG_ORDER - log_size)); }

typedef struct { uint32_t initial_index; int log_size; uint32_t step_size; cpoint initial, step; } coset_t;
/* circle.ts:199-211 Coset constructor */
static coset_t coset_new(uint32_t initial_index, int log_size)
{
    coset_t c;
    c.initial_index = initial_index & IDX_MASK;
    c.log_size = log_size;
    c.step_size = subgroup_gen(log_size);
    c.initial = idx_to_point(c.initial_index);
    c.step = idx_to_point(c.step_size);
    return c;
}
/* circle.ts:250-253 double: initial_index * 2, log_size - 1 */
static coset_t coset_double(const coset_t *c) { return coset_new((c->initial_index * 2u) & IDX_MASK, c->log_size - 1); }
/* circle.ts:274-276 index_at */
static inline uint32_t coset_index_at(uint32_t initial_index, int log_size, uint32_t i)
{
    return (initial_index + subgroup_gen(log_size) * i) & IDX_MASK;
}
/* circle.ts:264-266 is_doubling_of */
static int coset_is_doubling_of(uint32_t a_init, int a_log, uint32_t b_init, int b_log)
{
    if (a_log > b_log) return 0;
    uint32_t init = b_init & IDX_MASK;
    for (int i = 0; i < b_log - a_log; i++) init = (init * 2u) & IDX_MASK;
    return init == (a_init & IDX_MASK);
}
/* poly/circle/domain.ts:76-88 CircleDomain.indexAt (the CORRECT enumeration) */
static uint32_t domain_index_at(uint32_t half_init, int half_log, uint32_t i)
{
    uint32_t half = 1u << half_log;
    if (i < half) return coset_index_at(half_init, half_log, i);
    return (0u - coset_index_at(half_init, half_log, i - half)) & IDX_MASK;
}
/* poly/circle/domain.ts:35-38 CircleDomain.iter: half coset, then `halfCoset.conjugate()` which in TS
 * (circle.ts:288-290) negates only the initial index and keeps +step. */
static uint32_t domain_iter_index(uint32_t half_init, int half_log, uint32_t i)
{
    uint32_t half = 1u << half_log;
    if (i < half) return coset_index_at(half_init, half_log, i);
    return coset_index_at((0u - half_init) & IDX_MASK, half_log, i - half);
}
EXPORT void orc_index_to_point(uint32_t index, uint32_t *xy) { cpoint p = idx_to_point(index); xy[0] = p.x; xy[1] = p.y; }
EXPORT uint32_t orc_domain_index_at(uint32_t half_init, int half_log, uint32_t i) { return domain_index_at(half_init, half_log, i); }
EXPORT uint32_t orc_domain_iter_index(uint32_t half_init, int half_log, uint32_t i) { return domain_iter_index(half_init, half_log, i); }

/* utils.ts:15-22 bitReverseIndex */
static inline uint32_t bit_reverse_index(uint32_t idx, int log_size)
{
    uint32_t rev = 0;
    for (int i = 0; i < log_size; i++) { rev = (rev << 1) | (idx & 1); idx >>= 1; }
    return rev;
}
/* backend/cpu/index.ts:62-79 bitReverse (in place) */
EXPORT int orc_bit_reverse(uint32_t *a, uint64_t n)
{
    if (n == 0 || (n & (n - 1))) return -1;
    int log_n = 0;
    while ((1ull << log_n) < n) log_n++;
    for (uint64_t i = 0; i < n; i++) {
        uint64_t j = bit_reverse_index((uint32_t)i, log_n);
        if (j > i) { uint32_t t = a[i]; a[i] = a[j]; a[j] = t; }
    }
    return 0;
}

/* ------------------------------------------------------------------ twiddles
 * backend/cpu/circle.ts:210-221 slowPrecomputeTwiddles + :223-239 _precomputeTwiddles.
 * tw, itw: 2^log_size words each. */
EXPORT int orc_precompute_twiddles(uint32_t initial_index, int log_size, uint32_t *tw, uint32_t *itw)
{
    coset_t c = coset_new(initial_index, log_size);
    uint64_t off = 0;
    for (int i = 0; i < log_size; i++) {
        uint64_t half = (1ull << c.log_size) / 2;
        cpoint cur = c.initial;                      /* CosetIterator: cur, cur+step, ... (circle.ts:293-309) */
        for (uint64_t k = 0; k < half; k++) { tw[off + k] = cur.x; cur = cp_add(cur, c.step); }
        if (half > 1) orc_bit_reverse(tw + off, half);
        off += half;
        c = coset_double(&c);
    }
    tw[off] = 1;                                     /* pad (circle.ts:220) */
    uint64_t n = 1ull << log_size;
    if (itw) for (uint64_t i = 0; i < n; i++) itw[i] = m31_inv(tw[i]);   /* elementwise == batch inverse */
    return 0;
}

/* ------------------------------------------------------------------ circle FFT
 * backend/cpu/circle.ts:243-257 fftLayerLoop, fft.ts:12-30 butterflies */
static inline void butterfly(m31 *v0, m31 *v1, m31 t)
{
    m31 tmp = m31_mul(*v1, t), a = *v0;
    *v1 = m31_sub(a, tmp);
    *v0 = m31_add(a, tmp);
}
static inline void ibutterfly(m31 *v0, m31 *v1, m31 it)
{
    m31 a = *v0, b = *v1;
    *v0 = m31_add(a, b);
    *v1 = m31_mul(m31_sub(a, b), it);
}
static void fft_layer(m31 *v, uint64_t n, int i, const m31 *tw, int circle_layer, int inverse)
{
    /* for every block h of size 2^(i+1): twiddle t_h; pairs (idx0, idx0 + 2^i) */
    uint64_t nblocks = n >> (i + 1);
    for (uint64_t h = 0; h < nblocks; h++) {
        m31 t;
        if (circle_layer) {
            /* backend/cpu/circle.ts:270-278 circleTwiddlesFromLineTwiddles: [y, -y, -x, x] per (x, y) pair */
            m31 x = tw[(h >> 2) * 2], y = tw[(h >> 2) * 2 + 1];
            switch (h & 3) { case 0: t = y; break; case 1: t = m31_neg(y); break; case 2: t = m31_neg(x); break; default: t = x; }
        } else t = tw[h];
        m31 *b = v + (h << (i + 1));
        uint64_t half = 1ull << i;
        if (inverse) for (uint64_t l = 0; l < half; l++) ibutterfly(b + l, b + l + half, t);
        else for (uint64_t l = 0; l < half; l++) butterfly(b + l, b + l + half, t);
    }
}
/* poly/utils.ts:78-99 domainLineTwiddlesFromTree: layer k (k = 0 .. log_half-1) is the slice
 * [len_total - 2*len, len_total - len) of the root buffer with len = 2^(log_half-1-k). */
static inline const m31 *line_twiddles(const m31 *buf, uint64_t buf_len, int log_half, int k)
{
    uint64_t len = 1ull << (log_half - 1 - k);
    return buf + (buf_len - 2 * len);
}

/* backend/cpu/circle.ts:84-134 CpuCirclePoly.evaluate (with extend :71-82).
 * coeffs: 2^poly_log words; out: 2^domain_log words. tw = root-coset twiddle buffer (2^tw_log words). */
EXPORT int orc_evaluate(const uint32_t *coeffs, int poly_log, uint32_t *out, int domain_log,
                        uint32_t half_init, uint32_t tw_root_init, int tw_log, const uint32_t *tw, int quirks)
{
    int half_log = domain_log - 1;
    if (domain_log < 1 || poly_log > domain_log) return -2;
    if (!coset_is_doubling_of(half_init, half_log, tw_root_init, tw_log)) return -3;   /* "twiddle tree mismatch" */
    uint64_t n = 1ull << domain_log, np = 1ull << poly_log;
    memcpy(out, coeffs, np * 4);
    memset(out + np, 0, (n - np) * 4);
    cpoint init = idx_to_point(half_init & IDX_MASK);
    if (domain_log == 1) { butterfly(&out[0], &out[1], init.y); return 0; }
    if (domain_log == 2) {
        butterfly(&out[0], &out[2], init.x); butterfly(&out[1], &out[3], init.x);
        butterfly(&out[0], &out[1], init.y); butterfly(&out[2], &out[3], m31_neg(init.y));
        return 0;
    }
    uint64_t buf_len = 1ull << tw_log;
    for (int k = half_log - 1; k >= 0; k--) fft_layer(out, n, k + 1, line_twiddles(tw, buf_len, half_log, k), 0, 0);
    fft_layer(out, n, 0, line_twiddles(tw, buf_len, half_log, 0), 1, 0);
    if (quirks && domain_log == 3) { m31 t = out[5]; out[5] = out[7]; out[7] = t; }   /* circle.ts:127-131 (F9) */
    return 0;
}
/* backend/cpu/circle.ts:136-207 CpuCirclePoly.interpolate; values in place (2^domain_log words). */
EXPORT int orc_interpolate(uint32_t *v, int domain_log, uint32_t half_init, uint32_t tw_root_init, int tw_log,
                           const uint32_t *itw, int quirks)
{
    int half_log = domain_log - 1;
    if (domain_log < 1) return -2;
    if (!coset_is_doubling_of(half_init, half_log, tw_root_init, tw_log)) return -3;
    uint64_t n = 1ull << domain_log;
    if (quirks && domain_log == 3) { m31 t = v[5]; v[5] = v[7]; v[7] = t; }           /* circle.ts:147-151 */
    cpoint init = idx_to_point(half_init & IDX_MASK);
    if (domain_log == 1) {
        m31 y = init.y, nn = 2, yn_inv = m31_inv(m31_mul(y, nn));
        m31 y_inv = m31_mul(yn_inv, nn), n_inv = m31_mul(yn_inv, y);
        ibutterfly(&v[0], &v[1], y_inv);
        v[0] = m31_mul(v[0], n_inv); v[1] = m31_mul(v[1], n_inv);
        return 0;
    }
    if (domain_log == 2) {
        m31 x = init.x, y = init.y, nn = 4, xyn_inv = m31_inv(m31_mul(m31_mul(x, y), nn));
        m31 x_inv = m31_mul(m31_mul(xyn_inv, y), nn), y_inv = m31_mul(m31_mul(xyn_inv, x), nn);
        m31 n_inv = m31_mul(m31_mul(xyn_inv, x), y);
        ibutterfly(&v[0], &v[1], y_inv); ibutterfly(&v[2], &v[3], m31_neg(y_inv));
        ibutterfly(&v[0], &v[2], x_inv); ibutterfly(&v[1], &v[3], x_inv);
        for (int i = 0; i < 4; i++) v[i] = m31_mul(v[i], n_inv);
        return 0;
    }
    uint64_t buf_len = 1ull << tw_log;
    fft_layer(v, n, 0, line_twiddles(itw, buf_len, half_log, 0), 1, 1);
    for (int k = 0; k < half_log; k++) fft_layer(v, n, k + 1, line_twiddles(itw, buf_len, half_log, k), 0, 1);
    m31 inv = m31_inv((m31)n);
    for (uint64_t i = 0; i < n; i++) v[i] = m31_mul(v[i], inv);
    return 0;
}

/* backend/cpu/circle.ts:52-69 eval_at_point + poly/utils.ts:36-59 fold.
 * point = (x, y) as two QM31 (4 words each); out = 4 words. */
static qm31 fold_rec(const m31 *vals, uint64_t n, const qm31 *factors)
{
    if (n == 1) return qm_from_m31(vals[0]);
    qm31 l = fold_rec(vals, n / 2, factors + 1), r = fold_rec(vals + n / 2, n / 2, factors + 1);
    return qm_add(l, qm_mul(r, factors[0]));
}
EXPORT int orc_eval_at_point(const uint32_t *coeffs, int log_size, const uint32_t *px, const uint32_t *py, uint32_t *out)
{
    if (log_size == 0) { qm_st(out, qm_from_m31(coeffs[0])); return 0; }
    qm31 maps[64];
    qm31 x = qm_ld(px);
    const qm31 one = qm(1, 0, 0, 0);
    maps[0] = qm_ld(py);
    for (int i = 1; i < log_size; i++) {
        maps[i] = x;
        qm31 sx = qm_mul(x, x);                      /* circle.ts:37-40 double_x = 2x^2 - 1 */
        x = qm_sub(qm_add(sx, sx), one);
    }
    for (int i = 0; i < log_size / 2; i++) { qm31 t = maps[i]; maps[i] = maps[log_size - 1 - i]; maps[log_size - 1 - i] = t; }
    qm_st(out, fold_rec(coeffs, 1ull << log_size, maps));
    return 0;
}

/* ------------------------------------------------------------------ BLAKE2s-256 (RFC 7693; the reference
 * calls @noble/hashes 1.8.0 blake2s at vcs/blake2_hash.ts:1,53-71 and backend/cpu/blake2.ts:37,124) */
static const uint32_t B2S_IV[8] = { 0x6A09E667, 0xBB67AE85, 0x3C6EF372, 0xA54FF53A, 0x510E527F, 0x9B05688C, 0x1F83D9AB, 0x5BE0CD19 };
static const uint8_t B2S_SIGMA[10][16] = {
    { 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15 }, { 14, 10, 4, 8, 9, 15, 13, 6, 1, 12, 0, 2, 11, 7, 5, 3 },
    { 11, 8, 12, 0, 5, 2, 15, 13, 10, 14, 3, 6, 7, 1, 9, 4 }, { 7, 9, 3, 1, 13, 12, 11, 14, 2, 6, 5, 10, 4, 0, 15, 8 },
    { 9, 0, 5, 7, 2, 4, 10, 15, 14, 1, 11, 12, 6, 8, 3, 13 }, { 2, 12, 6, 10, 0, 11, 8, 3, 4, 13, 7, 5, 15, 14, 1, 9 },
    { 12, 5, 1, 15, 14, 13, 4, 10, 0, 7, 6, 3, 9, 2, 8, 11 }, { 13, 11, 7, 14, 12, 1, 3, 9, 5, 0, 15, 4, 8, 6, 2, 10 },
    { 6, 15, 14, 9, 11, 3, 0, 8, 12, 2, 13, 7, 1, 4, 10, 5 }, { 10, 2, 8, 4, 7, 6, 1, 5, 15, 11, 9, 14, 3, 12, 13, 0 } };
static inline uint32_t rotr32(uint32_t x, int n) { return (x >> n) | (x << (32 - n)); }
#define B2S_G(a, b, c, d, x, y) do { a = a + b + (x); d = rotr32(d ^ a, 16); c = c + d; b = rotr32(b ^ c, 12); \
                                      a = a + b + (y); d = rotr32(d ^ a, 8);  c = c + d; b = rotr32(b ^ c, 7); } while (0)
static void b2s_compress(uint32_t h[8], const uint32_t m[16], uint64_t t, int last)
{
    uint32_t v[16];
    for (int i = 0; i < 8; i++) { v[i] = h[i]; v[i + 8] = B2S_IV[i]; }
    v[12] ^= (uint32_t)t; v[13] ^= (uint32_t)(t >> 32);
    if (last) v[14] = ~v[14];
    for (int r = 0; r < 10; r++) {
        const uint8_t *s = B2S_SIGMA[r];
        B2S_G(v[0], v[4], v[8], v[12], m[s[0]], m[s[1]]);   B2S_G(v[1], v[5], v[9], v[13], m[s[2]], m[s[3]]);
        B2S_G(v[2], v[6], v[10], v[14], m[s[4]], m[s[5]]);  B2S_G(v[3], v[7], v[11], v[15], m[s[6]], m[s[7]]);
        B2S_G(v[0], v[5], v[10], v[15], m[s[8]], m[s[9]]);  B2S_G(v[1], v[6], v[11], v[12], m[s[10]], m[s[11]]);
        B2S_G(v[2], v[7], v[8], v[13], m[s[12]], m[s[13]]); B2S_G(v[3], v[4], v[9], v[14], m[s[14]], m[s[15]]);
    }
    for (int i = 0; i < 8; i++) h[i] ^= v[i] ^ v[i + 8];
}
/* Hash a message given as little-endian 32-bit words (all messages on this path are word streams). */
static void b2s_hash_words(const uint32_t *w, uint64_t nwords, uint8_t out[32])
{
    uint32_t h[8], m[16];
    for (int i = 0; i < 8; i++) h[i] = B2S_IV[i];
    h[0] ^= 0x01010020u;                             /* digest 32, no key, fanout 1, depth 1 */
    uint64_t done = 0;
    while (nwords - done > 16) { b2s_compress(h, w + done, (done + 16) * 4, 0); done += 16; }
    uint64_t rem = nwords - done;
    memset(m, 0, sizeof m);
    memcpy(m, w + done, rem * 4);
    b2s_compress(h, m, nwords * 4, 1);
    memcpy(out, h, 32);                              /* little-endian host */
}
EXPORT void orc_blake2s(const uint8_t *msg, uint64_t len, uint8_t out[32])
{
    /* general byte-string entry for the KATs */
    uint32_t h[8], m[16];
    for (int i = 0; i < 8; i++) h[i] = B2S_IV[i];
    h[0] ^= 0x01010020u;
    uint64_t done = 0;
    while (len - done > 64) { memcpy(m, msg + done, 64); b2s_compress(h, m, done + 64, 0); done += 64; }
    memset(m, 0, sizeof m);
    memcpy(m, msg + done, len - done);
    b2s_compress(h, m, len, 1);
    memcpy(out, h, 32);
}
EXPORT void orc_blake2s_compress(uint32_t h[8], const uint32_t m[16], uint64_t t, int last) { b2s_compress(h, m, t, last); }

/* vcs/ops.ts:16-26 commitOnLayer with vcs/blake2_merkle.ts:9-24 hashNode semantics:
 * node_i = blake2s( [prev[2i] || prev[2i+1]] || LE32(col_0[i]) || ... ).  prev may be NULL (leaf layer).
 * cols: n_cols pointers to 2^log_size words.  out: 2^log_size * 32 bytes. */
EXPORT int orc_merkle_commit_layer(int log_size, const uint8_t *prev, const uint32_t *const *cols, int n_cols, uint8_t *out)
{
    int64_t size = 1ll << log_size;
#pragma omp parallel
    {
        uint32_t *msg = (uint32_t *)malloc((16 + (size_t)n_cols) * 4);
#pragma omp for schedule(static)
        for (int64_t i = 0; i < size; i++) {
            uint64_t w = 0;
            if (prev) { memcpy(msg, prev + 64 * i, 64); w = 16; }
            for (int c = 0; c < n_cols; c++) msg[w++] = cols[c][i];
            b2s_hash_words(msg, w, out + 32 * i);
        }
        free(msg);
    }
    return 0;
}
/* vcs/prover.ts:13-30 MerkleProver.commit.  col_logs[c] = log2(len of column c).  layers_out receives the
 * layers concatenated root-first: layer k (2^k hashes) at byte offset 32*(2^k - 1).  Returns max_log. */
EXPORT int orc_merkle_commit(const uint32_t *const *cols, const int *col_logs, int n_cols, uint8_t *layers_out)
{
    if (n_cols == 0) { orc_merkle_commit_layer(0, NULL, NULL, 0, layers_out); return 0; }
    int max_log = 0;
    for (int c = 0; c < n_cols; c++) if (col_logs[c] > max_log) max_log = col_logs[c];
    const uint32_t **layer_cols = (const uint32_t **)malloc(sizeof(void *) * (size_t)n_cols);
    const uint8_t *prev = NULL;
    for (int log = max_log; log >= 0; log--) {
        int k = 0;                                   /* stable: original order among equal lengths (prover.ts:21) */
        for (int c = 0; c < n_cols; c++) if (col_logs[c] == log) layer_cols[k++] = cols[c];
        uint8_t *dst = layers_out + 32ull * ((1ull << log) - 1);
        orc_merkle_commit_layer(log, prev, layer_cols, k, dst);
        prev = dst;
    }
    free(layer_cols);
    return max_log;
}

/* ------------------------------------------------------------------ FRI (backend/cpu/fri.ts, fri.ts:120-192)
 * SecureColumnByCoords layout: 4 coordinate columns (fields/secure_columns.ts:124-217). */
static inline qm31 sc_at(const uint32_t *const c[4], uint64_t i) { return qm(c[0][i], c[1][i], c[2][i], c[3][i]); }
static inline void sc_set(uint32_t *const c[4], uint64_t i, qm31 v) { c[0][i] = v.c0.re; c[1][i] = v.c0.im; c[2][i] = v.c1.re; c[3][i] = v.c1.im; }
static inline void qm_ibutterfly(qm31 *v0, qm31 *v1, m31 it)
{
    qm31 a = *v0, b = *v1;
    *v0 = qm_add(a, b);
    *v1 = qm_mul_m31(qm_sub(a, b), it);
}
/* backend/cpu/fri.ts:23-51 foldLine.  src: 4 x 2^log_size; dst: 4 x 2^(log_size-1).
 * Line domain = coset (init, log_size); x = coset.at(bitrev(2i, log)).x */
EXPORT int orc_fold_line(const uint32_t *s0, const uint32_t *s1, const uint32_t *s2, const uint32_t *s3, int log_size,
                         uint32_t coset_init, const uint32_t *alpha4, uint32_t *d0, uint32_t *d1, uint32_t *d2, uint32_t *d3)
{
    if (log_size < 1) return -2;
    const uint32_t *const src[4] = { s0, s1, s2, s3 };
    uint32_t *const dst[4] = { d0, d1, d2, d3 };
    qm31 alpha = qm_ld(alpha4);
    int64_t n = 1ll << log_size;
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n / 2; i++) {
        qm31 f_x = sc_at(src, 2 * i), f_nx = sc_at(src, 2 * i + 1);
        uint32_t idx = coset_index_at(coset_init, log_size, bit_reverse_index((uint32_t)(i << 1), log_size));
        m31 x = idx_to_point(idx).x;
        qm_ibutterfly(&f_x, &f_nx, m31_inv(x));
        sc_set(dst, i, qm_add(f_x, qm_mul(alpha, f_nx)));
    }
    return 0;
}
/* backend/cpu/fri.ts:61-92 foldCircleIntoLine.  src on circle domain (half coset init, domain log_size);
 * dst (2^(log_size-1)) is read-modify-written: dst = dst * alpha^2 + (f0 + alpha f1). */
EXPORT int orc_fold_circle_into_line(uint32_t *d0, uint32_t *d1, uint32_t *d2, uint32_t *d3,
                                     const uint32_t *s0, const uint32_t *s1, const uint32_t *s2, const uint32_t *s3,
                                     int log_size, uint32_t half_init, const uint32_t *alpha4)
{
    if (log_size < 1) return -2;
    const uint32_t *const src[4] = { s0, s1, s2, s3 };
    uint32_t *const dst[4] = { d0, d1, d2, d3 };
    const uint32_t *const dstc[4] = { d0, d1, d2, d3 };
    qm31 alpha = qm_ld(alpha4), alpha_sq = qm_mul(alpha, alpha);
    int64_t n = 1ll << log_size;
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n / 2; i++) {
        qm31 f_p = sc_at(src, 2 * i), f_np = sc_at(src, 2 * i + 1);
        uint32_t idx = domain_index_at(half_init, log_size - 1, bit_reverse_index((uint32_t)(i << 1), log_size));
        m31 y = idx_to_point(idx).y;
        qm_ibutterfly(&f_p, &f_np, m31_inv(y));
        qm31 f_prime = qm_add(qm_mul(alpha, f_np), f_p);
        sc_set(dst, i, qm_add(qm_mul(sc_at(dstc, i), alpha_sq), f_prime));
    }
    return 0;
}
/* backend/cpu/fri.ts:97-164 decompose: lambda = (sum first half - sum second half) / n; g = f -/+ lambda */
EXPORT int orc_fri_decompose(const uint32_t *s0, const uint32_t *s1, const uint32_t *s2, const uint32_t *s3, int log_size,
                             uint32_t *g0, uint32_t *g1, uint32_t *g2, uint32_t *g3, uint32_t *lambda4)
{
    const uint32_t *const src[4] = { s0, s1, s2, s3 };
    uint32_t *const g[4] = { g0, g1, g2, g3 };
    uint64_t n = 1ull << log_size, half = n / 2;
    qm31 lambda;
    if (n == 1) lambda = qm_sub(qm(0, 0, 0, 0), sc_at(src, 0));
    else {
        qm31 a = qm(0, 0, 0, 0), b = qm(0, 0, 0, 0);
        for (uint64_t i = 0; i < half; i++) a = qm_add(a, sc_at(src, i));
        for (uint64_t i = half; i < n; i++) b = qm_add(b, sc_at(src, i));
        lambda = qm_mul_m31(qm_sub(a, b), m31_inv((m31)n));
    }
    if (n == 1) sc_set(g, 0, qm_sub(sc_at(src, 0), lambda));
    else {
        for (uint64_t i = 0; i < half; i++) sc_set(g, i, qm_sub(sc_at(src, i), lambda));
        for (uint64_t i = half; i < n; i++) sc_set(g, i, qm_add(sc_at(src, i), lambda));
    }
    qm_st(lambda4, lambda);
    return 0;
}

/* ------------------------------------------------------------------ accumulation (backend/cpu/accumulation.ts) */
EXPORT void orc_accumulate(uint32_t *d0, uint32_t *d1, uint32_t *d2, uint32_t *d3,
                           const uint32_t *s0, const uint32_t *s1, const uint32_t *s2, const uint32_t *s3, uint64_t len)
{
    for (uint64_t i = 0; i < len; i++) {             /* accumulation.ts:38-49 */
        d0[i] = m31_add(d0[i], s0[i]); d1[i] = m31_add(d1[i], s1[i]);
        d2[i] = m31_add(d2[i], s2[i]); d3[i] = m31_add(d3[i], s3[i]);
    }
}
EXPORT void orc_generate_secure_powers(const uint32_t *felt4, int n, uint32_t *out)
{
    qm31 acc = qm(1, 0, 0, 0), f = qm_ld(felt4);     /* accumulation.ts:52-63 */
    for (int i = 0; i < n; i++) { qm_st(out + 4 * i, acc); acc = qm_mul(acc, f); }
}

/* Wide-Fibonacci constraint quotients over the evaluation domain.  The constraint is examples/fibonacci.ts:72-91
 * (c - (a^2 + b^2) on three consecutive columns); the loop over the domain is the one the TS code leaves as a placeholder
 * (constraint_framework/index.ts:131-152) and Stwo's FrameworkComponent::evaluate_constraint_quotients_on_domain does:
 * per row, row_res = sum_i coeffs[i] * constraint_i(row), then col[row] += row_res * denom_inv[row >> trace_log_size].
 * parity unpinned: no reference golden exists for this function. */
EXPORT void orc_widefib_constraints(int trace_log_size, int eval_log_size, const uint32_t *const *cols, int n_cols,
                                    const uint32_t *coeffs, const uint32_t *denom_inv,
                                    uint32_t *o0, uint32_t *o1, uint32_t *o2, uint32_t *o3)
{
    const uint64_t n = 1ull << eval_log_size;
    for (uint64_t row = 0; row < n; row++) {
        qm31 res = qm(0, 0, 0, 0);
        for (int i = 0; i + 2 < n_cols; i++) {
            uint32_t a = cols[i][row], b = cols[i + 1][row], c = cols[i + 2][row];
            uint32_t con = m31_sub(c, m31_add(m31_mul(a, a), m31_mul(b, b)));
            res = qm_add(res, qm_mul_m31(qm_ld(coeffs + 4 * i), con));
        }
        res = qm_mul_m31(res, denom_inv[row >> trace_log_size]);
        o0[row] = m31_add(o0[row], res.c0.re); o1[row] = m31_add(o1[row], res.c0.im);
        o2[row] = m31_add(o2[row], res.c1.re); o3[row] = m31_add(o3[row], res.c1.im);
    }
}

/* backend/cpu/quotients.ts:52-178 accumulateQuotients.
 * Batches are flattened: batch b has point (px,py: 8 words at points[8b]) and entries
 * [batch_offsets[b], batch_offsets[b+1]) of (col_index[], values[4*]).  Follows the TS code, including its
 * divergence from Stwo: denominatorInverses uses only the M31 parts c0.real / c0.imag (quotients.ts:160-178). */
static int g_quotient_ts_denominators = 1;
/* 1 (default): denominators as the TS code computes them; 0: Stwo's CM31 denominators (Rust `denominator_inverses`:
 * (Re(p.x) - d.x) * Im(p.y) - (Re(p.y) - d.y) * Im(p.x) over the CM31 halves of the sample point), for which quotients are low degree. */
EXPORT void orc_set_quotient_ts_denominators(int on) { g_quotient_ts_denominators = on; }
EXPORT int orc_accumulate_quotients(int log_size, uint32_t half_init, const uint32_t *const *cols, int n_cols,
                                    const uint32_t *random_coeff4, int n_batches, const uint32_t *points,
                                    const int *batch_offsets, const int *col_index, const uint32_t *values,
                                    uint32_t *o0, uint32_t *o1, uint32_t *o2, uint32_t *o3)
{
    (void)n_cols;
    uint32_t *const out[4] = { o0, o1, o2, o3 };
    qm31 rc = qm_ld(random_coeff4);
    int total = batch_offsets[n_batches];
    qm31 *la = (qm31 *)malloc(sizeof(qm31) * (size_t)(total + 1)), *lb = (qm31 *)malloc(sizeof(qm31) * (size_t)(total + 1)),
         *lc = (qm31 *)malloc(sizeof(qm31) * (size_t)(total + 1)), *bc = (qm31 *)malloc(sizeof(qm31) * (size_t)(n_batches + 1));
    for (int b = 0; b < n_batches; b++) {
        qm31 px = qm_ld(points + 8 * b), py = qm_ld(points + 8 * b + 4);
        (void)px;
        qm31 alpha = qm(1, 0, 0, 0), pw = qm(1, 0, 0, 0);
        for (int j = batch_offsets[b]; j < batch_offsets[b + 1]; j++) {
            alpha = qm_mul(alpha, rc);               /* quotients.ts:124-139 columnLineCoeffs */
            pw = qm_mul(pw, rc);
            qm31 v = qm_ld(values + 4 * j);
            /* constraints.ts:117-129 complexConjugateLineCoeffs; the TS QM31.complexConjugate conjugates both CM31 halves
             * (qm31.ts:433-435) where Stwo maps u -> -u, i.e. (c0, -c1) */
            qm31 cv = g_quotient_ts_denominators ? qm_conj(v) : qm(v.c0.re, v.c0.im, m31_neg(v.c1.re), m31_neg(v.c1.im));
            qm31 cpy = g_quotient_ts_denominators ? qm_conj(py) : qm(py.c0.re, py.c0.im, m31_neg(py.c1.re), m31_neg(py.c1.im));
            qm31 a = qm_sub(cv, v);
            qm31 c = qm_sub(cpy, py);
            qm31 bb = qm_sub(qm_mul(v, c), qm_mul(a, py));
            la[j] = qm_mul(alpha, a); lb[j] = qm_mul(alpha, bb); lc[j] = qm_mul(alpha, c);
        }
        bc[b] = pw;                                  /* quotients.ts:147-152 random_coeff^len */
    }
    int64_t n = 1ll << log_size;
#pragma omp parallel for schedule(static)
    for (int64_t row = 0; row < n; row++) {
        uint32_t idx = domain_index_at(half_init, log_size - 1, bit_reverse_index((uint32_t)row, log_size));
        cpoint dp = idx_to_point(idx);
        qm31 acc = qm(0, 0, 0, 0);
        for (int b = 0; b < n_batches; b++) {
            const uint32_t *pt = points + 8 * b;
            m31 prx = pt[0], pix = pt[1], pry = pt[4], piy = pt[5];
            m31 den = m31_sub(m31_mul(m31_sub(prx, dp.x), piy), m31_mul(m31_sub(pry, dp.y), pix));
            cm31 den_inv = cm(m31_inv(den), 0);
            if (!g_quotient_ts_denominators) {
                cm31 rx = cm(m31_sub(pt[0], dp.x), pt[1]), ix = cm(pt[2], pt[3]), ry = cm(m31_sub(pt[4], dp.y), pt[5]), iy = cm(pt[6], pt[7]);
                den_inv = cm_inv(cm_sub(cm_mul(rx, iy), cm_mul(ry, ix)));
            }
            qm31 num = qm(0, 0, 0, 0);
            for (int j = batch_offsets[b]; j < batch_offsets[b + 1]; j++) {
                qm31 value = qm_mul(qm_from_m31(cols[col_index[j]][row]), lc[j]);
                qm31 lin = qm_add(qm_mul(la[j], qm_from_m31(dp.y)), lb[j]);
                num = qm_add(num, qm_sub(value, lin));
            }
            acc = qm_add(qm_mul(acc, bc[b]), qm_mul_cm31(num, den_inv));
        }
        out[0][row] = acc.c0.re; out[1][row] = acc.c0.im; out[2][row] = acc.c1.re; out[3][row] = acc.c1.im;
    }
    free(la); free(lb); free(lc); free(bc);
    return 0;
}

/* ------------------------------------------------------------------ LDE + commit pipeline (the metric)
 * Stwo's CommitmentTreeProver::new as quoted at pcs/prover.ts:215-235 / poly/circle/ops.ts:73-101:
 * interpolate every column on CanonicCoset(log_size), evaluate on CanonicCoset(log_size+log_blowup),
 * MerkleProver.commit.  cols_in: n_cols x 2^log_size (bit-reversed evaluations), evals_out: n_cols x
 * 2^(log_size+log_blowup), coeffs_out (may be NULL), layers_out as orc_merkle_commit.  tw/itw: twiddle tree of
 * the half coset of the LDE domain (root log = log_size+log_blowup-1).  OpenMP over columns / rows. */
EXPORT int orc_lde_commit(const uint32_t *cols_in, int n_cols, int log_size, int log_blowup, const uint32_t *tw,
                          const uint32_t *itw, uint32_t *coeffs_out, uint32_t *evals_out, uint8_t *layers_out, int quirks)
{
    int ext_log = log_size + log_blowup, tw_log = ext_log - 1;
    uint32_t tw_root_init = subgroup_gen(tw_log + 2);          /* half_odds(tw_log): canonic.ts half_coset */
    uint32_t trace_half_init = subgroup_gen(log_size - 1 + 2), ext_half_init = tw_root_init;
    uint64_t n = 1ull << log_size, ne = 1ull << ext_log;
    int rc = 0;
#pragma omp parallel for schedule(dynamic)
    for (int c = 0; c < n_cols; c++) {
        uint32_t *tmp = coeffs_out ? coeffs_out + (uint64_t)c * n : (uint32_t *)malloc(n * 4);
        memcpy(tmp, cols_in + (uint64_t)c * n, n * 4);
        int r1 = orc_interpolate(tmp, log_size, trace_half_init, tw_root_init, tw_log, itw, quirks);
        int r2 = orc_evaluate(tmp, log_size, evals_out + (uint64_t)c * ne, ext_log, ext_half_init, tw_root_init, tw_log, tw, quirks);
        if (r1 || r2) rc = r1 ? r1 : r2;
        if (!coeffs_out) free(tmp);
    }
    if (rc) return rc;
    if (layers_out) {
        const uint32_t **cols = (const uint32_t **)malloc(sizeof(void *) * (size_t)n_cols);
        int *logs = (int *)malloc(sizeof(int) * (size_t)n_cols);
        for (int c = 0; c < n_cols; c++) { cols[c] = evals_out + (uint64_t)c * ne; logs[c] = ext_log; }
        orc_merkle_commit(cols, logs, n_cols, layers_out);
        free(cols); free(logs);
    }
    return 0;
}
/* torchrun exports OMP_NUM_THREADS=1 to its workers; the timed CPU baseline asks for the cores it may use explicitly. */
EXPORT void orc_set_num_threads(int n)
{
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}
EXPORT int orc_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
