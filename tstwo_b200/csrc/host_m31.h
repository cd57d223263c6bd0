// Host-side M31 / circle-group scalar helpers used by the launchers (twiddle table seeds, closed-form sizes).
#pragma once
#include <cstdint>

namespace tstwo { namespace host {

constexpr uint32_t P = 0x7fffffffu;
inline uint32_t add(uint32_t a, uint32_t b) { uint32_t s = a + b; return s >= P ? s - P : s; }
inline uint32_t sub(uint32_t a, uint32_t b) { return a >= b ? a - b : a + P - b; }
inline uint32_t neg(uint32_t a) { return a ? P - a : 0; }
inline uint32_t mul(uint32_t a, uint32_t b) { return (uint32_t)(((uint64_t)a * b) % P); }
inline uint32_t pow(uint32_t a, uint32_t e) { uint32_t r = 1; while (e) { if (e & 1) r = mul(r, a); a = mul(a, a); e >>= 1; } return r; }
inline uint32_t inv(uint32_t a) { return pow(a, P - 2); }
struct Pt { uint32_t x, y; };
inline Pt padd(Pt p, Pt q) { return { sub(mul(p.x, q.x), mul(p.y, q.y)), add(mul(p.x, q.y), mul(p.y, q.x)) }; }
// circle.ts:137 M31_CIRCLE_GEN; circle.ts:172-174 CirclePointIndex.to_point
inline Pt index_to_point(uint32_t index)
{
    Pt res = { 1, 0 }, cur = { 2u, 1268011823u };
    index &= 0x7fffffffu;
    while (index) { if (index & 1) res = padd(res, cur); cur = padd(cur, cur); index >>= 1; }
    return res;
}
inline uint32_t subgroup_gen(int log_size) { return log_size == 0 ? 0u : (1u << (31 - log_size)); }

}}  // namespace tstwo::host

// ---- host QM31 (fields/qm31.ts) for the constants of the elementwise launchers ----
namespace tstwo { namespace host {
struct CM { uint32_t re, im; };
struct QM { uint32_t a, b, c, d; };
inline CM cadd(CM x, CM y) { return { add(x.re, y.re), add(x.im, y.im) }; }
inline CM cmul(CM x, CM y) { return { sub(mul(x.re, y.re), mul(x.im, y.im)), add(mul(x.re, y.im), mul(x.im, y.re)) }; }
inline QM qadd(QM x, QM y) { return { add(x.a, y.a), add(x.b, y.b), add(x.c, y.c), add(x.d, y.d) }; }
inline QM qsub(QM x, QM y) { return { sub(x.a, y.a), sub(x.b, y.b), sub(x.c, y.c), sub(x.d, y.d) }; }
inline QM qmul(QM x, QM y)
{
    CM x0 = { x.a, x.b }, x1 = { x.c, x.d }, y0 = { y.a, y.b }, y1 = { y.c, y.d };
    CM r0 = cadd(cmul(x0, y0), cmul(cmul({ 2, 1 }, x1), y1));
    CM r1 = cadd(cmul(x0, y1), cmul(x1, y0));
    return { r0.re, r0.im, r1.re, r1.im };
}
inline QM qconj(QM x) { return { x.a, neg(x.b), x.c, neg(x.d) }; }
inline QM qmul_m(QM x, uint32_t m) { return { mul(x.a, m), mul(x.b, m), mul(x.c, m), mul(x.d, m) }; }
inline QM qld(const uint32_t *p) { return { p[0], p[1], p[2], p[3] }; }
inline void qst(uint32_t *p, QM q) { p[0] = q.a; p[1] = q.b; p[2] = q.c; p[3] = q.d; }
}}  // namespace tstwo::host
