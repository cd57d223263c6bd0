// M31 / CM31 / QM31 device arithmetic for sm_100a.
// Semantics follow packages/core/src/fields/{m31,cm31,qm31}.ts of the reference: canonical residues in
// [0, P), P = 2^31 - 1.  Every function takes and returns canonical values, so any algebraically
// equivalent evaluation order is bit-exact with the reference's BigInt arithmetic.
//
// Instruction budget (checked with cuobjdump, see profiles/): add/sub = IADD + VIADDMNMX.U32,
// mul by a pre-doubled twiddle = IMAD.WIDE.U32 + LEA.HI + VIADDMNMX.U32  => 7 instr / butterfly.
#pragma once
#include <cstdint>

namespace m31 {

constexpr uint32_t P = 0x7fffffffu;
constexpr uint32_t NEG_P = 0x80000001u;   // -P mod 2^32

// r in [0, 2P) -> [0, P)   (single VIADDMNMX.U32: min(r - P, r) as unsigned)
__device__ __forceinline__ uint32_t csub(uint32_t r) { return __viaddmin_u32(r, NEG_P, r); }

__device__ __forceinline__ uint32_t add(uint32_t a, uint32_t b) { return csub(a + b); }
// a - b in (-P, P): min(d + P, d) as unsigned picks the in-range one
__device__ __forceinline__ uint32_t sub(uint32_t a, uint32_t b) { uint32_t d = a - b; return __viaddmin_u32(d, P, d); }
__device__ __forceinline__ uint32_t neg(uint32_t a) { return a == 0 ? 0u : P - a; }

// a * t where t2 = 2*t (pre-doubled, < 2^32).  hi = (a*t) >> 31, lo>>1 = (a*t) & P.
__device__ __forceinline__ uint32_t mul2(uint32_t a, uint32_t t2)
{
    uint64_t p = (uint64_t)a * t2;
    uint32_t s = (uint32_t)(p >> 32) + ((uint32_t)p >> 1);   // LEA.HI
    return csub(s);
}
__device__ __forceinline__ uint32_t mul(uint32_t a, uint32_t b) { return mul2(a, b + b); }
__device__ __forceinline__ uint32_t sqr(uint32_t a) { return mul2(a, a + a); }

__device__ __forceinline__ uint32_t sqn(uint32_t v, int n)
{
    for (int i = 0; i < n; i++) v = sqr(v);
    return v;
}
// fields/m31.ts:305-326 pow2147483645 (x^(P-2)); inv(0) is left as 0 (the reference throws).
__device__ __forceinline__ uint32_t inv(uint32_t v)
{
    uint32_t t0 = mul(sqn(v, 2), v);
    uint32_t t1 = mul(sqn(t0, 1), t0);
    uint32_t t2 = mul(sqn(t1, 3), t0);
    uint32_t t3 = mul(sqn(t2, 1), t0);
    uint32_t t4 = mul(sqn(t3, 8), t3);
    uint32_t t5 = mul(sqn(t4, 8), t3);
    return mul(sqn(t5, 7), t2);
}

// fft.ts:12-17 butterfly: (v0 + v1 t, v0 - v1 t), t2 = 2t
__device__ __forceinline__ void butterfly(uint32_t &v0, uint32_t &v1, uint32_t t2)
{
    uint32_t m = mul2(v1, t2), a = v0;
    v0 = add(a, m);
    v1 = sub(a, m);
}
// fft.ts:25-30 ibutterfly: (v0 + v1, (v0 - v1) it), it2 = 2 it
__device__ __forceinline__ void ibutterfly(uint32_t &v0, uint32_t &v1, uint32_t it2)
{
    uint32_t a = v0, b = v1;
    v0 = add(a, b);
    v1 = mul2(sub(a, b), it2);
}

// Pipe-balanced butterflies for the FFT passes.  On sm_100 ptxas turns every two-input add into IMAD.IADD (FMA pipe),
// which is also where IMAD.WIDE (~6 issue cycles) lives; the ALU pipe (LEA.HI, VIADDMNMX) is then under-used
// (profiles/r01).  `z` is a run-time zero (kernel argument): a + b + z can only be an IADD3, i.e. an ALU-pipe add.
__device__ __forceinline__ void butterfly_z(uint32_t &v0, uint32_t &v1, uint32_t t2, uint32_t z)
{
    uint32_t m = mul2(v1, t2), a = v0;
    v0 = csub(a + m + z);
    v1 = sub(a, m);
}
__device__ __forceinline__ void ibutterfly_z(uint32_t &v0, uint32_t &v1, uint32_t it2, uint32_t z)
{
    uint32_t a = v0, b = v1;
    v0 = csub(a + b + z);
    v1 = mul2(sub(a, b), it2);
}

// ---- CM31 = M31[i]/(i^2+1)  (fields/cm31.ts) ----
struct CM { uint32_t re, im; };
__device__ __forceinline__ CM cadd(CM a, CM b) { return { add(a.re, b.re), add(a.im, b.im) }; }
__device__ __forceinline__ CM csub2(CM a, CM b) { return { sub(a.re, b.re), sub(a.im, b.im) }; }
__device__ __forceinline__ CM cmul(CM a, CM b)
{
    return { sub(mul(a.re, b.re), mul(a.im, b.im)), add(mul(a.re, b.im), mul(a.im, b.re)) };
}
__device__ __forceinline__ CM cmul_m(CM a, uint32_t b) { uint32_t b2 = b + b; return { mul2(a.re, b2), mul2(a.im, b2) }; }

// ---- QM31 = CM31[u]/(u^2 - 2 - i)  (fields/qm31.ts; R = 2 + i at qm31.ts:9) ----
struct QM { uint32_t a, b, c, d; };   // (a + bi) + (c + di)u  == to_m31_array order (qm31.ts:139-141)
__device__ __forceinline__ QM qadd(QM x, QM y) { return { add(x.a, y.a), add(x.b, y.b), add(x.c, y.c), add(x.d, y.d) }; }
__device__ __forceinline__ QM qsub(QM x, QM y) { return { sub(x.a, y.a), sub(x.b, y.b), sub(x.c, y.c), sub(x.d, y.d) }; }
__device__ __forceinline__ QM qmul_m(QM x, uint32_t m)
{
    uint32_t m2 = m + m;
    return { mul2(x.a, m2), mul2(x.b, m2), mul2(x.c, m2), mul2(x.d, m2) };
}
__device__ __forceinline__ QM qmul(QM x, QM y)
{
    CM x0 = { x.a, x.b }, x1 = { x.c, x.d }, y0 = { y.a, y.b }, y1 = { y.c, y.d };
    CM bd = cmul(x1, y1);
    // R * bd = (2 + i)(p + qi) = (2p - q) + (p + 2q) i
    CM rbd = { sub(add(bd.re, bd.re), bd.im), add(bd.re, add(bd.im, bd.im)) };
    CM r0 = cadd(cmul(x0, y0), rbd);
    CM r1 = cadd(cmul(x0, y1), cmul(x1, y0));
    return { r0.re, r0.im, r1.re, r1.im };
}
__device__ __forceinline__ QM qmul_cm(QM x, CM y)
{
    CM r0 = cmul({ x.a, x.b }, y), r1 = cmul({ x.c, x.d }, y);
    return { r0.re, r0.im, r1.re, r1.im };
}
__device__ __forceinline__ QM qfrom(uint32_t m) { return { m, 0, 0, 0 }; }
__device__ __forceinline__ CM cneg(CM a) { return { neg(a.re), neg(a.im) }; }
__device__ __forceinline__ QM qneg(QM x) { return { neg(x.a), neg(x.b), neg(x.c), neg(x.d) }; }
// fields/cm31.ts:237-251: (a - bi) / (a^2 + b^2); inverse of 0 is left as 0 (the reference throws)
__device__ __forceinline__ CM cinv(CM a) { return cmul_m({ a.re, neg(a.im) }, inv(add(sqr(a.re), sqr(a.im)))); }
// fields/qm31.ts:411-428: with b2 = c1^2, denom = c0^2 - (2 b2 + i b2) = c0^2 - R b2; 1/(c0 + c1 u) = (c0 - c1 u) / denom
__device__ __forceinline__ QM qinv(QM x)
{
    const CM c0 = { x.a, x.b }, c1 = { x.c, x.d };
    const CM b2 = cmul(c1, c1);
    const CM rb2 = { sub(add(b2.re, b2.re), b2.im), add(b2.re, add(b2.im, b2.im)) };
    const CM di = cinv(csub2(cmul(c0, c0), rb2));
    const CM r0 = cmul(c0, di), r1 = cneg(cmul(c1, di));
    return { r0.re, r0.im, r1.re, r1.im };
}
// fields/m31.ts:39-44 partialReduce ([0, 2P) -> [0, P)) and :60-101 reduce (any u64): ((((v >> 31) + v + 1) >> 31) + v) & P
__device__ __forceinline__ uint32_t partial_reduce(uint32_t v) { return v >= P ? v - P : v; }
__device__ __forceinline__ uint32_t reduce(uint64_t v)
{
    // v < 2^64: fold 31 bits at a time (2^31 = 1 mod P), then the reference's closing step on a value < 2^33
    uint64_t s = (v & P) + ((v >> 31) & P) + (v >> 62);
    return (uint32_t)(((((s >> 31) + s + 1) >> 31) + s) & P);
}

// ---- circle group (circle.ts:102-106) ----
struct Pt { uint32_t x, y; };
__device__ __forceinline__ Pt padd(Pt p, Pt q)
{
    return { sub(mul(p.x, q.x), mul(p.y, q.y)), add(mul(p.x, q.y), mul(p.y, q.x)) };
}

__device__ __forceinline__ uint32_t bitrev(uint32_t i, int log_n) { return log_n == 0 ? 0u : (__brev(i) >> (32 - log_n)); }

}  // namespace m31
