// Elementwise QM31 kernels for sm_100a: FriOps (fold_line, fold_circle_into_line, decompose), AccumulationOps,
// QuotientOps and eval_at_point of the reference's CpuBackend
// (packages/core/src/backend/cpu/{fri,accumulation,quotients}.ts, backend/cpu/circle.ts:52-69).
// SecureField columns are 4 coordinate columns (SecureColumnByCoords); all kernels are HBM-bound streams with
// uint2/uint4 vector access where pairs are adjacent.
#include <cstring>
#include <vector>

#include "circle_dev.cuh"
#include "common.cuh"

namespace tstwo {

cudaError_t elementwise_init() { return upload_gen_table(); }

struct Q4 { uint32_t *c[4]; };
struct CQ4 { const uint32_t *c[4]; };

// Inverse twiddle of pair i, pre-doubled.  mode 0: direct table it2[i]; mode 1: circle twiddle derived from the
// line table T1 (backend/cpu/circle.ts:270-278): +-T1[(i>>1)^1], sign = bit0(i ^ (i>>1)).
__device__ __forceinline__ uint32_t pair_itw2(const uint32_t *tab, int mode, uint64_t i)
{
    if (mode == 0) return __ldg(tab + i);
    uint32_t t = __ldg(tab + ((i >> 1) ^ 1));
    return ((i ^ (i >> 1)) & 1) ? 0xfffffffeu - t : t;
}

// fri.ts:120-152 / backend/cpu/fri.ts:23-51: out[i] = f0 + alpha f1, (f0, f1) = ibutterfly(f(x), f(-x), 1/x)
// alpha_dev != nullptr: the folding coefficient is read from device memory (written by fri_channel_kernel: the transcript stays
// on the device between the layers of tstwo_fri_commit_layers)
// V = outputs per thread: 2 when the columns allow 16-byte loads / 8-byte stores (n_out even, aligned), else 1.
template <int V>
__global__ void __launch_bounds__(256) fold_line_kernel(CQ4 src, Q4 dst, uint64_t n_out, const uint32_t *itab, int mode, m31::QM alpha,
                                                        const uint32_t *alpha_dev = nullptr)
{
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x, i = V * t;
    if (i >= n_out) return;
    if (alpha_dev) alpha = { alpha_dev[0], alpha_dev[1], alpha_dev[2], alpha_dev[3] };
    uint32_t it2[V];
#pragma unroll
    for (int v = 0; v < V; v++) it2[v] = pair_itw2(itab, mode, i + v);
    uint32_t f0[V][4], f1[V][4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        if (V == 2) {
            const uint4 p = __ldcs(reinterpret_cast<const uint4 *>(src.c[k]) + t);
            f0[0][k] = p.x; f1[0][k] = p.y; f0[V - 1][k] = p.z; f1[V - 1][k] = p.w;
        } else {
            const uint2 p = __ldcs(reinterpret_cast<const uint2 *>(src.c[k]) + t);
            f0[0][k] = p.x; f1[0][k] = p.y;
        }
#pragma unroll
        for (int v = 0; v < V; v++) m31::ibutterfly(f0[v][k], f1[v][k], it2[v]);
    }
    m31::QM r[V];
#pragma unroll
    for (int v = 0; v < V; v++)
        r[v] = m31::qadd({ f0[v][0], f0[v][1], f0[v][2], f0[v][3] }, m31::qmul(alpha, { f1[v][0], f1[v][1], f1[v][2], f1[v][3] }));
    if (V == 2) {
        reinterpret_cast<uint2 *>(dst.c[0])[t] = make_uint2(r[0].a, r[V - 1].a); reinterpret_cast<uint2 *>(dst.c[1])[t] = make_uint2(r[0].b, r[V - 1].b);
        reinterpret_cast<uint2 *>(dst.c[2])[t] = make_uint2(r[0].c, r[V - 1].c); reinterpret_cast<uint2 *>(dst.c[3])[t] = make_uint2(r[0].d, r[V - 1].d);
    } else {
        dst.c[0][i] = r[0].a; dst.c[1][i] = r[0].b; dst.c[2][i] = r[0].c; dst.c[3][i] = r[0].d;
    }
}

// fri.ts:162-192 / backend/cpu/fri.ts:61-92: dst[i] = dst[i] alpha^2 + (alpha f1 + f0), twiddle 1/p.y
template <int V>
__global__ void __launch_bounds__(256) fold_circle_kernel(CQ4 src, Q4 dst, uint64_t n_out, const uint32_t *itab, int mode, m31::QM alpha,
                                                          m31::QM alpha_sq, const uint32_t *alpha_dev = nullptr)
{
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x, i = V * t;
    if (i >= n_out) return;
    if (alpha_dev) { alpha = { alpha_dev[0], alpha_dev[1], alpha_dev[2], alpha_dev[3] }; alpha_sq = m31::qmul(alpha, alpha); }
    uint32_t it2[V];
#pragma unroll
    for (int v = 0; v < V; v++) it2[v] = pair_itw2(itab, mode, i + v);
    uint32_t f0[V][4], f1[V][4], cur[V][4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        if (V == 2) {
            const uint4 p = __ldcs(reinterpret_cast<const uint4 *>(src.c[k]) + t);
            f0[0][k] = p.x; f1[0][k] = p.y; f0[V - 1][k] = p.z; f1[V - 1][k] = p.w;
            const uint2 c = reinterpret_cast<const uint2 *>(dst.c[k])[t];
            cur[0][k] = c.x; cur[V - 1][k] = c.y;
        } else {
            const uint2 p = __ldcs(reinterpret_cast<const uint2 *>(src.c[k]) + t);
            f0[0][k] = p.x; f1[0][k] = p.y;
            cur[0][k] = dst.c[k][i];
        }
#pragma unroll
        for (int v = 0; v < V; v++) m31::ibutterfly(f0[v][k], f1[v][k], it2[v]);
    }
    m31::QM r[V];
#pragma unroll
    for (int v = 0; v < V; v++) {
        const m31::QM fp = m31::qadd(m31::qmul(alpha, { f1[v][0], f1[v][1], f1[v][2], f1[v][3] }), { f0[v][0], f0[v][1], f0[v][2], f0[v][3] });
        r[v] = m31::qadd(m31::qmul({ cur[v][0], cur[v][1], cur[v][2], cur[v][3] }, alpha_sq), fp);
    }
    if (V == 2) {
        reinterpret_cast<uint2 *>(dst.c[0])[t] = make_uint2(r[0].a, r[V - 1].a); reinterpret_cast<uint2 *>(dst.c[1])[t] = make_uint2(r[0].b, r[V - 1].b);
        reinterpret_cast<uint2 *>(dst.c[2])[t] = make_uint2(r[0].c, r[V - 1].c); reinterpret_cast<uint2 *>(dst.c[3])[t] = make_uint2(r[0].d, r[V - 1].d);
    } else {
        dst.c[0][i] = r[0].a; dst.c[1][i] = r[0].b; dst.c[2][i] = r[0].c; dst.c[3][i] = r[0].d;
    }
}
// two outputs per thread when every column allows it
static bool fold_vec2(const CQ4 &s, const Q4 &d, uint64_t n_out)
{
    if (n_out < 2 || (n_out & 1)) return false;
    for (int k = 0; k < 4; k++)
        if ((reinterpret_cast<uintptr_t>(s.c[k]) & 15) || (reinterpret_cast<uintptr_t>(d.c[k]) & 7)) return false;
    return true;
}
#define LAUNCH_FOLD(kernel, s, d, n_out, ...)                                                                          \
    do {                                                                                                               \
        if (fold_vec2(s, d, n_out)) TS_LAUNCH(kernel<2>, (unsigned)((n_out / 2 + 255) / 256), 256, 0, s, d, n_out, __VA_ARGS__); \
        else TS_LAUNCH(kernel<1>, (unsigned)((n_out + 255) / 256), 256, 0, s, d, n_out, __VA_ARGS__);                 \
    } while (0)

// backend/cpu/fri.ts:97-123 decompositionCoefficient: 8 sums (4 coordinates x 2 halves) as u64
__global__ void half_sums_kernel(CQ4 src, uint64_t n, unsigned long long *sums)
{
    __shared__ unsigned long long sh[8];
    if (threadIdx.x < 8) sh[threadIdx.x] = 0;
    __syncthreads();
    unsigned long long acc[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
    const uint64_t half = n / 2;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        const int hi = i >= half;
#pragma unroll
        for (int k = 0; k < 4; k++) acc[2 * k + hi] += src.c[k][i];
    }
#pragma unroll
    for (int k = 0; k < 8; k++) {
        unsigned long long v = acc[k];
        for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0) atomicAdd(&sh[k], v);
    }
    __syncthreads();
    if (threadIdx.x < 8) atomicAdd(&sums[threadIdx.x], sh[threadIdx.x]);
}
// the same for n >= 8 with 16-byte loads (a uint4 never straddles the halves), 4 loads in flight per thread
__global__ void __launch_bounds__(256) half_sums_vec_kernel(CQ4 src, uint64_t n4, unsigned long long *sums)
{
    __shared__ unsigned long long sh[8];
    if (threadIdx.x < 8) sh[threadIdx.x] = 0;
    __syncthreads();
    unsigned long long acc[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
    const uint64_t half4 = n4 / 2;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (uint64_t)gridDim.x * blockDim.x) {
        const int hi = i >= half4;
        uint4 v[4];
#pragma unroll
        for (int k = 0; k < 4; k++) v[k] = __ldcs(reinterpret_cast<const uint4 *>(src.c[k]) + i);
#pragma unroll
        for (int k = 0; k < 4; k++) acc[2 * k + hi] += (unsigned long long)v[k].x + v[k].y + v[k].z + v[k].w;
    }
#pragma unroll
    for (int k = 0; k < 8; k++) {
        unsigned long long v = acc[k];
        for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0) atomicAdd(&sh[k], v);
    }
    __syncthreads();
    if (threadIdx.x < 8) atomicAdd(&sums[threadIdx.x], sh[threadIdx.x]);
}
__global__ void __launch_bounds__(256) decompose_apply_vec_kernel(CQ4 src, Q4 g, uint64_t n4, m31::QM lambda)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n4) return;
    const bool second = i >= n4 / 2;
    const uint32_t lam[4] = { lambda.a, lambda.b, lambda.c, lambda.d };
    uint4 v[4];
#pragma unroll
    for (int k = 0; k < 4; k++) v[k] = __ldcs(reinterpret_cast<const uint4 *>(src.c[k]) + i);
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const uint32_t l = second ? lam[k] : m31::sub(0, lam[k]);     // f + lambda on the second half, f - lambda on the first
        reinterpret_cast<uint4 *>(g.c[k])[i] = make_uint4(m31::add(v[k].x, l), m31::add(v[k].y, l), m31::add(v[k].z, l), m31::add(v[k].w, l));
    }
}
// backend/cpu/fri.ts:133-164: g = f - lambda on the first half, f + lambda on the second
__global__ void decompose_apply_kernel(CQ4 src, Q4 g, uint64_t n, m31::QM lambda)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const bool second = n > 1 && i >= n / 2;
    const uint32_t lam[4] = { lambda.a, lambda.b, lambda.c, lambda.d };
#pragma unroll
    for (int k = 0; k < 4; k++) {
        uint32_t f = src.c[k][i];
        g.c[k][i] = second ? m31::add(f, lam[k]) : m31::sub(f, lam[k]);
    }
}

// backend/cpu/accumulation.ts:38-49
__global__ void accumulate_kernel(Q4 dst, CQ4 src, uint64_t n)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
#pragma unroll
    for (int k = 0; k < 4; k++) dst.c[k][i] = m31::add(dst.c[k][i], src.c[k][i]);
}
__global__ void __launch_bounds__(256) accumulate_vec_kernel(Q4 dst, CQ4 src, uint64_t n4)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n4) return;
    uint4 a[4], b[4];
#pragma unroll
    for (int k = 0; k < 4; k++) { a[k] = reinterpret_cast<const uint4 *>(dst.c[k])[i]; b[k] = __ldcs(reinterpret_cast<const uint4 *>(src.c[k]) + i); }
#pragma unroll
    for (int k = 0; k < 4; k++)
        reinterpret_cast<uint4 *>(dst.c[k])[i] = make_uint4(m31::add(a[k].x, b[k].x), m31::add(a[k].y, b[k].y), m31::add(a[k].z, b[k].z), m31::add(a[k].w, b[k].w));
}

// backend/cpu/quotients.ts:52-116.  One thread per row.  The reference evaluates, per sampled column j of a batch,
//   c_j f_j(row) - (a_j y + b_j)            (a_j, b_j, c_j: QM31 line coefficients, quotients.ts:124-152)
// and sums over j.  Field arithmetic is exact, so the sum is regrouped as  sum_j c_j f_j  -  (y A + B)  with A = sum a_j and
// B = sum b_j computed once on the host: the inner loop is then 4 multiply-accumulates per column value instead of ~50
// instructions, which turns the kernel from integer-bound (6.3 ms for 256 x 2^22) into an HBM stream.  The 4 sums are kept as
// u64 (products < 2^62) and folded to < 2^33 every third term.  consts: c_j at cvals[4j..]; per batch (A, B) at ab[8b..], the
// batch coefficient at bcoef[4b..] and the sample point (x[4], y[4]) at pts[8b..].  TS_DENOM: the denominator as the TS code
// computes it, from the M31 parts x.c0.re, x.c0.im, y.c0.re, y.c0.im only (quotients.ts:160-178) -- bit-exact with CpuBackend but
// not a quotient by the line through the point and its conjugate; otherwise Stwo's CM31 denominator
// (Re(p.x) - d.x) Im(p.y) - (Re(p.y) - d.y) Im(p.x) with Re/Im the CM31 halves, which is what makes the quotient low degree.
// Domain point of bit-reversed position j of a circle domain with 2^half_log points per half (CircleDomain.at,
// poly/circle/domain.ts:76-98) from host-built digit tables: T[0][v] = point(half_init + step v), T[d][v] = point(step v 256^d),
// so that the point is at most 3 group additions away instead of a 31-step divergent ladder per row.
constexpr int PT_DIGITS = 4, PT_WORDS = PT_DIGITS * 256 * 2;
__device__ __forceinline__ m31::Pt table_point(const uint2 *__restrict__ T, int half_log, uint32_t j)
{
    const uint32_t jm = j & ((1u << half_log) - 1u);
    uint2 t = __ldg(T + (jm & 255u));
    m31::Pt r = { t.x, t.y };
    for (int d = 1; 8 * d < half_log; d++) {
        t = __ldg(T + 256 * d + ((jm >> (8 * d)) & 255u));
        r = m31::padd(r, m31::Pt{ t.x, t.y });
    }
    if (j >> half_log) r.y = m31::sub(0, r.y);
    return r;
}
__device__ __forceinline__ uint64_t fold62(uint64_t s) { return (s >> 31) + (s & m31::P); }
__device__ __forceinline__ uint32_t reduce64(uint64_t s)
{
    s = fold62(fold62(s));                 // < 2^31 + 4
    uint32_t r = (uint32_t)s;
    r = (r >> 31) + (r & m31::P);          // < P + 5
    return r >= m31::P ? r - m31::P : r;
}
constexpr int QUOT_THREADS = 256, QUOT_SMEM_ENTRIES = 1024;
template <bool TS_DENOM>
__global__ void __launch_bounds__(QUOT_THREADS) quotients_kernel(int log_size, uint32_t half_init, ColSet cols, int n_batches,
                                                               const uint4 *__restrict__ cvals, const uint32_t *__restrict__ ab,
                                                               const uint32_t *__restrict__ bcoef, const uint32_t *__restrict__ pts,
                                                               const int *__restrict__ offsets, const int *__restrict__ col_index,
                                                               const uint2 *__restrict__ ptab, Q4 out)
{
    __shared__ uint4 sc[QUOT_SMEM_ENTRIES];
    __shared__ const uint32_t *scol[QUOT_SMEM_ENTRIES];   // column base pointers of the staged entries
    const uint64_t row = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = !(row >> log_size);
    const uint64_t r = live ? row : 0;
    m31::Pt dp = table_point(ptab, log_size - 1, m31::bitrev((uint32_t)r, log_size));
    m31::QM acc = { 0, 0, 0, 0 };
    for (int b = 0; b < n_batches; b++) {
        const uint32_t *pt = pts + 8 * b;
        m31::CM den_inv;
        if (TS_DENOM) {
            const uint32_t prx = pt[0], pix = pt[1], pry = pt[4], piy = pt[5];
            den_inv = { m31::inv(m31::sub(m31::mul(m31::sub(prx, dp.x), piy), m31::mul(m31::sub(pry, dp.y), pix))), 0 };
        } else {
            const m31::CM prx = { m31::sub(pt[0], dp.x), pt[1] }, pix = { pt[2], pt[3] }, pry = { m31::sub(pt[4], dp.y), pt[5] }, piy = { pt[6], pt[7] };
            const m31::CM den = m31::csub2(m31::cmul(prx, piy), m31::cmul(pry, pix));
            const uint32_t norm_inv = m31::inv(m31::add(m31::mul(den.re, den.re), m31::mul(den.im, den.im)));
            den_inv = { m31::mul(den.re, norm_inv), m31::mul(m31::sub(0, den.im), norm_inv) };
        }
        uint64_t s0 = 0, s1 = 0, s2 = 0, s3 = 0;
        for (int j0 = offsets[b]; j0 < offsets[b + 1]; j0 += QUOT_SMEM_ENTRIES) {
            const int cnt = min(QUOT_SMEM_ENTRIES, offsets[b + 1] - j0);
            __syncthreads();
            for (int e = threadIdx.x; e < cnt; e += QUOT_THREADS) { sc[e] = cvals[j0 + e]; scol[e] = cols.col(col_index[j0 + e]); }
            __syncthreads();
            int e = 0;
            for (; e + 12 <= cnt; e += 12) {        // 12 column loads in flight, then 48 multiply-accumulates
                uint32_t f[12];
#pragma unroll
                for (int u = 0; u < 12; u++) f[u] = __ldcs(scol[e + u] + r);
#pragma unroll
                for (int u = 0; u < 12; u += 3) {
                    const uint4 c0 = sc[e + u], c1 = sc[e + u + 1], c2 = sc[e + u + 2];
                    s0 = fold62(s0 + (uint64_t)c0.x * f[u] + (uint64_t)c1.x * f[u + 1] + (uint64_t)c2.x * f[u + 2]);
                    s1 = fold62(s1 + (uint64_t)c0.y * f[u] + (uint64_t)c1.y * f[u + 1] + (uint64_t)c2.y * f[u + 2]);
                    s2 = fold62(s2 + (uint64_t)c0.z * f[u] + (uint64_t)c1.z * f[u + 1] + (uint64_t)c2.z * f[u + 2]);
                    s3 = fold62(s3 + (uint64_t)c0.w * f[u] + (uint64_t)c1.w * f[u + 1] + (uint64_t)c2.w * f[u + 2]);
                }
            }
            for (; e < cnt; e++) {
                const uint32_t f = __ldcs(scol[e] + r);
                const uint4 c = sc[e];
                s0 = fold62(s0 + (uint64_t)c.x * f); s1 = fold62(s1 + (uint64_t)c.y * f);
                s2 = fold62(s2 + (uint64_t)c.z * f); s3 = fold62(s3 + (uint64_t)c.w * f);
            }
        }
        const uint32_t *k = ab + 8 * b;
        const m31::QM A = { k[0], k[1], k[2], k[3] }, B = { k[4], k[5], k[6], k[7] };
        m31::QM num = { reduce64(s0), reduce64(s1), reduce64(s2), reduce64(s3) };
        num = m31::qsub(num, m31::qadd(m31::qmul_m(A, dp.y), B));
        m31::QM bc = { bcoef[4 * b], bcoef[4 * b + 1], bcoef[4 * b + 2], bcoef[4 * b + 3] };
        acc = m31::qadd(m31::qmul(acc, bc), m31::qmul_cm(num, den_inv));
    }
    if (live) { out.c[0][row] = acc.a; out.c[1][row] = acc.b; out.c[2][row] = acc.c; out.c[3][row] = acc.d; }
}

// Tensor-core form of the same sum.  sum_j c_j f_j(row) over the columns of a batch is a (rows x columns) x (columns x 4)
// matrix product over M31; it is made exact on the INT8 tensor cores by splitting both factors into bytes:
//   f = sum_i 2^(8i) f_i,   (2^(8i) c mod P) = sum_j' 2^(8j') w_(i,j')     =>     c f = sum_j' 2^(8j') sum_i f_i w_(i,j')   (mod P)
// A column value loaded from HBM is already the 4 consecutive K-elements (column, byte i) of an mma.m16n8k32.u8 operand, so the
// data goes from the load straight into the B fragment: no unpacking, 2 LDG.128 + 1 LDS.128 + 4 IMMA per 32 rows x 8 columns
// instead of ~20 integer instructions per value.  The A operand holds the 16 weight rows m = 4 j' + coordinate; the host lays
// the table out as one uint4 (a0..a3) per lane per 8-column step.  Row r of the thread's 4 mma's: n = groupID <-> row
// base + 4 groupID + m, so that each thread's loads are 16 contiguous bytes of a column.  Byte products are < 2^16 and the s32
// accumulators are flushed (transposed through shared memory to one row per lane, recombined mod P) at least every
// QM_FLUSH columns: 4 QM_FLUSH 255^2 < 2^31.
constexpr int QM_WARPS = 8, QM_CHUNK = 256, QM_PAD = 17, QM_FLUSH = 4096, QM_R = 2, QM_D = 2;
__device__ __forceinline__ void imma_u8(int (&c)[4], const uint4 &a, uint32_t b0, uint32_t b1)
{
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3]) : "r"(a.x), "r"(a.y), "r"(a.z), "r"(a.w), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint4 ldcs4(const uint32_t *p) { return p ? __ldcs(reinterpret_cast<const uint4 *>(p)) : make_uint4(0, 0, 0, 0); }
template <bool TS_DENOM, int R, int D>
__global__ void __launch_bounds__(QM_WARPS * 32, 3) quotients_mma_kernel(int log_size, uint32_t half_init, int n_batches,
                                                                     const uint4 *__restrict__ wtab, const uint32_t *const *__restrict__ colptr,
                                                                     const uint32_t *__restrict__ ab, const uint32_t *__restrict__ bcoef,
                                                                     const uint32_t *__restrict__ pts, const int *__restrict__ offsets,
                                                                     const uint2 *__restrict__ ptab, Q4 out)
{
    __shared__ uint4 sw[QM_CHUNK / 8 * 32];
    __shared__ const uint32_t *sp[QM_CHUNK];
    __shared__ uint32_t st[QM_WARPS][32 * QM_PAD];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, tig = lane & 3;
    const uint64_t row0 = ((uint64_t)blockIdx.x * QM_WARPS + warp) * (32 * R);   // grid covers the domain exactly (host checks)
    m31::Pt dp[R];
    m31::QM acc[R];
#pragma unroll
    for (int q = 0; q < R; q++) {
        dp[q] = table_point(ptab, log_size - 1, m31::bitrev((uint32_t)(row0 + 32 * q + lane), log_size));
        acc[q] = { 0, 0, 0, 0 };
    }
    uint32_t *tr = st[warp];
    for (int b = 0; b < n_batches; b++) {
        int c[R][4][4];
        uint32_t num[R][4];
#pragma unroll
        for (int q = 0; q < R; q++)
#pragma unroll
            for (int m = 0; m < 4; m++) { num[q][m] = 0; c[q][m][0] = c[q][m][1] = c[q][m][2] = c[q][m][3] = 0; }
        auto flush = [&]() {
#pragma unroll
            for (int q = 0; q < R; q++) {
#pragma unroll
                for (int m = 0; m < 4; m++) {
                    tr[(8 * tig + m) * QM_PAD + g] = c[q][m][0];     tr[(8 * tig + 4 + m) * QM_PAD + g] = c[q][m][1];
                    tr[(8 * tig + m) * QM_PAD + g + 8] = c[q][m][2]; tr[(8 * tig + 4 + m) * QM_PAD + g + 8] = c[q][m][3];
                    c[q][m][0] = c[q][m][1] = c[q][m][2] = c[q][m][3] = 0;
                }
                __syncwarp();
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const uint32_t *t = tr + lane * QM_PAD + k;
                    const uint64_t v = (uint64_t)t[0] + ((uint64_t)t[4] << 8) + ((uint64_t)t[8] << 16) + ((uint64_t)t[12] << 24);
                    num[q][k] = m31::add(num[q][k], reduce64(v));
                }
                __syncwarp();
            }
        };
        int since = 0;
        for (int j0 = offsets[b]; j0 < offsets[b + 1]; j0 += QM_CHUNK) {
            const int cnt = min(QM_CHUNK, offsets[b + 1] - j0);     // a multiple of 8
            __syncthreads();
            for (int e = threadIdx.x; e < cnt * 4; e += QM_WARPS * 32) sw[e] = wtab[(size_t)j0 * 4 + e];
            for (int e = threadIdx.x; e < cnt; e += QM_WARPS * 32) sp[e] = colptr[j0 + e];
            __syncthreads();
            if (since + cnt > QM_FLUSH) { flush(); since = 0; }
            since += cnt;
            // D steps of column loads in flight per warp ahead of the mma's that consume them
            const int steps = cnt / 8;
            uint4 x[D][R], y[D][R];
            auto fetch = [&](int s, uint4 (&xs)[R], uint4 (&ys)[R]) {
                const uint32_t *p0 = sp[8 * s + tig], *p1 = sp[8 * s + 4 + tig];
#pragma unroll
                for (int q = 0; q < R; q++) {
                    xs[q] = ldcs4(p0 ? p0 + row0 + 32 * q + 4 * g : nullptr);
                    ys[q] = ldcs4(p1 ? p1 + row0 + 32 * q + 4 * g : nullptr);
                }
            };
#pragma unroll
            for (int d = 0; d < D - 1; d++)
                if (d < steps) fetch(d, x[d], y[d]);
            for (int s0 = 0; s0 < steps; s0 += D) {
#pragma unroll
                for (int d = 0; d < D; d++) {
                    const int s = s0 + d;
                    if (s < steps) {
                        if (s + D - 1 < steps) fetch(s + D - 1, x[(d + D - 1) % D], y[(d + D - 1) % D]);
                        const uint4 w = sw[s * 32 + lane];
#pragma unroll
                        for (int q = 0; q < R; q++) {
                            imma_u8(c[q][0], w, x[d][q].x, y[d][q].x); imma_u8(c[q][1], w, x[d][q].y, y[d][q].y);
                            imma_u8(c[q][2], w, x[d][q].z, y[d][q].z); imma_u8(c[q][3], w, x[d][q].w, y[d][q].w);
                        }
                    }
                }
            }
        }
        flush();
        const uint32_t *pt = pts + 8 * b, *k = ab + 8 * b;
        const m31::QM A = { k[0], k[1], k[2], k[3] }, B = { k[4], k[5], k[6], k[7] };
        const m31::QM bc = { bcoef[4 * b], bcoef[4 * b + 1], bcoef[4 * b + 2], bcoef[4 * b + 3] };
        // the M31 value to invert per row (the denominator itself, or the norm of the CM31 denominator); the R rows of a thread
        // share one inversion (Montgomery), with the reference's inv(0) = 0 kept by a rare second path
        uint32_t dv[R], di[R];
        m31::CM den[R];
#pragma unroll
        for (int q = 0; q < R; q++) {
            if (TS_DENOM) {
                dv[q] = m31::sub(m31::mul(m31::sub(pt[0], dp[q].x), pt[5]), m31::mul(m31::sub(pt[4], dp[q].y), pt[1]));
            } else {
                const m31::CM prx = { m31::sub(pt[0], dp[q].x), pt[1] }, pix = { pt[2], pt[3] }, pry = { m31::sub(pt[4], dp[q].y), pt[5] }, piy = { pt[6], pt[7] };
                den[q] = m31::csub2(m31::cmul(prx, piy), m31::cmul(pry, pix));
                dv[q] = m31::add(m31::mul(den[q].re, den[q].re), m31::mul(den[q].im, den[q].im));
            }
        }
        if (R == 2) {
            const uint32_t prod = m31::mul(dv[0], dv[R - 1]);
            if (prod != 0) {
                const uint32_t t = m31::inv(prod);
                di[0] = m31::mul(t, dv[R - 1]); di[R - 1] = m31::mul(t, dv[0]);
            } else {
                di[0] = m31::inv(dv[0]); di[R - 1] = m31::inv(dv[R - 1]);
            }
        } else {
#pragma unroll
            for (int q = 0; q < R; q++) di[q] = m31::inv(dv[q]);
        }
#pragma unroll
        for (int q = 0; q < R; q++) {
            m31::CM den_inv;
            if (TS_DENOM) den_inv = { di[q], 0 };
            else den_inv = { m31::mul(den[q].re, di[q]), m31::mul(m31::sub(0, den[q].im), di[q]) };
            m31::QM nm = { num[q][0], num[q][1], num[q][2], num[q][3] };
            nm = m31::qsub(nm, m31::qadd(m31::qmul_m(A, dp[q].y), B));
            const m31::QM term = m31::qmul_cm(nm, den_inv);
            acc[q] = b == 0 ? term : m31::qadd(m31::qmul(acc[q], bc), term);     // 0 * bc + term for the first batch (uniform branch)
        }
    }
#pragma unroll
    for (int q = 0; q < R; q++) {
        const uint64_t row = row0 + 32 * q + lane;
        out.c[0][row] = acc[q].a; out.c[1][row] = acc[q].b; out.c[2][row] = acc[q].c; out.c[3][row] = acc[q].d;
    }
}

// Wide-Fibonacci constraint quotients on the evaluation domain (examples/fibonacci.ts:72-91 c - (a^2 + b^2); the domain
// evaluation is Stwo's FrameworkComponent::evaluate_constraint_quotients_on_domain, a placeholder in
// constraint_framework/index.ts:131-152):   out[row] += denom_inv[row >> trace_log] * sum_i coeff_i (f_{i+2} - f_{i+1}^2 - f_i^2)(row).
// Reference form, one thread per row (small or unaligned inputs):
__global__ void __launch_bounds__(256) widefib_kernel(uint64_t n, int trace_log, ColSet cols, int n_cols, const uint4 *__restrict__ coeffs,
                                                      const uint32_t *__restrict__ denom_inv, Q4 out)
{
    const uint64_t row = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= n) return;
    uint32_t a = cols.col(0)[row], b = cols.col(1)[row];
    uint32_t sa = m31::mul(a, a), sb = m31::mul(b, b);
    m31::QM acc = { 0, 0, 0, 0 };
    for (int k = 2; k < n_cols; k++) {
        const uint32_t c = cols.col(k)[row];
        const uint32_t g = m31::sub(c, m31::add(sa, sb));
        const uint4 w = coeffs[k - 2];
        acc = m31::qadd(acc, m31::qmul_m(m31::QM{ w.x, w.y, w.z, w.w }, g));
        sa = sb; sb = m31::mul(c, c);
    }
    acc = m31::qmul_m(acc, denom_inv[row >> trace_log]);
    out.c[0][row] = m31::add(out.c[0][row], acc.a); out.c[1][row] = m31::add(out.c[1][row], acc.b);
    out.c[2][row] = m31::add(out.c[2][row], acc.c); out.c[3][row] = m31::add(out.c[3][row], acc.d);
}
// Tensor-core form.  The sum is linear in the column values and in their squares:
//   sum_i coeff_i (f_{i+2} - f_{i+1}^2 - f_i^2) = sum_k U_k f_k + sum_k V_k f_k^2,   U_k = coeff_{k-2},  V_k = -(coeff_{k-1} + coeff_k)
// so each thread squares the values it loaded (no neighbouring columns needed) and the two sums are the byte-split INT8 products
// of quotients_mma_kernel with two weight tables.  8 mma's per 32 rows x 8 columns, flush every WF_FLUSH columns.
constexpr int WF_CHUNK = 128, WF_FLUSH = 2048, WF_R = 2;
template <int R>
__global__ void __launch_bounds__(QM_WARPS * 32, 3) widefib_mma_kernel(int trace_log, int n_entries, const uint4 *__restrict__ wtab_f,
                                                                      const uint4 *__restrict__ wtab_s, const uint32_t *const *__restrict__ colptr,
                                                                      const uint32_t *__restrict__ denom_inv, Q4 out)
{
    __shared__ uint4 swf[WF_CHUNK / 8 * 32], sws[WF_CHUNK / 8 * 32];
    __shared__ const uint32_t *sp[WF_CHUNK];
    __shared__ uint32_t st[QM_WARPS][32 * QM_PAD];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, tig = lane & 3;
    const uint64_t row0 = ((uint64_t)blockIdx.x * QM_WARPS + warp) * (32 * R);
    uint32_t *tr = st[warp];
    int c[R][4][4];
    uint32_t num[R][4];
#pragma unroll
    for (int q = 0; q < R; q++)
#pragma unroll
        for (int m = 0; m < 4; m++) { num[q][m] = 0; c[q][m][0] = c[q][m][1] = c[q][m][2] = c[q][m][3] = 0; }
    auto flush = [&]() {
#pragma unroll
        for (int q = 0; q < R; q++) {
#pragma unroll
            for (int m = 0; m < 4; m++) {
                tr[(8 * tig + m) * QM_PAD + g] = c[q][m][0];     tr[(8 * tig + 4 + m) * QM_PAD + g] = c[q][m][1];
                tr[(8 * tig + m) * QM_PAD + g + 8] = c[q][m][2]; tr[(8 * tig + 4 + m) * QM_PAD + g + 8] = c[q][m][3];
                c[q][m][0] = c[q][m][1] = c[q][m][2] = c[q][m][3] = 0;
            }
            __syncwarp();
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const uint32_t *t = tr + lane * QM_PAD + k;
                const uint64_t v = (uint64_t)t[0] + ((uint64_t)t[4] << 8) + ((uint64_t)t[8] << 16) + ((uint64_t)t[12] << 24);
                num[q][k] = m31::add(num[q][k], reduce64(v));
            }
            __syncwarp();
        }
    };
    // v^2 mod P up to one multiple of P: any 32-bit representative works as an operand of the integer byte products
    auto sq = [](uint32_t v) { const uint64_t p = (uint64_t)v * v; return ((uint32_t)p & m31::P) + (uint32_t)(p >> 31); };
    auto sq4 = [&](uint4 v) { return make_uint4(sq(v.x), sq(v.y), sq(v.z), sq(v.w)); };
    int since = 0;
    for (int j0 = 0; j0 < n_entries; j0 += WF_CHUNK) {
        const int cnt = min(WF_CHUNK, n_entries - j0);     // a multiple of 8
        __syncthreads();
        for (int e = threadIdx.x; e < cnt * 4; e += QM_WARPS * 32) { swf[e] = wtab_f[(size_t)j0 * 4 + e]; sws[e] = wtab_s[(size_t)j0 * 4 + e]; }
        for (int e = threadIdx.x; e < cnt; e += QM_WARPS * 32) sp[e] = colptr[j0 + e];
        __syncthreads();
        if (since + cnt > WF_FLUSH) { flush(); since = 0; }
        since += cnt;
        const int steps = cnt / 8;
        constexpr int D = 2;      // load steps in flight ahead of the squares and mma's that consume them
        uint4 x[D][R], y[D][R];
        auto fetch = [&](int s, uint4 (&xs)[R], uint4 (&ys)[R]) {
            const uint32_t *p0 = sp[8 * s + tig], *p1 = sp[8 * s + 4 + tig];
#pragma unroll
            for (int q = 0; q < R; q++) {
                xs[q] = ldcs4(p0 ? p0 + row0 + 32 * q + 4 * g : nullptr);
                ys[q] = ldcs4(p1 ? p1 + row0 + 32 * q + 4 * g : nullptr);
            }
        };
        fetch(0, x[0], y[0]);
        for (int s0 = 0; s0 < steps; s0 += D) {
#pragma unroll
            for (int d = 0; d < D; d++) {
                const int s = s0 + d;
                if (s < steps) {
                    if (s + 1 < steps) fetch(s + 1, x[(d + 1) % D], y[(d + 1) % D]);
                    const uint4 wf = swf[s * 32 + lane], ws = sws[s * 32 + lane];
#pragma unroll
                    for (int q = 0; q < R; q++) {
                        imma_u8(c[q][0], wf, x[d][q].x, y[d][q].x); imma_u8(c[q][1], wf, x[d][q].y, y[d][q].y);
                        imma_u8(c[q][2], wf, x[d][q].z, y[d][q].z); imma_u8(c[q][3], wf, x[d][q].w, y[d][q].w);
                        const uint4 xs = sq4(x[d][q]), ys = sq4(y[d][q]);
                        imma_u8(c[q][0], ws, xs.x, ys.x); imma_u8(c[q][1], ws, xs.y, ys.y);
                        imma_u8(c[q][2], ws, xs.z, ys.z); imma_u8(c[q][3], ws, xs.w, ys.w);
                    }
                }
            }
        }
    }
    flush();
#pragma unroll
    for (int q = 0; q < R; q++) {
        const uint64_t row = row0 + 32 * q + lane;
        const uint32_t di = denom_inv[row >> trace_log];
        out.c[0][row] = m31::add(out.c[0][row], m31::mul(num[q][0], di)); out.c[1][row] = m31::add(out.c[1][row], m31::mul(num[q][1], di));
        out.c[2][row] = m31::add(out.c[2][row], m31::mul(num[q][2], di)); out.c[3][row] = m31::add(out.c[3][row], m31::mul(num[q][3], di));
    }
}

// eval_at_point (backend/cpu/circle.ts:52-69, poly/utils.ts:36-59 fold).  The reference folds bottom-up: level l combines
// (w[2k], w[2k+1]) -> w[2k] + f_l w[2k+1] with f_0 = y, f_1 = x, f_2 = pi(x), ...  Field arithmetic is exact, so the same value is
//     sum_k c_k B_k,   B_k = prod_{bit b of k} f_b                                    (16 coefficients per thread, B_k on the host)
// i.e. per thread a 16 x 4 dot product of M31 coefficients with constant QM31 coordinates: 4 multiplications per coefficient
// instead of 9 for the pairwise tree, and they accumulate in u64 (IMAD.WIDE with a 64-bit addend) -- four products of canonical
// values plus a folded carry stay below 2^64, so a reduction is needed only every four terms.  The levels above multiply QM31
// values by the constant f_l, which is the 4 x 4 M31 matrix of qmat(): again four products per output coordinate, one reduction.
struct FoldFactors { m31::QM f[32]; };
struct BaseConsts { uint32_t b[16][4]; };           // B_k, k = 0..15 (B_0 = 1 is not multiplied)
struct QMat { uint32_t m[4][4]; };                  // y = f * x  <=>  y_r = sum_c m[r][c] x_c
struct QMats { QMat l[4]; };
__device__ __forceinline__ uint64_t fold31(uint64_t a) { return (a & m31::P) + (a >> 31); }   // same residue, < 2^31 + 2^33
// blockIdx.y = polynomial of the batch; out / in of polynomial p at + p * stride words
__global__ void __launch_bounds__(128) fold_base16_kernel(ColSet polys, uint64_t n_out, BaseConsts bc, uint32_t *out, uint64_t out_stride)
{
    uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_out) return;
    const uint32_t *c = polys.col(blockIdx.y) + t * 16;
    uint32_t v[16];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const uint4 q = __ldcs(reinterpret_cast<const uint4 *>(c) + k);
        v[4 * k] = q.x; v[4 * k + 1] = q.y; v[4 * k + 2] = q.z; v[4 * k + 3] = q.w;
    }
    uint32_t r[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
        uint64_t acc = j == 0 ? v[0] : 0;
#pragma unroll
        for (int k = 1; k < 16; k++) {
            acc += (uint64_t)v[k] * bc.b[k][j];
            if ((k & 3) == 0) acc = fold31(acc);      // after 4 products (k = 1..4, 5..8, 9..12); the last 3 go straight to reduce
        }
        r[j] = m31::reduce(acc);
    }
    out += blockIdx.y * out_stride;
    reinterpret_cast<uint4 *>(out)[t] = make_uint4(r[0], r[1], r[2], r[3]);
}
// Eleven levels in one launch, every coefficient read once and coalesced: a warp takes 2048 consecutive coefficients (lane l loads
// the 16-byte units l, l + 32, ..., l + 480), multiplies each by its constant B_p = prod_{bit b of p} f_b (p = position inside the
// block; 2048 x 4 words in shared memory, laid out so that the lanes of a warp read consecutive 16-byte entries), accumulates in
// u64 as above and the 32 lane sums are added mod P by shuffles: one QM31 per 2048 coefficients leaves the kernel, so the levels
// above it cost nothing.  (Giving each THREAD 64 consecutive coefficients instead was tried: 256-byte lane stride, 1.78 ms
// against 1.21 ms for 100 x 2^23.)
constexpr int WARP_BLOCK_LOG = 11, WARP_BLOCK = 1 << WARP_BLOCK_LOG;
// the constants of a point: entry (i, e, l) = B_p, p = 128 i + 4 l + e, as a product of at most 11 of the fold factors
__global__ void warp_table_kernel(FoldFactors ff, uint4 *table)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= WARP_BLOCK) return;
    const int l = k & 31, e = (k >> 5) & 3, i = k >> 7, p = 128 * i + 4 * l + e;
    m31::QM b = { 1, 0, 0, 0 };
    for (int bit = 0; bit < WARP_BLOCK_LOG; bit++)
        if (p >> bit & 1) b = m31::qmul(b, ff.f[bit]);
    table[k] = make_uint4(b.a, b.b, b.c, b.d);
}
__global__ void __launch_bounds__(256) fold_warp_block_kernel(ColSet polys, uint32_t blocks_per_poly, const uint4 *__restrict__ table, uint32_t *out,
                                                              uint64_t out_stride)
{
    extern __shared__ uint4 sh_tab[];              // entry (i, e, l) at (4 i + e) * 32 + l: constant of position 128 i + 4 l + e
    for (int k = threadIdx.x; k < WARP_BLOCK; k += blockDim.x) sh_tab[k] = table[k];
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
    const uint32_t *poly = polys.col(blockIdx.y);
    uint32_t *dst = out + blockIdx.y * out_stride;
    for (uint32_t blk = blockIdx.x * n_warps + warp; blk < blocks_per_poly; blk += gridDim.x * n_warps) {
        const uint4 *c = reinterpret_cast<const uint4 *>(poly + (uint64_t)blk * WARP_BLOCK) + lane;
        uint64_t acc[4] = { 0, 0, 0, 0 };
#pragma unroll 4
        for (int i = 0; i < 16; i++) {
            const uint4 q = __ldcs(c + 32 * i);
            const uint32_t v[4] = { q.x, q.y, q.z, q.w };
#pragma unroll
            for (int e = 0; e < 4; e++) {
                const uint4 w = sh_tab[(4 * i + e) * 32 + lane];
                acc[0] += (uint64_t)v[e] * w.x; acc[1] += (uint64_t)v[e] * w.y; acc[2] += (uint64_t)v[e] * w.z; acc[3] += (uint64_t)v[e] * w.w;
            }
#pragma unroll
            for (int j = 0; j < 4; j++) acc[j] = fold31(acc[j]);      // a folded value + four products < 2^64
        }
        uint32_t r[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            r[j] = m31::reduce(acc[j]);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) r[j] = m31::add(r[j], __shfl_xor_sync(0xffffffffu, r[j], o));
        }
        if (lane == 0) reinterpret_cast<uint4 *>(dst)[blk] = make_uint4(r[0], r[1], r[2], r[3]);
    }
}
// the pairwise form, for polynomials with fewer than 16 coefficients
__global__ void fold_base_kernel(ColSet polys, uint64_t n_out, int levels, FoldFactors ff, uint32_t *out, uint64_t out_stride)
{
    uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_out) return;
    const int cnt = 1 << levels;   // < 16
    m31::QM w[8];
    const uint32_t *c = polys.col(blockIdx.y) + t * cnt;
    uint32_t v[16];
#pragma unroll
    for (int k = 0; k < 16; k++) v[k] = k < cnt ? c[k] : 0u;
#pragma unroll
    for (int k = 0; k < 8; k++)
        if (2 * k + 1 < cnt) w[k] = m31::qadd(m31::qfrom(v[2 * k]), m31::qmul_m(ff.f[0], v[2 * k + 1]));
#pragma unroll
    for (int l = 1; l < 4; l++)
        if (l < levels) {
#pragma unroll
            for (int k = 0; k < (8 >> l); k++)
                if (k < (cnt >> (l + 1))) w[k] = m31::qadd(w[2 * k], m31::qmul(ff.f[l], w[2 * k + 1]));
        }
    out += blockIdx.y * out_stride;
    reinterpret_cast<uint4 *>(out)[t] = make_uint4(w[0].a, w[0].b, w[0].c, w[0].d);
}
// w_even + f * w_odd with f as its matrix: 4 products + one canonical addend < 2^64 per coordinate
__device__ __forceinline__ m31::QM fold_pair(const m31::QM &e, const m31::QM &o, const QMat &m)
{
    const uint32_t x[4] = { o.a, o.b, o.c, o.d }, add[4] = { e.a, e.b, e.c, e.d };
    uint32_t r[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
        uint64_t acc = add[j];
#pragma unroll
        for (int c = 0; c < 4; c++) acc += (uint64_t)x[c] * m.m[j][c];
        r[j] = m31::reduce(acc);
    }
    return { r[0], r[1], r[2], r[3] };
}
__global__ void __launch_bounds__(128) fold_qm_kernel(const uint32_t *in, uint64_t in_stride, uint64_t n_out, int levels, QMats mats, uint32_t *out,
                                                      uint64_t out_stride)
{
    uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_out) return;
    const int cnt = 1 << levels;   // <= 16
    m31::QM w[16];
    const uint4 *c = reinterpret_cast<const uint4 *>(in + blockIdx.y * in_stride) + t * cnt;
#pragma unroll
    for (int k = 0; k < 16; k++)
        if (k < cnt) { uint4 a = c[k]; w[k] = { a.x, a.y, a.z, a.w }; }
#pragma unroll
    for (int l = 0; l < 4; l++)
        if (l < levels) {
#pragma unroll
            for (int k = 0; k < (8 >> l); k++)
                if (k < (cnt >> (l + 1))) w[k] = fold_pair(w[2 * k], w[2 * k + 1], mats.l[l]);
        }
    out += blockIdx.y * out_stride;
    reinterpret_cast<uint4 *>(out)[t] = make_uint4(w[0].a, w[0].b, w[0].c, w[0].d);
}
// multiplication by the constant f = (f0 + f1 i) + (f2 + f3 i) u, u^2 = 2 + i (fields/qm31.ts:9), as a matrix over M31
static QMat qmat(const host::QM &f)
{
    using namespace host;
    const uint32_t n1 = sub(0, f.b), n3 = sub(0, f.d);
    const uint32_t t = sub(add(f.c, f.c), f.d);                // 2 f2 - f3
    const uint32_t s = add(f.c, add(f.d, f.d));                // f2 + 2 f3
    QMat q = { { { f.a, n1, t, sub(0, s) }, { f.b, f.a, s, t }, { f.c, n3, f.a, n1 }, { f.d, f.c, f.b, f.a } } };
    return q;
}

static m31::QM to_dev(const uint32_t a[4]) { return { a[0], a[1], a[2], a[3] }; }

// scratch buffer (grown on demand, freed at shutdown by process exit)
// per device (Context::ew_scratch)
#define g_scratch (ctx().ew_scratch)
#define g_scratch_bytes (ctx().ew_scratch_bytes)
static int scratch(size_t bytes, uint32_t **out)
{
    if (bytes > g_scratch_bytes) {
        if (g_scratch) { TS_CUDA(cudaStreamSynchronize(ctx().stream)); TS_CUDA(cudaFree(g_scratch)); g_scratch = nullptr; g_scratch_bytes = 0; }
        cudaError_t e = cudaMalloc(&g_scratch, bytes);
        if (e == cudaErrorMemoryAllocation) { cudaGetLastError(); return fail(TSTWO_ERR_OOM, "scratch out of memory"); }
        TS_CUDA(e);
        g_scratch_bytes = bytes;
    }
    *out = g_scratch;
    return 0;
}

// ---- Fiat-Shamir channel on the device (channel/blake2.ts:25-224, vcs/blake2_merkle.ts:27-32), one thread: what
// FriProver.commitInnerLayers (fri.ts:676-716) does between two layers -- mix_root(root) then draw_felt() -- without a host round
// trip.  State layout (13 words): digest[8], n_challenges, n_sent, queue_len, then queue[8] (the base-field draws draw_felt keeps
// between calls, blake2.ts:172-181; NOT cleared by a digest update, as in the reference).
__device__ void b2s_hash64(const uint32_t m[16], uint32_t out[8])   // unkeyed BLAKE2s-256 of one full 64-byte block (RFC 7693)
{
    const uint32_t IV[8] = { 0x6A09E667u, 0xBB67AE85u, 0x3C6EF372u, 0xA54FF53Au, 0x510E527Fu, 0x9B05688Cu, 0x1F83D9ABu, 0x5BE0CD19u };
    const uint8_t SIG[10][16] = {
        { 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15 }, { 14, 10, 4, 8, 9, 15, 13, 6, 1, 12, 0, 2, 11, 7, 5, 3 },
        { 11, 8, 12, 0, 5, 2, 15, 13, 10, 14, 3, 6, 7, 1, 9, 4 }, { 7, 9, 3, 1, 13, 12, 11, 14, 2, 6, 5, 10, 4, 0, 15, 8 },
        { 9, 0, 5, 7, 2, 4, 10, 15, 14, 1, 11, 12, 6, 8, 3, 13 }, { 2, 12, 6, 10, 0, 11, 8, 3, 4, 13, 7, 5, 15, 14, 1, 9 },
        { 12, 5, 1, 15, 14, 13, 4, 10, 0, 7, 6, 3, 9, 2, 8, 11 }, { 13, 11, 7, 14, 12, 1, 3, 9, 5, 0, 15, 4, 8, 6, 2, 10 },
        { 6, 15, 14, 9, 11, 3, 0, 8, 12, 2, 13, 7, 1, 4, 10, 5 }, { 10, 2, 8, 4, 7, 6, 1, 5, 15, 11, 9, 14, 3, 12, 13, 0 } };
    uint32_t h[8], v[16];
    for (int i = 0; i < 8; i++) h[i] = IV[i];
    h[0] ^= 0x01010020u;                               // digest length 32, fanout 1, depth 1
    for (int i = 0; i < 8; i++) { v[i] = h[i]; v[8 + i] = IV[i]; }
    v[12] ^= 64u;                                      // t = 64 bytes
    v[14] = ~v[14];                                    // final block
    auto rotr = [](uint32_t x, int r) { return (x >> r) | (x << (32 - r)); };
    for (int r = 0; r < 10; r++) {
        auto G = [&](int a, int b, int c, int d, int x, int y) {
            v[a] = v[a] + v[b] + m[SIG[r][x]]; v[d] = rotr(v[d] ^ v[a], 16);
            v[c] = v[c] + v[d];                v[b] = rotr(v[b] ^ v[c], 12);
            v[a] = v[a] + v[b] + m[SIG[r][y]]; v[d] = rotr(v[d] ^ v[a], 8);
            v[c] = v[c] + v[d];                v[b] = rotr(v[b] ^ v[c], 7);
        };
        G(0, 4, 8, 12, 0, 1); G(1, 5, 9, 13, 2, 3); G(2, 6, 10, 14, 4, 5); G(3, 7, 11, 15, 6, 7);
        G(0, 5, 10, 15, 8, 9); G(1, 6, 11, 12, 10, 11); G(2, 7, 8, 13, 12, 13); G(3, 4, 9, 14, 14, 15);
    }
    for (int i = 0; i < 8; i++) out[i] = h[i] ^ v[i] ^ v[8 + i];
}
constexpr int CH_WORDS = 19;
__global__ void fri_channel_kernel(uint32_t *state, const uint32_t *root, uint32_t *alpha_out, uint32_t *root_out)
{
    if (threadIdx.x || blockIdx.x) return;
    uint32_t m[16], d[8];
    for (int i = 0; i < 8; i++) { m[i] = state[i]; m[8 + i] = root[i]; root_out[i] = root[i]; }
    b2s_hash64(m, d);                                  // Blake2sMerkleChannel.mix_root: digest = H(digest || root)
    for (int i = 0; i < 8; i++) state[i] = d[i];
    state[8] += 1;                                     // updateDigest: n_challenges++, n_sent = 0
    state[9] = 0;
    uint32_t qlen = state[10];
    while (qlen < 4) {                                 // draw_felt -> _draw_base_felts (retry when a word is >= 2P)
        for (;;) {
            for (int i = 0; i < 8; i++) m[i] = state[i];
            m[8] = state[9];
            for (int i = 9; i < 16; i++) m[i] = 0;
            state[9] += 1;
            b2s_hash64(m, d);
            bool ok = true;
            for (int i = 0; i < 8; i++) ok = ok && d[i] < 2u * m31::P;
            if (ok) break;
        }
        for (int i = 0; i < 8; i++) state[11 + qlen + i] = d[i] >= m31::P ? d[i] - m31::P : d[i];
        qlen += 8;
    }
    for (int i = 0; i < 4; i++) alpha_out[i] = state[11 + i];
    for (uint32_t i = 4; i < qlen; i++) state[11 + i - 4] = state[11 + i];
    state[10] = qlen - 4;
}

// ---- tstwo_debug_field_op: the device field arithmetic of m31.cuh, one operation per thread, so that the reference's
// golden vectors (test-vectors/*.json, replayed by test-equivalence/fields/*.test.ts) run on the DEVICE code one to one
__global__ void field_op_kernel(int field, int op, const uint32_t *a, const uint32_t *b, uint32_t *out, int n)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t *x = a + 4 * i, *y = b + 4 * i;
    uint32_t r[4] = { 0, 0, 0, 0 };
    if (field == 0) {
        switch (op) {
        case 0: r[0] = m31::add(x[0], y[0]); break;
        case 1: r[0] = m31::sub(x[0], y[0]); break;
        case 2: r[0] = m31::mul(x[0], y[0]); break;
        case 3: r[0] = m31::neg(x[0]); break;
        case 4: r[0] = m31::inv(x[0]); break;
        case 7: r[0] = m31::reduce((uint64_t)x[0] | ((uint64_t)x[1] << 32)); break;
        case 8: r[0] = m31::partial_reduce(x[0]); break;
        case 9: r[0] = m31::sqr(x[0]); break;
        }
    } else if (field == 1) {
        const m31::CM u = { x[0], x[1] }, v = { y[0], y[1] };
        m31::CM w = { 0, 0 };
        switch (op) {
        case 0: w = m31::cadd(u, v); break;
        case 1: w = m31::csub2(u, v); break;
        case 2: w = m31::cmul(u, v); break;
        case 3: w = m31::cneg(u); break;
        case 4: w = m31::cinv(u); break;
        case 6: w = m31::cmul_m(u, y[0]); break;
        }
        r[0] = w.re; r[1] = w.im;
    } else {
        const m31::QM u = { x[0], x[1], x[2], x[3] }, v = { y[0], y[1], y[2], y[3] };
        m31::QM w = { 0, 0, 0, 0 };
        switch (op) {
        case 0: w = m31::qadd(u, v); break;
        case 1: w = m31::qsub(u, v); break;
        case 2: w = m31::qmul(u, v); break;
        case 3: w = m31::qneg(u); break;
        case 4: w = m31::qinv(u); break;
        case 5: w = m31::qmul_cm(u, { y[0], y[1] }); break;
        case 6: w = m31::qmul_m(u, y[0]); break;
        }
        r[0] = w.a; r[1] = w.b; r[2] = w.c; r[3] = w.d;
    }
    for (int k = 0; k < 4; k++) out[4 * i + k] = r[k];
}

}  // namespace tstwo

using namespace tstwo;

extern "C" {

int tstwo_debug_field_op(int field, int op, const uint32_t *a, const uint32_t *b, uint32_t *out, int n)
{
    TS_REQUIRE_INIT();
    if (field < 0 || field > 2 || n < 0 || (n && (!a || !b || !out))) return fail(TSTWO_ERR_BAD_ARG, "bad field_op arguments");
    if (n == 0) return 0;
    uint32_t *d;
    const size_t words = 4 * (size_t)n;
    int rc = scratch(3 * words * 4, &d);
    if (rc) return rc;
    cudaStream_t st = ctx().stream;
    TS_CUDA(cudaMemcpyAsync(d, a, words * 4, cudaMemcpyHostToDevice, st));
    TS_CUDA(cudaMemcpyAsync(d + words, b, words * 4, cudaMemcpyHostToDevice, st));
    TS_LAUNCH(field_op_kernel, (unsigned)((n + 127) / 128), 128, 0, field, op, d, d + words, d + 2 * words, n);
    TS_CUDA(cudaMemcpyAsync(out, d + 2 * words, words * 4, cudaMemcpyDeviceToHost, st));
    TS_CUDA(cudaStreamSynchronize(st));
    return 0;
}

int tstwo_fold_line(const tstwo_dptr src[4], int log_size, uint32_t coset_init, const uint32_t alpha[4],
                    const tstwo_dptr dst[4], tstwo_twiddles tw)
{
    TS_REQUIRE_INIT();
    if (log_size < 1 || log_size > 30) return fail(TSTWO_ERR_SIZE_MISMATCH, "fold_line: Evaluation too small, must have at least 2 elements.");
    const uint64_t n_out = 1ull << (log_size - 1);
    CQ4 s; Q4 d;
    for (int k = 0; k < 4; k++) { s.c[k] = (const uint32_t *)src[k]; d.c[k] = (uint32_t *)dst[k]; }
    const uint32_t *itab = nullptr;
    TwiddleTree *t = tw ? get_tree(tw) : nullptr;
    if (tw && !t) return fail(TSTWO_ERR_BAD_ARG, "unknown twiddle tree handle");
    if (t && is_doubling_of(coset_init, log_size, t)) {
        // layer of the tower whose coset is (coset_init, log_size): bitrev(x of first half) = what fold_line needs
        itab = t->itw2 + ((1ull << t->root_log) - (1ull << log_size));
    } else {
        uint32_t *tmp;
        int rc = scratch(n_out * 4, &tmp);
        if (rc) return rc;
        if ((rc = launch_coset_points(tmp, 0, coset_init & 0x7fffffffu, log_size, log_size - 1))) return rc;
        itab = tmp;
    }
    LAUNCH_FOLD(fold_line_kernel, s, d, n_out, itab, 0, to_dev(alpha), (const uint32_t *)nullptr);
    return 0;
}

int tstwo_fold_circle_into_line(const tstwo_dptr dst[4], const tstwo_dptr src[4], int log_size, uint32_t half_init,
                                const uint32_t alpha[4], tstwo_twiddles tw)
{
    TS_REQUIRE_INIT();
    if (log_size < 1 || log_size > 30) return fail(TSTWO_ERR_SIZE_MISMATCH, "fold_circle_into_line: Length mismatch between src and dst after considering fold step.");
    const uint64_t n_out = 1ull << (log_size - 1);
    CQ4 s; Q4 d;
    for (int k = 0; k < 4; k++) { s.c[k] = (const uint32_t *)src[k]; d.c[k] = (uint32_t *)dst[k]; }
    const uint32_t *itab = nullptr;
    int mode = 0;
    TwiddleTree *t = tw ? get_tree(tw) : nullptr;
    if (tw && !t) return fail(TSTWO_ERR_BAD_ARG, "unknown twiddle tree handle");
    const int half_log = log_size - 1;
    if (t && half_log >= 2 && is_doubling_of(half_init, half_log, t)) {
        itab = t->itw2 + ((1ull << t->root_log) - (1ull << half_log));   // T1 of the domain
        mode = 1;
    } else {
        uint32_t *tmp;
        int rc = scratch(n_out * 4, &tmp);
        if (rc) return rc;
        // p = half_coset.at(bitrev(i, half_log)), twiddle 1/p.y
        if ((rc = launch_coset_points(tmp, 1, half_init & 0x7fffffffu, half_log, half_log))) return rc;
        itab = tmp;
    }
    host::QM a = host::qld(alpha), a2 = host::qmul(a, a);
    uint32_t a2v[4];
    host::qst(a2v, a2);
    LAUNCH_FOLD(fold_circle_kernel, s, d, n_out, itab, mode, to_dev(alpha), to_dev(a2v), (const uint32_t *)nullptr);
    return 0;
}

// FriProver.commitInnerLayers (fri.ts:676-716) in ONE call, the transcript on the device: per layer  Merkle tree of the 4 coordinate
// columns -> mix_root -> draw_felt -> fold_line (-> fold_circle_into_line of the circle column whose folded size matches), all
// enqueued back to back; the host waits once at the end.  Same transcript, same layers as the per-layer calls.
int tstwo_fri_commit_layers(const tstwo_dptr first[4], int first_log, uint32_t first_coset_init, int last_log,
                            const tstwo_dptr *circle_cols, const int *circle_logs, const uint32_t *circle_half_inits, int n_circle, tstwo_twiddles tw,
                            uint32_t channel_state[19], tstwo_dptr *layer_evals, tstwo_dptr *layer_trees, uint8_t *roots,
                            uint32_t *alphas, int *n_layers_out)
{
    TS_REQUIRE_INIT();
    if (!first || first_log < 1 || first_log > 30 || last_log < 0 || !channel_state || !layer_evals || !layer_trees || !roots || !alphas || !n_layers_out ||
        n_circle < 0 || (n_circle && (!circle_cols || !circle_logs || !circle_half_inits)))
        return fail(TSTWO_ERR_BAD_ARG, "bad fri_commit_layers arguments");
    const int n_layers = first_log > last_log ? first_log - last_log : 0;
    *n_layers_out = n_layers;
    if (n_layers == 0) return 0;
    Context &c = ctx();
    uint32_t *d_state = nullptr, *d_alpha = nullptr, *d_roots = nullptr;
    TS_CUDA(cudaMallocAsync(&d_state, CH_WORDS * 4, c.stream));
    TS_CUDA(cudaMallocAsync(&d_alpha, (size_t)n_layers * 16, c.stream));
    TS_CUDA(cudaMallocAsync(&d_roots, (size_t)n_layers * 32, c.stream));
    auto cleanup = [&]() { cudaFreeAsync(d_state, c.stream); cudaFreeAsync(d_alpha, c.stream); cudaFreeAsync(d_roots, c.stream); };
    auto body = [&]() -> int {
        TS_CUDA(cudaMemcpyAsync(d_state, channel_state, CH_WORDS * 4, cudaMemcpyHostToDevice, c.stream));
        tstwo_dptr cur[4] = { first[0], first[1], first[2], first[3] };
        uint32_t coset = first_coset_init & 0x7fffffffu;
        int pending = 0;
        const uint32_t zero_alpha[4] = { 0, 0, 0, 0 };
        for (int l = 0; l < n_layers; l++) {
            const int log = first_log - l;
            const uint64_t n = 1ull << log;
            // Merkle tree of this layer's evaluation
            tstwo_dptr tree;
            int rc = tstwo_malloc(32ull * (2 * n - 1), &tree);
            if (rc) return rc;
            layer_trees[l] = tree;
            int logs[4] = { log, log, log, log };
            if ((rc = tstwo_merkle_commit(cur, logs, 4, tree, nullptr))) return rc;
            // mix_root + draw_felt on the device
            TS_LAUNCH(fri_channel_kernel, 1, 32, 0, d_state, (const uint32_t *)tree, d_alpha + 4 * l, d_roots + 8 * l);
            // fold_line into fresh columns (alpha read from device memory)
            tstwo_dptr nxt[4];
            // one block, coordinate k at + k * stride words (stride = n/2, at least 4 words: every coordinate stays 16-byte aligned)
            const uint64_t stride_words = n / 2 < 4 ? 4 : n / 2;
            if ((rc = tstwo_malloc(4 * 4ull * stride_words, &nxt[0]))) return rc;
            for (int k = 0; k < 4; k++) {
                nxt[k] = nxt[0] + (uint64_t)k * 4ull * stride_words;
                layer_evals[4 * l + k] = nxt[k];
            }
            {
                CQ4 s4; Q4 d4;
                for (int k = 0; k < 4; k++) { s4.c[k] = (const uint32_t *)cur[k]; d4.c[k] = (uint32_t *)nxt[k]; }
                TwiddleTree *t = tw ? get_tree(tw) : nullptr;
                if (tw && !t) return fail(TSTWO_ERR_BAD_ARG, "unknown twiddle tree handle");
                const uint32_t *itab;
                const uint64_t n_out = n / 2;
                if (t && is_doubling_of(coset, log, t)) itab = t->itw2 + ((1ull << t->root_log) - (1ull << log));
                else {
                    // every layer needs its own table while the launches are in flight: carve them from one scratch block
                    uint32_t *tmp;
                    if ((rc = scratch((4ull << first_log) * 4, &tmp))) return rc;
                    tmp += (2ull << first_log) - (2ull << log);
                    if ((rc = launch_coset_points(tmp, 0, coset, log, log - 1))) return rc;
                    itab = tmp;
                }
                LAUNCH_FOLD(fold_line_kernel, s4, d4, n_out, itab, 0, to_dev(zero_alpha), (const uint32_t *)(d_alpha + 4 * l));
            }
            coset = (coset * 2u) & 0x7fffffffu;
            // a circle column whose folded size equals the new layer joins it, with the same alpha (fri.ts:704-711)
            if (pending < n_circle && circle_logs[pending] == log) {
                const int clog = circle_logs[pending], half_log = clog - 1;
                const uint32_t half_init = circle_half_inits[pending] & 0x7fffffffu;
                const uint64_t n_out = 1ull << half_log;
                CQ4 s4; Q4 d4;
                for (int k = 0; k < 4; k++) { s4.c[k] = (const uint32_t *)circle_cols[4 * pending + k]; d4.c[k] = (uint32_t *)nxt[k]; }
                TwiddleTree *t = tw ? get_tree(tw) : nullptr;
                const uint32_t *itab;
                int mode = 0;
                if (t && half_log >= 2 && is_doubling_of(half_init, half_log, t)) { itab = t->itw2 + ((1ull << t->root_log) - (1ull << half_log)); mode = 1; }
                else {
                    uint32_t *tmp;
                    if ((rc = scratch((4ull << first_log) * 4, &tmp))) return rc;
                    tmp += (2ull << first_log) + (2ull << first_log) - (2ull << clog);     // second half of the scratch block
                    if ((rc = launch_coset_points(tmp, 1, half_init, half_log, half_log))) return rc;
                    itab = tmp;
                }
                LAUNCH_FOLD(fold_circle_kernel, s4, d4, n_out, itab, mode, to_dev(zero_alpha), to_dev(zero_alpha), (const uint32_t *)(d_alpha + 4 * l));
                pending++;
            }
            for (int k = 0; k < 4; k++) cur[k] = nxt[k];
        }
        TS_CUDA(cudaMemcpyAsync(channel_state, d_state, CH_WORDS * 4, cudaMemcpyDeviceToHost, c.stream));
        TS_CUDA(cudaMemcpyAsync(alphas, d_alpha, (size_t)n_layers * 16, cudaMemcpyDeviceToHost, c.stream));
        TS_CUDA(cudaMemcpyAsync(roots, d_roots, (size_t)n_layers * 32, cudaMemcpyDeviceToHost, c.stream));
        TS_CUDA(cudaStreamSynchronize(c.stream));
        return 0;
    };
    const int rc = body();
    cleanup();
    return rc;
}

int tstwo_fri_decompose(const tstwo_dptr src[4], int log_size, const tstwo_dptr g[4], uint32_t lambda[4])
{
    TS_REQUIRE_INIT();
    if (log_size < 0 || log_size > 30) return fail(TSTWO_ERR_BAD_ARG, "bad decompose size");
    const uint64_t n = 1ull << log_size;
    CQ4 s; Q4 d;
    for (int k = 0; k < 4; k++) { s.c[k] = (const uint32_t *)src[k]; d.c[k] = (uint32_t *)g[k]; }
    uint32_t *tmp;
    int rc = scratch(64, &tmp);
    if (rc) return rc;
    unsigned long long *sums = (unsigned long long *)tmp;
    TS_CUDA(cudaMemsetAsync(sums, 0, 64, ctx().stream));
    bool vec = n >= 8;
    for (int k = 0; k < 4; k++) vec = vec && !((src[k] | g[k]) & 15);
    unsigned grid = (unsigned)(((vec ? n / 4 : n) + 255) / 256);
    if (grid > 148 * 8) grid = 148 * 8;
    if (vec) TS_LAUNCH(half_sums_vec_kernel, grid, 256, 0, s, n / 4, sums);
    else TS_LAUNCH(half_sums_kernel, grid, 256, 0, s, n, sums);
    unsigned long long h[8];
    TS_CUDA(cudaMemcpyAsync(h, sums, 64, cudaMemcpyDeviceToHost, ctx().stream));
    TS_CUDA(cudaStreamSynchronize(ctx().stream));
    uint32_t lam[4];
    const uint32_t n_inv = host::inv((uint32_t)(n % host::P));
    for (int k = 0; k < 4; k++) {
        uint32_t a = (uint32_t)(h[2 * k] % host::P), b = (uint32_t)(h[2 * k + 1] % host::P);
        // n == 1: lambda = 0 - f[0] (all of it is in the "first half" sum by construction: half = 0 => hi = 1)
        if (n == 1) lam[k] = host::neg(b);
        else lam[k] = host::mul(host::sub(a, b), n_inv);
    }
    if (vec) TS_LAUNCH(decompose_apply_vec_kernel, (unsigned)((n / 4 + 255) / 256), 256, 0, s, d, n / 4, to_dev(lam));
    else TS_LAUNCH(decompose_apply_kernel, (unsigned)((n + 255) / 256), 256, 0, s, d, n, to_dev(lam));
    if (lambda) memcpy(lambda, lam, 16);
    return 0;
}

int tstwo_accumulate(const tstwo_dptr dst[4], const tstwo_dptr src[4], uint64_t len)
{
    TS_REQUIRE_INIT();
    if (len == 0) return 0;
    CQ4 s; Q4 d;
    for (int k = 0; k < 4; k++) { s.c[k] = (const uint32_t *)src[k]; d.c[k] = (uint32_t *)dst[k]; }
    bool vec = len >= 4 && (len & 3) == 0;
    for (int k = 0; k < 4; k++) vec = vec && !((reinterpret_cast<uintptr_t>(d.c[k]) | reinterpret_cast<uintptr_t>(s.c[k])) & 15);
    if (vec) TS_LAUNCH(accumulate_vec_kernel, (unsigned)((len / 4 + 255) / 256), 256, 0, d, s, len / 4);
    else TS_LAUNCH(accumulate_kernel, (unsigned)((len + 255) / 256), 256, 0, d, s, len);
    return 0;
}

// Weight words of entry e of an INT8 product table (quotients_mma_kernel, widefib_mma_kernel): rows m = 4 j' + coordinate,
// word = bytes j' of (2^(8i) c mod P), i = 0..3; lane (g, tig) of an 8-entry step holds a0 = W[tig][g], a1 = W[tig][g+8],
// a2 = W[4+tig][g], a3 = W[4+tig][g+8].
static void put_mma_weights(uint32_t *wt, int e, const uint32_t c4[4])
{
    const int step = e >> 3, le = e & 7;
    for (int k = 0; k < 4; k++) {
        const uint32_t cv = c4[k];
        uint32_t rot[4];
        for (int i = 0; i < 4; i++) rot[i] = i ? ((cv << (8 * i)) | (cv >> (31 - 8 * i))) & 0x7fffffffu : cv;   // 2^(8i) c mod P
        for (int jb = 0; jb < 4; jb++) {
            const int m = 4 * jb + k;
            uint32_t word = 0;
            for (int i = 0; i < 4; i++) word |= ((rot[i] >> (8 * jb)) & 0xffu) << (8 * i);
            const int lane = (m & 7) * 4 + (le & 3), reg = (le >> 2) * 2 + (m >> 3);
            wt[((size_t)step * 32 + lane) * 4 + reg] = word;
        }
    }
}
static int g_quot_mode = 0, g_last_quot = 0;   // 0 auto, 1 integer-pipe kernel, 2 tensor-core kernel
int tstwo_debug_set_quot_mode(int mode) { g_quot_mode = mode; return 0; }
int tstwo_debug_last_quot_mode(void) { return g_last_quot; }

int tstwo_accumulate_quotients(int log_size, uint32_t half_init, const tstwo_dptr *cols, int n_cols, const uint32_t rc4[4],
                               int n_batches, const uint32_t *points, const int *batch_offsets, const int *col_index,
                               const uint32_t *values, const tstwo_dptr out[4])
{
    return tstwo_accumulate_quotients_ex(log_size, half_init, cols, n_cols, rc4, n_batches, points, batch_offsets, col_index, values, out,
                                         TSTWO_FLAG_QUIRKS);
}

int tstwo_accumulate_quotients_ex(int log_size, uint32_t half_init, const tstwo_dptr *cols, int n_cols, const uint32_t rc4[4],
                                  int n_batches, const uint32_t *points, const int *batch_offsets, const int *col_index,
                                  const uint32_t *values, const tstwo_dptr out[4], uint32_t flags)
{
    TS_REQUIRE_INIT();
    if (log_size < 1 || log_size > 30 || n_batches < 0) return fail(TSTWO_ERR_BAD_ARG, "bad quotient arguments");
    const int total = n_batches ? batch_offsets[n_batches] : 0;
    for (int j = 0; j < total; j++)
        if (col_index[j] < 0 || col_index[j] >= n_cols) return fail(TSTWO_ERR_BAD_ARG, "column index %d out of range", col_index[j]);
    const bool ts_quirks = (flags & TSTWO_FLAG_QUIRKS) != 0;
    const uint64_t n = 1ull << log_size;
    // tensor-core form: whole CTAs of QM_WARPS x 32 QM_R rows, 16-byte aligned columns
    bool mma = g_quot_mode != 1 && n >= (uint64_t)QM_WARPS * 32 * QM_R && total > 0;
    for (int j = 0; j < total && mma; j++) mma = ((uintptr_t)cols[col_index[j]] & 15) == 0;
    if (g_quot_mode == 2 && !mma) return fail(TSTWO_ERR_BAD_ARG, "tensor-core quotient kernel does not apply");
    g_last_quot = mma ? 2 : 1;
    // padded batch layout of the tensor-core form: every batch starts at a multiple of 8 entries
    std::vector<int> poff(n_batches + 1, 0);
    for (int b = 0; b < n_batches; b++) poff[b + 1] = poff[b] + (batch_offsets[b + 1] - batch_offsets[b] + 7) / 8 * 8;
    const size_t ptotal = mma ? (size_t)poff[n_batches] : 0;

    // one host image of everything the kernel reads, one copy.  Sections (32-bit words, each a multiple of 4 words):
    auto round4 = [](size_t w) { return (w + 3) & ~(size_t)3; };
    const size_t o_ptab = 0, o_wt = o_ptab + PT_WORDS, o_ptr = o_wt + ptotal * 16, o_consts = o_ptr + round4(ptotal * 2 + 2),
                 o_ab = o_consts + (size_t)total * 4 + 4, o_bc = o_ab + (size_t)n_batches * 8 + 8, o_pts = o_bc + (size_t)n_batches * 4 + 4,
                 o_off = o_pts + (size_t)n_batches * 8 + 8, o_ci = o_off + round4((size_t)n_batches + 1), o_end = o_ci + round4((size_t)total + 1);
    std::vector<uint32_t> h(o_end, 0);
    {   // digit tables of the domain points (table_point)
        const int half_log = log_size - 1;
        const uint32_t step = half_log == 0 ? 0u : 1u << (31 - half_log);
        for (int d = 0; d < PT_DIGITS; d++) {
            const host::Pt gd = host::index_to_point((uint32_t)(((uint64_t)step << (8 * d)) & 0x7fffffffu));
            host::Pt cur = d ? host::Pt{ 1u, 0u } : host::index_to_point(half_init & 0x7fffffffu);
            for (int v = 0; v < 256; v++) { h[o_ptab + (256 * d + v) * 2] = cur.x; h[o_ptab + (256 * d + v) * 2 + 1] = cur.y; cur = host::padd(cur, gd); }
        }
    }
    // quotientConstants (quotients.ts:124-152, constraints.ts:117-129) on the host
    uint32_t *consts = &h[o_consts];
    host::QM rc = host::qld(rc4);
    for (int b = 0; b < n_batches; b++) {
        host::QM py = host::qld(points + 8 * b + 4);
        host::QM alpha = { 1, 0, 0, 0 }, sum_a = { 0, 0, 0, 0 }, sum_b = { 0, 0, 0, 0 };
        // TS: QM31.complexConjugate conjugates both CM31 halves (qm31.ts:433-435); Stwo: u -> -u, i.e. (c0, -c1)
        auto conj = [&](host::QM q) { return ts_quirks ? host::qconj(q) : host::QM{ q.a, q.b, host::neg(q.c), host::neg(q.d) }; };
        const host::QM c = host::qsub(conj(py), py);
        for (int j = batch_offsets[b]; j < batch_offsets[b + 1]; j++) {
            alpha = host::qmul(alpha, rc);
            host::QM v = host::qld(values + 4 * j);
            host::QM a = host::qsub(conj(v), v);
            host::QM bb = host::qsub(host::qmul(v, c), host::qmul(a, py));
            sum_a = host::qadd(sum_a, host::qmul(alpha, a));
            sum_b = host::qadd(sum_b, host::qmul(alpha, bb));
            host::qst(&consts[4 * j], host::qmul(alpha, c));
        }
        host::qst(&h[o_ab + 8 * b], sum_a);
        host::qst(&h[o_ab + 8 * b + 4], sum_b);
        host::qst(&h[o_bc + 4 * b], alpha);   // random_coeff ^ len
        for (int k = 0; k < 8; k++) h[o_pts + 8 * b + k] = points[8 * b + k];
    }
    if (mma) {
        uint64_t *ptrs = reinterpret_cast<uint64_t *>(&h[o_ptr]);
        for (int b = 0; b < n_batches; b++)
            for (int j = batch_offsets[b]; j < batch_offsets[b + 1]; j++) {
                const int e = poff[b] + (j - batch_offsets[b]);
                ptrs[e] = (uint64_t)cols[col_index[j]];
                put_mma_weights(&h[o_wt], e, &consts[4 * (size_t)j]);
            }
        for (int b = 0; b <= n_batches; b++) h[o_off + b] = (uint32_t)poff[b];
    } else {
        for (int b = 0; b <= n_batches; b++) h[o_off + b] = n_batches ? (uint32_t)batch_offsets[b] : 0u;
        for (int j = 0; j < total; j++) h[o_ci + j] = (uint32_t)col_index[j];
    }
    uint32_t *d;
    int r = scratch(o_end * 4, &d);
    if (r) return r;
    cudaStream_t st = ctx().stream;
    TS_CUDA(cudaMemcpyAsync(d, h.data(), o_end * 4, cudaMemcpyHostToDevice, st));
    Q4 o;
    for (int k = 0; k < 4; k++) o.c[k] = (uint32_t *)out[k];
    const uint32_t hi = half_init & 0x7fffffffu;
    const uint2 *d_ptab = (const uint2 *)(d + o_ptab);
    if (mma) {
        const unsigned grid = (unsigned)(n / (QM_WARPS * 32 * QM_R));
        const uint4 *d_wt = (const uint4 *)(d + o_wt);
        const uint32_t *const *d_ptr = (const uint32_t *const *)(d + o_ptr);
        if (ts_quirks)
            TS_LAUNCH((quotients_mma_kernel<true, QM_R, QM_D>), grid, QM_WARPS * 32, 0, log_size, hi, n_batches, d_wt, d_ptr, d + o_ab, d + o_bc,
                      d + o_pts, (const int *)(d + o_off), d_ptab, o);
        else
            TS_LAUNCH((quotients_mma_kernel<false, QM_R, QM_D>), grid, QM_WARPS * 32, 0, log_size, hi, n_batches, d_wt, d_ptr, d + o_ab, d + o_bc,
                      d + o_pts, (const int *)(d + o_off), d_ptab, o);
    } else {
        ColSet cs;
        if ((r = make_colset(cols, n_cols, 0, &cs))) return r;
        const unsigned grid = (unsigned)((n + QUOT_THREADS - 1) / QUOT_THREADS);
        if (ts_quirks)
            TS_LAUNCH(quotients_kernel<true>, grid, QUOT_THREADS, 0, log_size, hi, cs, n_batches, (const uint4 *)(d + o_consts), d + o_ab, d + o_bc,
                      d + o_pts, (const int *)(d + o_off), (const int *)(d + o_ci), d_ptab, o);
        else
            TS_LAUNCH(quotients_kernel<false>, grid, QUOT_THREADS, 0, log_size, hi, cs, n_batches, (const uint4 *)(d + o_consts), d + o_ab, d + o_bc,
                      d + o_pts, (const int *)(d + o_off), (const int *)(d + o_ci), d_ptab, o);
    }
    TS_CUDA(cudaStreamSynchronize(st));   // the host image goes out of scope
    return 0;
}

int tstwo_widefib_constraints(int trace_log_size, int eval_log_size, const tstwo_dptr *cols, int n_cols, const uint32_t *coeffs,
                              const uint32_t *denom_inv, const tstwo_dptr out[4])
{
    TS_REQUIRE_INIT();
    if (trace_log_size < 0 || eval_log_size < trace_log_size || eval_log_size > 30 || eval_log_size - trace_log_size > 16 || n_cols < 2)
        return fail(TSTWO_ERR_BAD_ARG, "bad wide-Fibonacci arguments");
    if (n_cols == 2) return 0;   // no constraints
    const uint64_t n = 1ull << eval_log_size;
    const int n_den = 1 << (eval_log_size - trace_log_size), n_con = n_cols - 2;
    bool mma = g_quot_mode != 1 && n >= (uint64_t)QM_WARPS * 32 * WF_R;
    for (int k = 0; k < n_cols && mma; k++) mma = ((uintptr_t)cols[k] & 15) == 0;
    if (g_quot_mode == 2 && !mma) return fail(TSTWO_ERR_BAD_ARG, "tensor-core wide-Fibonacci kernel does not apply");
    g_last_quot = mma ? 2 : 1;
    const size_t pcols = ((size_t)n_cols + 7) / 8 * 8;
    auto round4 = [](size_t w) { return (w + 3) & ~(size_t)3; };
    const size_t o_wf = 0, o_ws = o_wf + (mma ? pcols * 16 : 0), o_ptr = o_ws + (mma ? pcols * 16 : 0), o_co = o_ptr + (mma ? pcols * 2 : 0),
                 o_den = o_co + (size_t)n_con * 4, o_end = o_den + round4(n_den);
    std::vector<uint32_t> h(o_end, 0);
    memcpy(&h[o_co], coeffs, (size_t)n_con * 16);
    memcpy(&h[o_den], denom_inv, (size_t)n_den * 4);
    for (size_t i = 0; i < (size_t)n_con * 4 + n_den; i++)
        if (h[o_co + i] >= host::P) return fail(TSTWO_ERR_BAD_ARG, "field element out of range");
    if (mma) {
        uint64_t *ptrs = reinterpret_cast<uint64_t *>(&h[o_ptr]);
        auto coeff = [&](int i) { return i >= 0 && i < n_con ? host::qld(coeffs + 4 * i) : host::QM{ 0, 0, 0, 0 }; };
        for (int k = 0; k < n_cols; k++) {
            ptrs[k] = (uint64_t)cols[k];
            uint32_t u[4], v[4];
            host::qst(u, coeff(k - 2));
            host::qst(v, host::qsub(host::QM{ 0, 0, 0, 0 }, host::qadd(coeff(k - 1), coeff(k))));
            put_mma_weights(&h[o_wf], k, u);
            put_mma_weights(&h[o_ws], k, v);
        }
    }
    uint32_t *d;
    int r = scratch(o_end * 4, &d);
    if (r) return r;
    cudaStream_t st = ctx().stream;
    TS_CUDA(cudaMemcpyAsync(d, h.data(), o_end * 4, cudaMemcpyHostToDevice, st));
    Q4 o;
    for (int k = 0; k < 4; k++) o.c[k] = (uint32_t *)out[k];
    if (mma) {
        TS_LAUNCH((widefib_mma_kernel<WF_R>), (unsigned)(n / (QM_WARPS * 32 * WF_R)), QM_WARPS * 32, 0, trace_log_size, (int)pcols,
                  (const uint4 *)(d + o_wf), (const uint4 *)(d + o_ws), (const uint32_t *const *)(d + o_ptr), d + o_den, o);
    } else {
        ColSet cs;
        if ((r = make_colset(cols, n_cols, 0, &cs))) return r;
        TS_LAUNCH(widefib_kernel, (unsigned)((n + 255) / 256), 256, 0, n, trace_log_size, cs, n_cols, (const uint4 *)(d + o_co), d + o_den, o);
    }
    TS_CUDA(cudaStreamSynchronize(st));   // the host image goes out of scope
    return 0;
}

int tstwo_eval_at_point(tstwo_dptr coeffs, int log_size, const uint32_t px[4], const uint32_t py[4], uint32_t out[4])
{
    return tstwo_eval_at_point_many(&coeffs, 1, log_size, px, py, out);
}

int tstwo_eval_at_point_many(const tstwo_dptr *coeffs, int n_polys, int log_size, const uint32_t px[4], const uint32_t py[4], uint32_t *out)
{
    TS_REQUIRE_INIT();
    Span span("Evaluate columns out of domain");
    if (log_size < 0 || log_size > 30 || n_polys < 0 || (n_polys && (!out || !coeffs))) return fail(TSTWO_ERR_BAD_ARG, "bad eval_at_point arguments");
    if (n_polys == 0) return 0;
    if (log_size == 0) {
        for (int p = 0; p < n_polys; p++) {
            uint32_t c0;
            int rc = tstwo_download(&c0, coeffs[p], 4);
            if (rc) return rc;
            out[4 * p] = c0; out[4 * p + 1] = out[4 * p + 2] = out[4 * p + 3] = 0;
        }
        return 0;
    }
    FoldFactors ff;
    host::QM hf[32];
    host::QM x = host::qld(px), one = { 1, 0, 0, 0 };
    hf[0] = host::qld(py);
    for (int i = 1; i < log_size; i++) {
        hf[i] = x;
        host::QM sx = host::qmul(x, x);
        x = host::qsub(host::qadd(sx, sx), one);     // circle.ts:37-40 double_x
    }
    for (int i = 0; i < log_size; i++) ff.f[i] = { hf[i].a, hf[i].b, hf[i].c, hf[i].d };
    BaseConsts bc;
    memset(&bc, 0, sizeof bc);
    if (log_size >= 4)
        for (int k = 0; k < 16; k++) {               // B_k = product of the factors of k's set bits
            host::QM b = one;
            for (int bit = 0; bit < 4; bit++) if (k >> bit & 1) b = host::qmul(b, hf[bit]);
            bc.b[k][0] = b.a; bc.b[k][1] = b.b; bc.b[k][2] = b.c; bc.b[k][3] = b.d;
        }
    const uint64_t n = 1ull << log_size;
    // levels of the first launch: 11 (fold_warp_block_kernel) when there is enough work to fill the GPU with 2048-coefficient blocks
    const bool big = log_size >= 14 && ((uint64_t)n_polys << log_size) >= (1ull << 22);
    const int lv = big ? WARP_BLOCK_LOG : log_size < 4 ? log_size : 4;
    uint32_t *d_table = nullptr;
    const uint64_t cnt0 = n >> lv, per_poly = 4 * (cnt0 + (cnt0 >> 1) + 64);      // words of scratch per polynomial (two ping-pong buffers)
    // polynomials per pass: at most 256 MiB of scratch, at most 65535 (gridDim.y), at most the pointer-table slot
    uint64_t fit = (256ull << 20) / (per_poly * 4);
    if (fit < 1) fit = 1;
    if (fit > 4096) fit = 4096;
    const int batch = (int)(fit < (uint64_t)n_polys ? fit : (uint64_t)n_polys);
    uint32_t *buf;
    int rc = scratch((uint64_t)batch * per_poly * 4 + (lv == WARP_BLOCK_LOG ? 16 * (size_t)WARP_BLOCK : 0), &buf);
    if (rc) return rc;
    if (lv == WARP_BLOCK_LOG) {                      // the constants table sits behind the partial sums
        d_table = buf + (uint64_t)batch * per_poly;
        TS_LAUNCH(warp_table_kernel, WARP_BLOCK / 256, 256, 0, ff, (uint4 *)d_table);
    }
    for (int p0 = 0; p0 < n_polys; p0 += batch) {
        const int nb = batch < n_polys - p0 ? batch : n_polys - p0;
        ColSet cs;
        if ((rc = make_colset(coeffs + p0, nb, 0, &cs))) return rc;
        uint64_t cnt = cnt0;
        uint32_t *cur = buf, *nxt = buf + 4 * cnt0;
        if (lv == WARP_BLOCK_LOG) {
            unsigned gx = (unsigned)((cnt + 7) / 8);                                // 8 warps per CTA, one 2048-coefficient block per warp and turn
            const unsigned want = (unsigned)((4 * ctx().sm_count + nb - 1) / nb);  // ~4 CTAs per SM over the batch: the 32 KiB table is loaded once per CTA
            if (gx > want) gx = want < 1 ? 1 : want;
            TS_LAUNCH(fold_warp_block_kernel, dim3(gx, nb), 256, 4 * 4 * WARP_BLOCK, cs, (uint32_t)cnt, (const uint4 *)d_table, cur, per_poly);
        } else if (lv == 4) TS_LAUNCH(fold_base16_kernel, dim3((unsigned)((cnt + 127) / 128), nb), 128, 0, cs, cnt, bc, cur, per_poly);
        else TS_LAUNCH(fold_base_kernel, dim3((unsigned)((cnt + 127) / 128), nb), 128, 0, cs, cnt, lv, ff, cur, per_poly);
        int level = lv;
        while (cnt > 1) {
            const int rem = log_size - level, l = rem < 4 ? rem : 4;
            const uint64_t ncnt = cnt >> l;
            QMats mats;
            memset(&mats, 0, sizeof mats);
            for (int k = 0; k < l; k++) mats.l[k] = qmat(hf[level + k]);
            TS_LAUNCH(fold_qm_kernel, dim3((unsigned)((ncnt + 127) / 128), nb), 128, 0, cur, per_poly, ncnt, l, mats, nxt, per_poly);
            uint32_t *t = cur; cur = nxt; nxt = t;
            cnt = ncnt; level += l;
        }
        // the results sit at cur + p * per_poly: one strided copy
        TS_CUDA(cudaMemcpy2DAsync(out + 4 * (size_t)p0, 16, cur, per_poly * 4, 16, nb, cudaMemcpyDeviceToHost, ctx().stream));
        TS_CUDA(cudaStreamSynchronize(ctx().stream));
    }
    return 0;
}

}  // extern "C"

// ---- eval_at_point WITHOUT any allocation (tstwo_eval_at_point_noalloc) -------------------------------------------------------
// No counterpart in the reference: added for callers that cannot call cudaMalloc (code
// under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same kernels and launch sequence as
// tstwo_eval_at_point_many for one polynomial, but the two ping-pong partial buffers live in a file-scope device array instead of
// the library's grow-on-demand scratch (scratch() above calls cudaMalloc).  2^20 coefficients need
// 4 * (2^16 + 2^15 + 64) = 393 472 words = 1.5 MiB: that is the array, held by every GPU that loads the library until exit.
//   * In HOST code the array's name is a host-side stand-in, not a device address: the address is resolved with
//     cudaGetSymbolAddress, once per device slot, with that device current, and only that pointer is given to kernels.
//   * Every call shares the array and it keeps the last call's contents.  Nothing in it has to start at zero: each launch writes
//     the partials the next launch reads (fold_base* fills buffer 0 completely up to cnt, every fold_qm reads only what the launch
//     before it wrote), so there is nothing to clear.  Calls that overlap on the GPU (two streams, two host threads) RACE on it.
constexpr int NOALLOC_MAX_LOG = 20;
constexpr uint64_t NOALLOC_WORDS = 4ull * ((1ull << (NOALLOC_MAX_LOG - 4)) + (1ull << (NOALLOC_MAX_LOG - 5)) + 64);
static __device__ __align__(256) uint32_t g_noalloc_scratch[NOALLOC_WORDS];
static uint32_t *g_noalloc_ptr[MAX_SLOTS];       // device address of the array on the device of each slot (nullptr = not resolved yet)

static int noalloc_scratch(uint32_t **out)
{
    Context &c = ctx();
    if (c.slot < 0 || c.slot >= MAX_SLOTS) return fail(TSTWO_ERR_BAD_ARG, "tstwo_eval_at_point_noalloc: bad device slot %d", c.slot);
    if (!g_noalloc_ptr[c.slot]) {                // once per device, with that device current (ctx_select made it so)
        int dev = -1;
        TS_CUDA(cudaGetDevice(&dev));
        if (dev != c.device) TS_CUDA(cudaSetDevice(c.device));
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            return fail(TSTWO_ERR_CUDA, "tstwo_eval_at_point_noalloc: cudaGetSymbolAddress failed on device %d: %s", c.device, cudaGetErrorString(e));
        }
        g_noalloc_ptr[c.slot] = static_cast<uint32_t *>(p);
    }
    *out = g_noalloc_ptr[c.slot];
    return 0;
}

extern "C" {

int tstwo_eval_at_point_noalloc_max_log(void) { return NOALLOC_MAX_LOG; }

int tstwo_eval_at_point_noalloc(tstwo_dptr coeffs, int log_size, const uint32_t px[4], const uint32_t py[4], uint32_t out[4])
{
    TS_REQUIRE_INIT();
    if (log_size < 0 || !coeffs || !px || !py || !out) return fail(TSTWO_ERR_BAD_ARG, "bad eval_at_point_noalloc arguments");
    if (log_size > NOALLOC_MAX_LOG)
        return fail(TSTWO_ERR_BAD_ARG, "tstwo_eval_at_point_noalloc: log_size %d exceeds the fixed scratch (at most 2^%d coefficients; "
                                       "use tstwo_eval_at_point for larger polynomials)", log_size, NOALLOC_MAX_LOG);
    Context &c = ctx();
    if (log_size == 0) {                          // a constant: its value, no kernel
        uint32_t c0 = 0;
        TS_CUDA(cudaMemcpyAsync(&c0, (const void *)coeffs, 4, cudaMemcpyDeviceToHost, c.stream));
        TS_CUDA(cudaStreamSynchronize(c.stream));
        out[0] = c0; out[1] = out[2] = out[3] = 0;
        return 0;
    }
    uint32_t *buf;
    int rc = noalloc_scratch(&buf);
    if (rc) return rc;
    FoldFactors ff;
    host::QM hf[32];
    host::QM x = host::qld(px), one = { 1, 0, 0, 0 };
    hf[0] = host::qld(py);
    for (int i = 1; i < log_size; i++) {
        hf[i] = x;
        host::QM sx = host::qmul(x, x);
        x = host::qsub(host::qadd(sx, sx), one);     // circle.ts:37-40 double_x
    }
    for (int i = 0; i < log_size; i++) ff.f[i] = { hf[i].a, hf[i].b, hf[i].c, hf[i].d };
    BaseConsts bc;
    memset(&bc, 0, sizeof bc);
    if (log_size >= 4)
        for (int k = 0; k < 16; k++) {
            host::QM b = one;
            for (int bit = 0; bit < 4; bit++) if (k >> bit & 1) b = host::qmul(b, hf[bit]);
            bc.b[k][0] = b.a; bc.b[k][1] = b.b; bc.b[k][2] = b.c; bc.b[k][3] = b.d;
        }
    const uint64_t n = 1ull << log_size;
    const int lv = log_size < 4 ? log_size : 4;
    const uint64_t cnt0 = n >> lv, per_poly = 4 * (cnt0 + (cnt0 >> 1) + 64);
    if (per_poly > NOALLOC_WORDS) return fail(TSTWO_ERR_BAD_ARG, "tstwo_eval_at_point_noalloc: internal size check failed");
    ColSet cs = { reinterpret_cast<const uint32_t *>(coeffs), 0, nullptr };      // one column, no pointer table
    cs.aligned16 = (coeffs & 15) == 0;
    if (lv == 4 && !cs.aligned16) return fail(TSTWO_ERR_BAD_ARG, "tstwo_eval_at_point_noalloc: coefficients must be 16-byte aligned");
    uint64_t cnt = cnt0;
    uint32_t *cur = buf, *nxt = buf + 4 * cnt0;
    if (lv == 4) TS_LAUNCH(fold_base16_kernel, dim3((unsigned)((cnt + 127) / 128), 1), 128, 0, cs, cnt, bc, cur, per_poly);
    else TS_LAUNCH(fold_base_kernel, dim3((unsigned)((cnt + 127) / 128), 1), 128, 0, cs, cnt, lv, ff, cur, per_poly);
    int level = lv;
    while (cnt > 1) {
        const int rem = log_size - level, l = rem < 4 ? rem : 4;
        const uint64_t ncnt = cnt >> l;
        QMats mats;
        memset(&mats, 0, sizeof mats);
        for (int k = 0; k < l; k++) mats.l[k] = qmat(hf[level + k]);
        TS_LAUNCH(fold_qm_kernel, dim3((unsigned)((ncnt + 127) / 128), 1), 128, 0, cur, per_poly, ncnt, l, mats, nxt, per_poly);
        uint32_t *t = cur; cur = nxt; nxt = t;
        cnt = ncnt; level += l;
    }
    TS_CUDA(cudaMemcpyAsync(out, cur, 16, cudaMemcpyDeviceToHost, c.stream));
    TS_CUDA(cudaStreamSynchronize(c.stream));
    return 0;
}

}  // extern "C"
