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
__global__ void fold_line_kernel(CQ4 src, Q4 dst, uint64_t n_out, const uint32_t *itab, int mode, m31::QM alpha)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_out) return;
    uint32_t it2 = pair_itw2(itab, mode, i);
    uint32_t f0[4], f1[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        uint2 p = __ldcs(reinterpret_cast<const uint2 *>(src.c[k]) + i);
        f0[k] = p.x; f1[k] = p.y;
        m31::ibutterfly(f0[k], f1[k], it2);
    }
    m31::QM r = m31::qadd({ f0[0], f0[1], f0[2], f0[3] }, m31::qmul(alpha, { f1[0], f1[1], f1[2], f1[3] }));
    dst.c[0][i] = r.a; dst.c[1][i] = r.b; dst.c[2][i] = r.c; dst.c[3][i] = r.d;
}

// fri.ts:162-192 / backend/cpu/fri.ts:61-92: dst[i] = dst[i] alpha^2 + (alpha f1 + f0), twiddle 1/p.y
__global__ void fold_circle_kernel(CQ4 src, Q4 dst, uint64_t n_out, const uint32_t *itab, int mode, m31::QM alpha,
                                   m31::QM alpha_sq)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_out) return;
    uint32_t it2 = pair_itw2(itab, mode, i);
    uint32_t f0[4], f1[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        uint2 p = __ldcs(reinterpret_cast<const uint2 *>(src.c[k]) + i);
        f0[k] = p.x; f1[k] = p.y;
        m31::ibutterfly(f0[k], f1[k], it2);
    }
    m31::QM fp = m31::qadd(m31::qmul(alpha, { f1[0], f1[1], f1[2], f1[3] }), { f0[0], f0[1], f0[2], f0[3] });
    m31::QM cur = { dst.c[0][i], dst.c[1][i], dst.c[2][i], dst.c[3][i] };
    m31::QM r = m31::qadd(m31::qmul(cur, alpha_sq), fp);
    dst.c[0][i] = r.a; dst.c[1][i] = r.b; dst.c[2][i] = r.c; dst.c[3][i] = r.d;
}

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
                                                               const int *__restrict__ offsets, const int *__restrict__ col_index, Q4 out)
{
    __shared__ uint4 sc[QUOT_SMEM_ENTRIES];
    __shared__ const uint32_t *scol[QUOT_SMEM_ENTRIES];   // column base pointers of the staged entries
    const uint64_t row = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = !(row >> log_size);
    const uint64_t r = live ? row : 0;
    m31::Pt dp = index_to_point(domain_index_at(half_init, log_size - 1, m31::bitrev((uint32_t)r, log_size)));
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

// eval_at_point (backend/cpu/circle.ts:52-69, poly/utils.ts:36-59 fold): bottom-up pairwise folding, 4 levels
// (16 values) per thread.  Level l combines (w[2k], w[2k+1]) -> w[2k] + f_l w[2k+1]; f_0 = y, f_1 = x, f_2 = pi(x)...
struct FoldFactors { m31::QM f[32]; };
__global__ void fold_base_kernel(const uint32_t *coeffs, uint64_t n_out, int levels, FoldFactors ff, uint32_t *out)
{
    uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_out) return;
    const int cnt = 1 << levels;   // <= 16
    m31::QM w[8];
    const uint32_t *c = coeffs + t * cnt;
#pragma unroll
    for (int k = 0; k < 8; k++)
        if (2 * k + 1 < cnt) w[k] = m31::qadd(m31::qfrom(c[2 * k]), m31::qmul_m(ff.f[0], c[2 * k + 1]));
#pragma unroll
    for (int l = 1; l < 4; l++)
        if (l < levels) {
#pragma unroll
            for (int k = 0; k < (8 >> l); k++)
                if (k < (cnt >> (l + 1))) w[k] = m31::qadd(w[2 * k], m31::qmul(ff.f[l], w[2 * k + 1]));
        }
    out[4 * t] = w[0].a; out[4 * t + 1] = w[0].b; out[4 * t + 2] = w[0].c; out[4 * t + 3] = w[0].d;
}
__global__ void fold_qm_kernel(const uint32_t *in, uint64_t n_out, int levels, int level0, FoldFactors ff, uint32_t *out)
{
    uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_out) return;
    const int cnt = 1 << levels;   // <= 16
    m31::QM w[16];
    const uint4 *c = reinterpret_cast<const uint4 *>(in) + t * cnt;
#pragma unroll
    for (int k = 0; k < 16; k++)
        if (k < cnt) { uint4 a = c[k]; w[k] = { a.x, a.y, a.z, a.w }; }
#pragma unroll
    for (int l = 0; l < 4; l++)
        if (l < levels) {
#pragma unroll
            for (int k = 0; k < (8 >> l); k++)
                if (k < (cnt >> (l + 1))) w[k] = m31::qadd(w[2 * k], m31::qmul(ff.f[level0 + l], w[2 * k + 1]));
        }
    out[4 * t] = w[0].a; out[4 * t + 1] = w[0].b; out[4 * t + 2] = w[0].c; out[4 * t + 3] = w[0].d;
}

static m31::QM to_dev(const uint32_t a[4]) { return { a[0], a[1], a[2], a[3] }; }

// scratch buffer (grown on demand, freed at shutdown by process exit)
static uint32_t *g_scratch = nullptr;
static size_t g_scratch_bytes = 0;
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

}  // namespace tstwo

using namespace tstwo;

extern "C" {

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
    TS_LAUNCH(fold_line_kernel, (unsigned)((n_out + 255) / 256), 256, 0, s, d, n_out, itab, 0, to_dev(alpha));
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
    TS_LAUNCH(fold_circle_kernel, (unsigned)((n_out + 255) / 256), 256, 0, s, d, n_out, itab, mode, to_dev(alpha), to_dev(a2v));
    return 0;
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
    unsigned grid = (unsigned)((n + 255) / 256);
    if (grid > 148 * 8) grid = 148 * 8;
    TS_LAUNCH(half_sums_kernel, grid, 256, 0, s, n, sums);
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
    TS_LAUNCH(decompose_apply_kernel, (unsigned)((n + 255) / 256), 256, 0, s, d, n, to_dev(lam));
    if (lambda) memcpy(lambda, lam, 16);
    return 0;
}

int tstwo_accumulate(const tstwo_dptr dst[4], const tstwo_dptr src[4], uint64_t len)
{
    TS_REQUIRE_INIT();
    if (len == 0) return 0;
    CQ4 s; Q4 d;
    for (int k = 0; k < 4; k++) { s.c[k] = (const uint32_t *)src[k]; d.c[k] = (uint32_t *)dst[k]; }
    TS_LAUNCH(accumulate_kernel, (unsigned)((len + 255) / 256), 256, 0, d, s, len);
    return 0;
}

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
    // quotientConstants (quotients.ts:124-152, constraints.ts:117-129) on the host
    const bool ts_quirks = (flags & TSTWO_FLAG_QUIRKS) != 0;
    std::vector<uint32_t> consts((size_t)total * 4 + 4), ab((size_t)n_batches * 8 + 8), bcoef((size_t)n_batches * 4 + 4), pts((size_t)n_batches * 8 + 8);
    host::QM rc = host::qld(rc4);
    for (int b = 0; b < n_batches; b++) {
        host::QM py = host::qld(points + 8 * b + 4);
        host::QM alpha = { 1, 0, 0, 0 }, sum_a = { 0, 0, 0, 0 }, sum_b = { 0, 0, 0, 0 };
        for (int j = batch_offsets[b]; j < batch_offsets[b + 1]; j++) {
            alpha = host::qmul(alpha, rc);
            host::QM v = host::qld(values + 4 * j);
            // TS: QM31.complexConjugate conjugates both CM31 halves (qm31.ts:433-435); Stwo: u -> -u, i.e. (c0, -c1)
            auto conj = [&](host::QM q) { return ts_quirks ? host::qconj(q) : host::QM{ q.a, q.b, host::neg(q.c), host::neg(q.d) }; };
            host::QM a = host::qsub(conj(v), v);
            host::QM c = host::qsub(conj(py), py);
            host::QM bb = host::qsub(host::qmul(v, c), host::qmul(a, py));
            sum_a = host::qadd(sum_a, host::qmul(alpha, a));
            sum_b = host::qadd(sum_b, host::qmul(alpha, bb));
            host::qst(&consts[4 * j], host::qmul(alpha, c));
        }
        host::qst(&ab[8 * b], sum_a);
        host::qst(&ab[8 * b + 4], sum_b);
        host::qst(&bcoef[4 * b], alpha);   // random_coeff ^ len
        for (int k = 0; k < 8; k++) pts[8 * b + k] = points[8 * b + k];
    }
    size_t w_consts = consts.size(), w_ab = ab.size(), w_bc = bcoef.size(), w_pts = pts.size(), w_off = (size_t)n_batches + 1, w_ci = (size_t)total + 1;
    uint32_t *tmp;
    int r = scratch((w_consts + w_ab + w_bc + w_pts + w_off + w_ci) * 4, &tmp);
    if (r) return r;
    uint32_t *d_consts = tmp, *d_ab = d_consts + w_consts, *d_bc = d_ab + w_ab, *d_pts = d_bc + w_bc;
    int *d_off = (int *)(d_pts + w_pts), *d_ci = d_off + w_off;
    cudaStream_t st = ctx().stream;
    TS_CUDA(cudaMemcpyAsync(d_consts, consts.data(), w_consts * 4, cudaMemcpyHostToDevice, st));
    TS_CUDA(cudaMemcpyAsync(d_ab, ab.data(), w_ab * 4, cudaMemcpyHostToDevice, st));
    TS_CUDA(cudaMemcpyAsync(d_bc, bcoef.data(), w_bc * 4, cudaMemcpyHostToDevice, st));
    TS_CUDA(cudaMemcpyAsync(d_pts, pts.data(), w_pts * 4, cudaMemcpyHostToDevice, st));
    std::vector<int> off0(1, 0);
    TS_CUDA(cudaMemcpyAsync(d_off, n_batches ? batch_offsets : off0.data(), w_off * 4, cudaMemcpyHostToDevice, st));
    if (total) TS_CUDA(cudaMemcpyAsync(d_ci, col_index, (size_t)total * 4, cudaMemcpyHostToDevice, st));
    ColSet cs;
    if ((r = make_colset(cols, n_cols, 0, &cs))) return r;
    Q4 o;
    for (int k = 0; k < 4; k++) o.c[k] = (uint32_t *)out[k];
    const uint64_t n = 1ull << log_size;
    const unsigned grid = (unsigned)((n + QUOT_THREADS - 1) / QUOT_THREADS);
    if (flags & TSTWO_FLAG_QUIRKS)
        TS_LAUNCH(quotients_kernel<true>, grid, QUOT_THREADS, 0, log_size, half_init & 0x7fffffffu, cs, n_batches, (const uint4 *)d_consts,
                  d_ab, d_bc, d_pts, d_off, d_ci, o);
    else
        TS_LAUNCH(quotients_kernel<false>, grid, QUOT_THREADS, 0, log_size, half_init & 0x7fffffffu, cs, n_batches, (const uint4 *)d_consts,
                  d_ab, d_bc, d_pts, d_off, d_ci, o);
    TS_CUDA(cudaStreamSynchronize(st));   // host vectors go out of scope
    return 0;
}

int tstwo_eval_at_point(tstwo_dptr coeffs, int log_size, const uint32_t px[4], const uint32_t py[4], uint32_t out[4])
{
    TS_REQUIRE_INIT();
    if (log_size < 0 || log_size > 30 || !out) return fail(TSTWO_ERR_BAD_ARG, "bad eval_at_point arguments");
    if (log_size == 0) {
        uint32_t c0;
        int rc = tstwo_download(&c0, coeffs, 4);
        if (rc) return rc;
        out[0] = c0; out[1] = out[2] = out[3] = 0;
        return 0;
    }
    FoldFactors ff;
    host::QM x = host::qld(px), one = { 1, 0, 0, 0 };
    { host::QM y = host::qld(py); ff.f[0] = { y.a, y.b, y.c, y.d }; }
    for (int i = 1; i < log_size; i++) {
        ff.f[i] = { x.a, x.b, x.c, x.d };
        host::QM sx = host::qmul(x, x);
        x = host::qsub(host::qadd(sx, sx), one);     // circle.ts:37-40 double_x
    }
    uint32_t *buf;
    const uint64_t n = 1ull << log_size;
    int lv = log_size < 4 ? log_size : 4;
    uint64_t cnt = n >> lv;
    int rc = scratch((cnt + (cnt >> 1) + 64) * 16, &buf);
    if (rc) return rc;
    uint32_t *cur = buf, *nxt = buf + 4 * cnt;
    TS_LAUNCH(fold_base_kernel, (unsigned)((cnt + 127) / 128), 128, 0, (const uint32_t *)coeffs, cnt, lv, ff, cur);
    int level = lv;
    while (cnt > 1) {
        int rem = log_size - level;
        int l = rem < 4 ? rem : 4;
        uint64_t ncnt = cnt >> l;
        TS_LAUNCH(fold_qm_kernel, (unsigned)((ncnt + 127) / 128), 128, 0, cur, ncnt, l, level, ff, nxt);
        uint32_t *t = cur; cur = nxt; nxt = t;
        cnt = ncnt; level += l;
    }
    return tstwo_download(out, (tstwo_dptr)cur, 16);
}

}  // extern "C"
