// A radix-16 circle-FFT stage (four butterfly layers, packages/core/src/fft.ts:12-30) on the 5th-generation tensor
// cores of sm_100a: tcgen05.mma kind::i8, operands u8, accumulators s32 in TMEM.
//
// The four layers of a stage whose twiddles are the same for every vector of a tile (the TOP stage of a pass: its
// twiddles depend only on index bits above the tile) are ONE 16 x 16 matrix W over M31 applied to many 16-element
// vectors.  Exactly, by byte splitting: with f = sum_i 2^(8i) f_i and (2^(8i) W[r][c]) mod P = sum_j 2^(8j) w[r][c][i][j],
//     out[r] = sum_j 2^(8j) * ( sum_(c,i) f_(c,i) * w[r][c][i][j] )      (mod P)
// i.e. a [vectors x 64] x [64 x 64] u8 product; every accumulator is < 64 * 255^2 < 2^22 and the four of an output are
// recombined mod P on the integer pipes (6 instructions, canonical result, against 14 for two butterflies).
// A u32 word of data IS four consecutive K-bytes (c, i = 0..3) of its vector, so with the A operand in TMEM (lane =
// vector, column = element c) the data need no byte transposition: registers -> tcgen05.st -> tcgen05.mma -> tcgen05.ld.
// Inputs may be ANY u32 representative of the residue (bytes are bytes); outputs are canonical.
//
// Measured in isolation (tools/tc_stage.cu, profiles/r02_tc_stage.txt): M = 128, N = 64, K = 32 issues every 45 clk
// (5.8 k MAC/clk/SM), TMEM reads 890 B/clk/SM, bit-exact against the scalar butterflies.
#pragma once
#include <cstdint>

#include "m31.cuh"

namespace tstwo {
namespace tc {

__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
// ---- TMEM management (one warp allocates all 512 columns of the SM for the CTA and frees them at the end)
__device__ __forceinline__ void tmem_alloc(uint32_t smem_dst, uint32_t ncols)
{
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_dst), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t ncols)
{
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// 32 lanes x 16 columns: thread i of the warp <-> TMEM lane 32 (warp % 4) + i
__device__ __forceinline__ void st16(uint32_t addr, const uint32_t (&v)[16])
{
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
                 ::"r"(addr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]),
                 "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]) : "memory");
}
__device__ __forceinline__ void ld16(uint32_t addr, uint32_t (&v)[16])
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
                   "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]) : "r"(addr) : "memory");
}
__device__ __forceinline__ void ld64(uint32_t addr, uint32_t (&v)[64])
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x64.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                 "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, "
                 "%32, %33, %34, %35, %36, %37, %38, %39, %40, %41, %42, %43, %44, %45, %46, %47, "
                 "%48, %49, %50, %51, %52, %53, %54, %55, %56, %57, %58, %59, %60, %61, %62, %63}, [%64];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
                   "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]),
                   "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]),
                   "=r"(v[30]), "=r"(v[31]), "=r"(v[32]), "=r"(v[33]), "=r"(v[34]), "=r"(v[35]), "=r"(v[36]), "=r"(v[37]), "=r"(v[38]), "=r"(v[39]),
                   "=r"(v[40]), "=r"(v[41]), "=r"(v[42]), "=r"(v[43]), "=r"(v[44]), "=r"(v[45]), "=r"(v[46]), "=r"(v[47]), "=r"(v[48]), "=r"(v[49]),
                   "=r"(v[50]), "=r"(v[51]), "=r"(v[52]), "=r"(v[53]), "=r"(v[54]), "=r"(v[55]), "=r"(v[56]), "=r"(v[57]), "=r"(v[58]), "=r"(v[59]),
                   "=r"(v[60]), "=r"(v[61]), "=r"(v[62]), "=r"(v[63]) : "r"(addr) : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem descriptor]; M = 128, K = 32 bytes, u8 x u8 -> s32
__device__ __forceinline__ void mma_i8_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate)
{
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n\t}"
                 ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void mma_commit(uint32_t bar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    uint32_t done, spins = 0;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(bar), "r"(parity) : "memory");
        if (!done && ++spins > (1u << 22)) __trap();   // a lost MMA is a bug: fail loudly instead of hanging the GPU
    } while (!done);
}
// instruction descriptor: D = s32 (2 << 4), A = B = u8 (format 0), both K-major, N >> 3 at bit 17, M >> 4 at bit 24
__host__ __device__ constexpr uint32_t make_idesc(int n) { return (2u << 4) | ((uint32_t)(n >> 3) << 17) | ((128u >> 4) << 24); }
// shared-memory matrix descriptor, no swizzle: start >> 4, leading (K) byte offset >> 4 at bit 16, stride (N) byte offset
// >> 4 at bit 32, descriptor version 1 at bit 46
__device__ __forceinline__ uint64_t make_sdesc(uint32_t addr, uint32_t lbo, uint32_t sbo)
{
    return (uint64_t)((addr >> 4) & 0x3fff) | ((uint64_t)((lbo >> 4) & 0x3fff) << 16) | ((uint64_t)((sbo >> 4) & 0x3fff) << 32) | (1ull << 46);
}

// ---- the weight tile: 64 (n = 4 r + j) x 64 (k = 4 c + i) bytes, K-major, no swizzle: core matrix = 8 n x 16 k bytes
constexpr uint32_t B_KCHUNK = 1024, B_NGROUP = 128, B_BYTES = 4096;
__host__ __device__ constexpr int b_off(int n, int k) { return (k >> 4) * B_KCHUNK + (n >> 3) * B_NGROUP + (n & 7) * 16 + (k & 15); }

// Builds the weight tile of one stage in shared memory, all `nthreads` (a multiple of 64) threads of the caller cooperating;
// ends with a __syncthreads, so the whole CTA must call it.  tw[(8 >> k) + j] = PRE-DOUBLED twiddle j of digit bit k
// (k = 0..3; 8 >> k twiddles per level: the unit-level heap of fft_tma.cu, levels sl..sl+3).  INV = inverse butterflies
// in the order k = 0..3, else forward butterflies k = 3..0.  scale2 != 0: every entry is multiplied by scale2 / 2 (the
// 1 / 2^n of the last inverse pass).  No scratch memory: thread t computes column c = t % 16 of W itself (32 butterflies
// on the unit vector e_c) and expands the rows r = t / 64 + 8 m with the byte shift i = (t / 16) % 4.
template <bool INV>
__device__ __forceinline__ void build_weights(uint8_t *sB, const uint32_t *tw, uint32_t scale2, int tid, int nthreads)
{
    const int c = tid & 15, i = (tid >> 4) & 3;
    uint32_t v[16];
#pragma unroll
    for (int d = 0; d < 16; d++) v[d] = d == c ? 1u : 0u;
#pragma unroll
    for (int kk = 0; kk < 4; kk++) {
        const int k = INV ? kk : 3 - kk;
#pragma unroll
        for (int d = 0; d < 16; d++) {
            if (d & (1 << k)) continue;
            const uint32_t t = tw[(8 >> k) + (d >> (k + 1))];
            if (INV) m31::ibutterfly(v[d], v[d | (1 << k)], t);
            else m31::butterfly(v[d], v[d | (1 << k)], t);
        }
    }
    for (int r0 = tid >> 6; r0 < 16; r0 += nthreads >> 6) {
        uint32_t x = 0;
#pragma unroll
        for (int r = 0; r < 16; r++) x = r == r0 ? v[r] : x;
        if (scale2) x = m31::mul2(x, scale2);
        const uint32_t w = m31::mul2(x, 2u << (8 * i));   // W[r0][c] * 2^(8 i) mod P
#pragma unroll
        for (int j = 0; j < 4; j++) sB[b_off(4 * r0 + j, 4 * c + i)] = (uint8_t)(w >> (8 * j));
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> tensor-core (async-proxy) reads
    __syncthreads();
}

// a0 + 2^8 a1 + 2^16 a2 + 2^24 a3 (mod P) in [0, 2P), for a_j <= 64 * 255^2 (< 2^22):
//   lo = a0 + 2^8 a1 and hi = a2 + 2^8 a3 are <= 257 * 64 * 255^2 < 2^30;
//   2^16 hi = 2^31 (hi >> 15) + 2^16 (hi & 0x7fff) = (hi >> 15) + ((hi << 17) mod 2^32 >> 1)   (mod P)
//   sum < 2^30 + 2^15 + 2^31 < 2P, so ONE conditional subtraction makes it canonical.
// 6 instructions: IMAD, IMAD, IMAD (FMA pipe); LEA.HI, LOP3, VIADDMNMX (ALU pipe) -- two butterflies are 14 (8 on the ALU pipe).
__device__ __forceinline__ uint32_t recombine(uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3)
{
    const uint32_t lo = a0 + (a1 << 8), hi = a2 + (a3 << 8);
    return m31::csub((hi & 0x7fffu) * 65536u + (lo + (hi >> 15)));
}

// What a 128-thread MMA group (four consecutive warps: one per TMEM lane quarter) needs for its stages.  All fields are
// warp-uniform; the TMEM addresses are passed through __shfl_sync by the caller so that ptxas keeps them in uniform
// registers (otherwise every tcgen05.ld/st is wrapped in a waterfall loop: 12.4 -> ~8 instructions per element).
struct Group {
    uint32_t tmem;        // column base of the group (lane field 0): A_e at + 16 e (e = 0..3), D at + 64; 128 columns
    uint32_t tmem_lane;   // tmem + (32 * (warp % 4)) << 16: the quarter this warp may access
    uint32_t bar;         // shared address of the group's mbarrier (MMA completion); 4 phases per stage call
    int bar_a, bar_d;     // named barrier ids: "A stored" (sync, 128), "D read" (issuing warp syncs, others arrive)
    bool issuer_warp;     // warp 0 of the group; its lane 0 issues the MMAs
};

// One stage on the 16 x 4 values a thread holds (v[d][e]: vector e of the thread, element d): in place, outputs canonical.
// b_desc0 / b_desc1 describe the K halves 0..31 / 32..63 of the weight tile.
__device__ __forceinline__ void stage16(uint32_t (&v)[16][4], const Group &gr, uint64_t b_desc0, uint64_t b_desc1, int lane)
{
    constexpr uint32_t IDESC = make_idesc(64);
#pragma unroll
    for (int e = 0; e < 4; e++) {
        uint32_t a[16];
#pragma unroll
        for (int d = 0; d < 16; d++) a[d] = v[d][e];
        st16(gr.tmem_lane + 16 * e, a);
    }
    wait_st();
    fence_before();
    asm volatile("bar.sync %0, 128;" ::"r"(gr.bar_a) : "memory");
#pragma unroll
    for (int e = 0; e < 4; e++) {
        if (gr.issuer_warp && lane == 0) {
            fence_after();
            mma_i8_ts(gr.tmem + 64, gr.tmem + 16 * e, b_desc0, IDESC, 0);
            mma_i8_ts(gr.tmem + 64, gr.tmem + 16 * e + 8, b_desc1, IDESC, 1);
            mma_commit(gr.bar);
        }
        __syncwarp();
        mbar_wait(gr.bar, e & 1);
        fence_after();
        uint32_t acc[16];
#pragma unroll
        for (int h = 0; h < 4; h++) {
            ld16(gr.tmem_lane + 64 + 16 * h, acc);
            wait_ld();
            if (h == 3 && e < 3) {                     // D has been read: the next MMA pair may overwrite it
                fence_before();
                if (gr.issuer_warp) asm volatile("bar.sync %0, 128;" ::"r"(gr.bar_d) : "memory");
                else asm volatile("bar.arrive %0, 128;" ::"r"(gr.bar_d) : "memory");
            }
#pragma unroll
            for (int r = 0; r < 4; r++) v[4 * h + r][e] = recombine(acc[4 * r], acc[4 * r + 1], acc[4 * r + 2], acc[4 * r + 3]);
        }
    }
}

// The same stage applied IN PLACE to vectors that live in shared memory, one vector per thread and batch: element c of the
// vector of (batch b, this thread) is the word base[b * stride_b + c * stride_c], where consecutive lanes of a warp own
// consecutive words (LDS.32 / STS.32 without bank conflicts, no register transposition, ~50 registers).  Software
// pipeline over the NB batches (NB even): A is double-buffered in TMEM (columns +0 / +16), so batch b + 1 is loaded and
// stored to TMEM while the MMA pair of batch b runs; ONE 128-thread barrier per batch says "A(b+1) stored and D(b) read".
// The caller orders the shared-memory accesses against the neighbouring stages (a group barrier before and after).
template <int NB>
__device__ __forceinline__ void stage16_smem(uint32_t *base, int stride_c, int stride_b, const Group &gr, uint64_t b_desc0, uint64_t b_desc1, int lane)
{
    static_assert(NB % 2 == 0, "the mbarrier phase must come back to 0");
    constexpr uint32_t IDESC = make_idesc(64);
    uint32_t v[16];
#pragma unroll
    for (int c = 0; c < 16; c++) v[c] = base[c * stride_c];
    st16(gr.tmem_lane, v);
    wait_st();
    fence_before();
    asm volatile("bar.sync %0, 128;" ::"r"(gr.bar_a) : "memory");
    if (gr.issuer_warp && lane == 0) {
        fence_after();
        mma_i8_ts(gr.tmem + 64, gr.tmem, b_desc0, IDESC, 0);
        mma_i8_ts(gr.tmem + 64, gr.tmem + 8, b_desc1, IDESC, 1);
        mma_commit(gr.bar);
    }
    __syncwarp();
#pragma unroll
    for (int b = 0; b < NB; b++) {
        uint32_t *cur = base + b * stride_b;
        if (b + 1 < NB) {
#pragma unroll
            for (int c = 0; c < 16; c++) v[c] = cur[stride_b + c * stride_c];
            st16(gr.tmem_lane + 16 * ((b + 1) & 1), v);
            wait_st();
        }
        mbar_wait(gr.bar, b & 1);
        fence_after();
        uint32_t acc[64];                              // all 64 accumulators of the vector in one TMEM round trip
        ld64(gr.tmem_lane + 64, acc);
        wait_ld();
#pragma unroll
        for (int r = 0; r < 16; r++) cur[r * stride_c] = recombine(acc[4 * r], acc[4 * r + 1], acc[4 * r + 2], acc[4 * r + 3]);
        if (b + 1 < NB) {
            fence_before();
            asm volatile("bar.sync %0, 128;" ::"r"(gr.bar_a) : "memory");
            if (gr.issuer_warp && lane == 0) {
                fence_after();
                mma_i8_ts(gr.tmem + 64, gr.tmem + 16 * ((b + 1) & 1), b_desc0, IDESC, 0);
                mma_i8_ts(gr.tmem + 64, gr.tmem + 16 * ((b + 1) & 1) + 8, b_desc1, IDESC, 1);
                mma_commit(gr.bar);
            }
            __syncwarp();
        }
    }
}

}  // namespace tc
}  // namespace tstwo
