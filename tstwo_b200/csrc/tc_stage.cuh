// A radix-16 circle-FFT stage (four butterfly layers, packages/core/src/fft.ts:12-30) on the 5th-generation tensor
// cores of sm_100a: tcgen05.mma kind::i8, operands u8, accumulators s32 in TMEM.
//
// The four layers of a stage whose twiddles are the same for every vector of a tile (the TOP stage of a pass: its
// twiddles depend only on index bits above the tile) are ONE 16 x 16 matrix W over M31 applied to many 16-element
// vectors.  Exactly, by byte splitting: with f = sum_i 2^(8i) f_i and (2^(8i) W[r][c]) mod P = sum_j 2^(8j) w[r][c][i][j],
//     out[r] = sum_j 2^(8j) * ( sum_(c,i) f_(c,i) * w[r][c][i][j] )      (mod P)
// i.e. a [vectors x 64] x [64 x 64] u8 product; every accumulator is < 64 * 255^2 < 2^22 and the four of an output are
// recombined mod P on the integer pipes (~6 instructions + 3 for the canonical form, against 14 for two butterflies).
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

// Builds the weight tile of one stage in shared memory, all `nthreads` threads of the caller cooperating (barrier id 0 =
// __syncthreads, so the whole CTA must call it).  tw[(8 >> k) + j] = PRE-DOUBLED twiddle j of digit bit k (k = 0..3;
// 8 >> k twiddles per level: the unit-level heap of fft_tma.cu, levels sl..sl+3).  INV = inverse butterflies in the order
// k = 0..3, else forward butterflies k = 3..0.  scale2 != 0: every entry is multiplied by scale2 / 2 (the 1 / 2^n of the
// last inverse pass).  `scratch` = 256 words of shared memory.
template <bool INV>
__device__ __forceinline__ void build_weights(uint8_t *sB, uint32_t *scratch, const uint32_t *tw, uint32_t scale2, int tid, int nthreads)
{
    if (tid < 16) {                                    // column c of W = the stage applied to the unit vector e_c
        uint32_t v[16];
#pragma unroll
        for (int d = 0; d < 16; d++) v[d] = d == tid ? 1u : 0u;
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
#pragma unroll
        for (int r = 0; r < 16; r++) scratch[r * 16 + tid] = scale2 ? m31::mul2(v[r], scale2) : v[r];
    }
    __syncthreads();
    for (int e = tid; e < 1024; e += nthreads) {       // e = (r, c, i): the four bytes j of W[r][c] * 2^(8 i) mod P
        const int i = e & 3, c = (e >> 2) & 15, r = e >> 6;
        const uint32_t w = m31::mul2(scratch[r * 16 + c], 2u << (8 * i));
#pragma unroll
        for (int j = 0; j < 4; j++) sB[b_off(4 * r + j, 4 * c + i)] = (uint8_t)(w >> (8 * j));
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> tensor-core (async-proxy) reads
    __syncthreads();
}

// any u32 congruent to a0 + 2^8 a1 + 2^16 a2 + 2^24 a3 (a_j < 2^22)
__device__ __forceinline__ uint32_t recombine(uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3)
{
    const uint32_t lo = a0 + (a1 << 8), hi = a2 + (a3 << 8);          // < 2^31 each
    const uint64_t p = (uint64_t)hi * 0x20000u;                        // 2 * (hi * 2^16): the fold of mul2()
    return lo + (uint32_t)(p >> 32) + ((uint32_t)p >> 1);              // < 2^32
}
__device__ __forceinline__ uint32_t canonical(uint32_t x) { return m31::csub((x & m31::P) + (x >> 31)); }

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
        uint32_t acc[2][16];
        ld16(gr.tmem_lane + 64, acc[0]);
#pragma unroll
        for (int h = 0; h < 4; h++) {
            wait_ld();
            if (h < 3) ld16(gr.tmem_lane + 64 + 16 * (h + 1), acc[(h + 1) & 1]);
            else if (e < 3) {                          // D has been read: the next MMA pair may overwrite it
                fence_before();
                if (gr.issuer_warp) asm volatile("bar.sync %0, 128;" ::"r"(gr.bar_d) : "memory");
                else asm volatile("bar.arrive %0, 128;" ::"r"(gr.bar_d) : "memory");
            }
#pragma unroll
            for (int r = 0; r < 4; r++) {
                const uint32_t *a = acc[h & 1] + 4 * r;
                v[4 * h + r][e] = canonical(recombine(a[0], a[1], a[2], a[3]));
            }
        }
    }
}

}  // namespace tc
}  // namespace tstwo
