// Blake2s Merkle layers for sm_100a: MerkleOps.commitOnLayer / MerkleProver.commit of the reference
// (packages/core/src/vcs/ops.ts:16-26, vcs/prover.ts:13-30) with Blake2sMerkleHasher.hashNode semantics
// (vcs/blake2_merkle.ts:9-24): node_i = BLAKE2s-256( prev[2i] || prev[2i+1] || LE32(col_0[i]) || LE32(col_1[i]) ... ).
// The hash itself is plain RFC 7693 BLAKE2s (the reference calls @noble/hashes 1.8.0, vcs/blake2_hash.ts:52-71).
//
// One thread per node.  The message words of a node come straight from column-major loads: for a fixed column
// the 32 lanes of a warp read 32 consecutive rows (128 B, fully coalesced).  Rotations by 16/8 are PRMT, by
// 12/7 funnel shifts; the ten rounds are fully unrolled with compile-time message schedules.
#include <cuda.h>

#include <vector>

#include "common.cuh"
#include "fft_params.cuh"

namespace tstwo {

__device__ __forceinline__ uint32_t rotr16(uint32_t x) { return __byte_perm(x, x, 0x1032); }
__device__ __forceinline__ uint32_t rotr8(uint32_t x) { return __byte_perm(x, x, 0x0321); }
__device__ __forceinline__ uint32_t rotr12(uint32_t x) { return __funnelshift_r(x, x, 12); }
__device__ __forceinline__ uint32_t rotr7(uint32_t x) { return __funnelshift_r(x, x, 7); }

// Additions are issued as IMAD (x * one + y, `one` = 1 passed as a kernel argument so ptxas cannot fold it back
// into IADD3): on sm_100 the xor/rotate work saturates the ALU pipe (64 lanes/clk/SM), the FMA pipe is idle, so
// moving all 6 two-input adds of a G there cuts the ALU work per G from 10 to 8 instructions (profiles/r01).
#define B2S_ADD(x, y) ((x) * one + (y))
#define B2S_G(a, b, c, d, x, y)                                     \
    do {                                                            \
        a = B2S_ADD(B2S_ADD(a, b), (x)); d = rotr16(d ^ a); c = B2S_ADD(c, d); b = rotr12(b ^ c); \
        a = B2S_ADD(B2S_ADD(a, b), (y)); d = rotr8(d ^ a);  c = B2S_ADD(c, d); b = rotr7(b ^ c);  \
    } while (0)

#define B2S_ROUND(s0, s1, s2, s3, s4, s5, s6, s7, s8, s9, s10, s11, s12, s13, s14, s15) \
    do {                                                                                  \
        B2S_G(v0, v4, v8, v12, m[s0], m[s1]);   B2S_G(v1, v5, v9, v13, m[s2], m[s3]);     \
        B2S_G(v2, v6, v10, v14, m[s4], m[s5]);  B2S_G(v3, v7, v11, v15, m[s6], m[s7]);    \
        B2S_G(v0, v5, v10, v15, m[s8], m[s9]);  B2S_G(v1, v6, v11, v12, m[s10], m[s11]);  \
        B2S_G(v2, v7, v8, v13, m[s12], m[s13]); B2S_G(v3, v4, v9, v14, m[s14], m[s15]);   \
    } while (0)

// one compression; t = byte counter (< 2^32 on this path), last = final-block flag
__device__ __forceinline__ void b2s_compress(uint32_t (&h)[8], const uint32_t (&m)[16], uint32_t t, bool last, uint32_t one)
{
    uint32_t v0 = h[0], v1 = h[1], v2 = h[2], v3 = h[3], v4 = h[4], v5 = h[5], v6 = h[6], v7 = h[7];
    uint32_t v8 = 0x6A09E667u, v9 = 0xBB67AE85u, v10 = 0x3C6EF372u, v11 = 0xA54FF53Au;
    uint32_t v12 = 0x510E527Fu ^ t, v13 = 0x9B05688Cu, v14 = last ? ~0x1F83D9ABu : 0x1F83D9ABu, v15 = 0x5BE0CD19u;
    B2S_ROUND(0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15);
    B2S_ROUND(14, 10, 4, 8, 9, 15, 13, 6, 1, 12, 0, 2, 11, 7, 5, 3);
    B2S_ROUND(11, 8, 12, 0, 5, 2, 15, 13, 10, 14, 3, 6, 7, 1, 9, 4);
    B2S_ROUND(7, 9, 3, 1, 13, 12, 11, 14, 2, 6, 5, 10, 4, 0, 15, 8);
    B2S_ROUND(9, 0, 5, 7, 2, 4, 10, 15, 14, 1, 11, 12, 6, 8, 3, 13);
    B2S_ROUND(2, 12, 6, 10, 0, 11, 8, 3, 4, 13, 7, 5, 15, 14, 1, 9);
    B2S_ROUND(12, 5, 1, 15, 14, 13, 4, 10, 0, 7, 6, 3, 9, 2, 8, 11);
    B2S_ROUND(13, 11, 7, 14, 12, 1, 3, 9, 5, 0, 15, 4, 8, 6, 2, 10);
    B2S_ROUND(6, 15, 14, 9, 11, 3, 0, 8, 12, 2, 13, 7, 1, 4, 10, 5);
    B2S_ROUND(10, 2, 8, 4, 7, 6, 1, 5, 15, 11, 9, 14, 3, 12, 13, 0);
    h[0] ^= v0 ^ v8;  h[1] ^= v1 ^ v9;  h[2] ^= v2 ^ v10; h[3] ^= v3 ^ v11;
    h[4] ^= v4 ^ v12; h[5] ^= v5 ^ v13; h[6] ^= v6 ^ v14; h[7] ^= v7 ^ v15;
}

__device__ __forceinline__ void b2s_init(uint32_t (&h)[8])
{
    h[0] = 0x6A09E667u ^ 0x01010020u; h[1] = 0xBB67AE85u; h[2] = 0x3C6EF372u; h[3] = 0xA54FF53Au;
    h[4] = 0x510E527Fu; h[5] = 0x9B05688Cu; h[6] = 0x1F83D9ABu; h[7] = 0x5BE0CD19u;
}

// HAS_PREV: the first 16 message words are the two child hashes.  The next block's 16 message words are loaded
// before the current block is compressed (software pipelining hides the column loads behind ~1200 instructions).
template <bool HAS_PREV>
__global__ void __launch_bounds__(256) merkle_layer_kernel(const uint4 *__restrict__ prev, ColSet cols, int n_cols,
                                                           uint64_t n_nodes, uint4 *__restrict__ out, uint32_t one)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_nodes) return;
    uint32_t h[8], m[16], nx[16];
    b2s_init(h);
    const uint32_t total_words = (HAS_PREV ? 16u : 0u) + (uint32_t)n_cols;
    const uint32_t n_blocks = total_words == 0 ? 1u : (total_words + 15u) / 16u;
    // block b covers message words [16b, 16b+16); word w is a child-hash word (w < 16, HAS_PREV) or column w - off
    const int off = HAS_PREV ? 16 : 0;
    auto load_block = [&](uint32_t b, uint32_t (&dst)[16]) {
        if (HAS_PREV && b == 0) {
            const uint4 *p = prev + 4 * i;   // 64 bytes = children 2i, 2i+1
            uint4 x0 = __ldcs(p), x1 = __ldcs(p + 1), x2 = __ldcs(p + 2), x3 = __ldcs(p + 3);
            dst[0] = x0.x; dst[1] = x0.y; dst[2] = x0.z; dst[3] = x0.w; dst[4] = x1.x; dst[5] = x1.y; dst[6] = x1.z; dst[7] = x1.w;
            dst[8] = x2.x; dst[9] = x2.y; dst[10] = x2.z; dst[11] = x2.w; dst[12] = x3.x; dst[13] = x3.y; dst[14] = x3.z; dst[15] = x3.w;
        } else {
            const int c0 = (int)(16 * b) - off;
            if (c0 + 16 <= n_cols) {
#pragma unroll
                for (int k = 0; k < 16; k++) dst[k] = __ldcs(cols.col(c0 + k) + i);
            } else {
#pragma unroll
                for (int k = 0; k < 16; k++) dst[k] = (c0 + k < n_cols) ? __ldcs(cols.col(c0 + k) + i) : 0u;
            }
        }
    };
    load_block(0, m);
    for (uint32_t b = 0; b + 1 < n_blocks; b++) {
        load_block(b + 1, nx);
        b2s_compress(h, m, (b + 1) * 64u, false, one);
#pragma unroll
        for (int k = 0; k < 16; k++) m[k] = nx[k];
    }
    b2s_compress(h, m, total_words * 4u, true, one);
    out[2 * i] = make_uint4(h[0], h[1], h[2], h[3]);
    out[2 * i + 1] = make_uint4(h[4], h[5], h[6], h[7]);
}

// Pure node layer (no column values at this size: every layer of a tree below its smallest column, all of a FRI layer tree but
// its leaves): one 64-byte block per node.  Its SHAPE is chosen for ptxas: with __launch_bounds__(256, 5) (<= 51 registers) and the
// compression inside a run-time loop (n_blocks = 1 at run time) the additions come out as IMAD R, R, UR, R -- `one` in a uniform
// register, 1345 clk per warp-compression -- where merkle_layer_kernel's straight-line copies get the three-register form (1556 clk,
// profiles/r01_microbench_blake2s_forms.txt).  tests/test_abi_cpu.py disassembles the built library and fails if the form is lost.
__global__ void __launch_bounds__(256, 5) merkle_node_kernel(const uint4 *__restrict__ prev, uint64_t n_nodes, uint4 *__restrict__ out,
                                                             uint32_t n_blocks, uint32_t one)
{
    const uint64_t i = (uint64_t)blockIdx.x * 256 + threadIdx.x;   // literal block size: with blockDim.x ptxas falls back to the register form
    if (i >= n_nodes) return;
    uint32_t h[8], m[16];
    const uint4 *p = prev + 4 * i;   // 64 bytes = children 2i, 2i+1
    const uint4 x0 = __ldcs(p), x1 = __ldcs(p + 1), x2 = __ldcs(p + 2), x3 = __ldcs(p + 3);
    m[0] = x0.x; m[1] = x0.y; m[2] = x0.z; m[3] = x0.w; m[4] = x1.x; m[5] = x1.y; m[6] = x1.z; m[7] = x1.w;
    m[8] = x2.x; m[9] = x2.y; m[10] = x2.z; m[11] = x2.w; m[12] = x3.x; m[13] = x3.y; m[14] = x3.z; m[15] = x3.w;
    b2s_init(h);
#pragma unroll 1
    for (uint32_t b = 0; b < n_blocks; b++) {        // n_blocks == 1: a node is exactly one final block
        const bool last = b + 1 == n_blocks;
        b2s_compress(h, m, (b + 1) * 64u, last, one);
    }
    out[2 * i] = make_uint4(h[0], h[1], h[2], h[3]);
    out[2 * i + 1] = make_uint4(h[4], h[5], h[6], h[7]);
}

// Node layer that also absorbs at most 16 evenly spaced columns (two blocks per node: the children, then the column values): the
// layer at which the smaller columns of a mixed-size tree join (FRI's first-layer tree over quotient columns of two sizes).  Same
// shape as merkle_node_kernel; the columns are addressed as cols.base + k * cols.stride on purpose -- through ColSet::col() (the pointer-table
// branch) ptxas falls back to the three-register IMAD form.  Scattered column tables keep merkle_layer_kernel<true>.
__global__ void __launch_bounds__(256, 5) merkle_node_cols_kernel(const uint4 *__restrict__ prev, ColSet cols, int n_cols, uint64_t n_nodes,
                                                                  uint4 *__restrict__ out, uint32_t n_blocks, uint32_t one)
{
    const uint64_t i = (uint64_t)blockIdx.x * 256 + threadIdx.x;   // literal block size (see merkle_node_kernel)
    if (i >= n_nodes) return;
    uint32_t h[8], m[16];
    const uint4 *p = prev + 4 * i;
    const uint4 x0 = __ldcs(p), x1 = __ldcs(p + 1), x2 = __ldcs(p + 2), x3 = __ldcs(p + 3);
    m[0] = x0.x; m[1] = x0.y; m[2] = x0.z; m[3] = x0.w; m[4] = x1.x; m[5] = x1.y; m[6] = x1.z; m[7] = x1.w;
    m[8] = x2.x; m[9] = x2.y; m[10] = x2.z; m[11] = x2.w; m[12] = x3.x; m[13] = x3.y; m[14] = x3.z; m[15] = x3.w;
    b2s_init(h);
#pragma unroll 1
    for (uint32_t b = 0; b < n_blocks; b++) {        // n_blocks == 2 at run time
        const bool last = b + 1 == n_blocks;
        b2s_compress(h, m, last ? 64u + 4u * (uint32_t)n_cols : 64u, last, one);
        if (!last) {
#pragma unroll
            for (int k = 0; k < 16; k++) m[k] = k < n_cols ? __ldcs(cols.base + (uint64_t)k * cols.stride + i) : 0u;
        }
    }
    out[2 * i] = make_uint4(h[0], h[1], h[2], h[3]);
    out[2 * i + 1] = make_uint4(h[4], h[5], h[6], h[7]);
}

// ---- leaf layer fed by bulk tensor copies -----------------------------------------------------------------------
// The leaf hash of row i absorbs column words c = 0, 1, ... of that row, 16 per compression.  With uniformly spaced
// columns the message blocks of 256 consecutive rows are boxes [16 columns x 256 rows] (16 KiB, sixteen 1 KiB runs) of
// one 2-D tensor (rows fastest).  A CTA keeps two boxes in shared memory: while it compresses block b the copy engine
// fetches block b + 1; the warp that is LAST to pull its message words out of a box (a shared-memory counter tells it)
// immediately asks for the box two blocks ahead, so no warp ever waits for a designated producer.  A thread's 16 message
// words are 16 conflict-free shared-memory loads with immediate offsets: no per-column address arithmetic (~10 % of
// the ALU slots of merkle_layer_kernel) and no register double buffer, which is what lets 5 CTAs (40 warps) fit an SM.
// Columns past n_cols are out of bounds of the tensor and arrive as zeros: exactly BLAKE2s's padding of the last block.
// (Per-warp boxes of [16 x 32] rows were tried: sixteen 128-byte runs per 2 KiB make the copy engine the bottleneck.)
constexpr int LEAF_ROWS = 256, LEAF_STAGES = 2, LEAF_STAGE_BYTES = 16 * LEAF_ROWS * 4;
constexpr int LEAF_SMEM = LEAF_STAGES * LEAF_STAGE_BYTES + 64;   // stages, then LEAF_STAGES barriers and LEAF_STAGES counters

__device__ __forceinline__ uint32_t leaf_smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
// The wait is one asm block (try_wait, branch back) rather than a C++ loop on purpose: with a C++ loop AND the lane-0 branch
// below in the hash loop, ptxas keeps `one` (B2S_ADD) in a vector register and allocates registers for three-operand IMADs,
// which leaves ~350 same-bank operand pairs per compression; with this form it uses IMAD R, R, UR, R and the pairs drop
// below 30 (cuobjdump -sass; dispatch stalls 0.9 -> 0.4 per issue).  A lost box traps after 2^22 polls instead of hanging.
__device__ __forceinline__ void leaf_mbar_wait(uint32_t bar, uint32_t parity)
{
    asm volatile("{\n\t.reg .pred p;\n\t.reg .u32 n;\n\tmov.u32 n, 0;\n"
                 "LEAF_WAIT:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra LEAF_DONE;\n\t"
                 "add.u32 n, n, 1;\n\tsetp.lt.u32 p, n, 0x400000;\n\t@p bra LEAF_WAIT;\n\ttrap;\n"
                 "LEAF_DONE:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void leaf_fetch(uint32_t dst, const CUtensorMap *map, uint32_t row0, uint32_t col0, uint32_t bar)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(LEAF_STAGE_BYTES) : "memory");
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(dst), "l"(map), "r"(row0), "r"(col0), "r"(bar) : "memory");
}
__device__ __forceinline__ uint32_t leaf_lds(uint32_t addr)
{
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
    return v;
}

__global__ void __launch_bounds__(LEAF_ROWS, 5) merkle_leaf_tma_kernel(const __grid_constant__ CUtensorMap map, int n_cols,
                                                                       uint4 *__restrict__ out, uint32_t one)
{
    extern __shared__ __align__(128) uint8_t leaf_smem[];
    const uint32_t base = leaf_smem_u32(leaf_smem);                 // stage s at base + s * LEAF_STAGE_BYTES
    const uint32_t ctl = base + LEAF_STAGES * LEAF_STAGE_BYTES;     // barrier s at ctl + 8 s, reader counter s at ctl + 32 + 4 s
    const uint32_t n_blocks = ((uint32_t)n_cols + 15u) / 16u;
    if (threadIdx.x == 0) {
#pragma unroll
        for (int s = 0; s < LEAF_STAGES; s++) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(ctl + 8 * s), "r"(1) : "memory");
            asm volatile("st.shared.u32 [%0], %1;" ::"r"(ctl + 32 + 4 * s), "r"(0) : "memory");
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
#pragma unroll
        for (uint32_t s = 0; s < LEAF_STAGES; s++)
            if (s < n_blocks) leaf_fetch(base + s * LEAF_STAGE_BYTES, &map, blockIdx.x * LEAF_ROWS, 16 * s, ctl + 8 * s);
    }
    __syncthreads();
    uint32_t h[8], m[16];
    b2s_init(h);
    uint32_t s = 0, ph = 0;          // stage and phase of block b
#pragma unroll 1
    for (uint32_t b = 0; b < n_blocks; b++) {
        leaf_mbar_wait(ctl + 8 * s, ph);
#pragma unroll
        for (int k = 0; k < 16; k++) m[k] = leaf_lds(base + s * LEAF_STAGE_BYTES + 4 * threadIdx.x + 4 * LEAF_ROWS * k);
        __syncwarp();
        if ((threadIdx.x & 31) == 0) {
            // shared-memory operations of a warp are performed in order: the loads above precede this counter update,
            // and the refill it may trigger is ordered after the update of every warp
            uint32_t old;
            asm volatile("atom.shared.add.u32 %0, [%1], 1;" : "=r"(old) : "r"(ctl + 32 + 4 * s) : "memory");
            if ((old & (LEAF_ROWS / 32 - 1)) == LEAF_ROWS / 32 - 1 && b + LEAF_STAGES < n_blocks)
                leaf_fetch(base + s * LEAF_STAGE_BYTES, &map, blockIdx.x * LEAF_ROWS, 16 * (b + LEAF_STAGES), ctl + 8 * s);
        }
        const bool last = b + 1 == n_blocks;
        b2s_compress(h, m, last ? (uint32_t)n_cols * 4u : (b + 1) * 64u, last, one);
        if (++s == LEAF_STAGES) { s = 0; ph ^= 1; }
    }
    const uint64_t i = (uint64_t)blockIdx.x * LEAF_ROWS + threadIdx.x;
    out[2 * i] = make_uint4(h[0], h[1], h[2], h[3]);
    out[2 * i + 1] = make_uint4(h[4], h[5], h[6], h[7]);
}

// The same kernel for a leaf layer that is hashed in COLUMN CHUNKS (multi-GPU: the first half of the columns of a rank's rows has
// arrived long before the last, tstwo_merkle_leaf_absorb): the map describes columns [c0, c0 + nc) only, the chaining value of
// every row lives in its slot of the leaf layer between chunks (a leaf hash is 32 bytes, so is a BLAKE2s state), the byte counter
// continues at 4 c0 and only the chunk that holds the last column finalises.  A separate kernel on purpose: merkle_leaf_tma_kernel's
// register allocation and IMAD form (see leaf_mbar_wait) are left exactly as they are.
__global__ void __launch_bounds__(LEAF_ROWS, 5) merkle_leaf_chunk_kernel(const __grid_constant__ CUtensorMap map, int nc, int c0, int total_cols,
                                                                         uint4 *__restrict__ out, uint32_t one)
{
    extern __shared__ __align__(128) uint8_t leaf_smem[];
    const uint32_t base = leaf_smem_u32(leaf_smem);
    const uint32_t ctl = base + LEAF_STAGES * LEAF_STAGE_BYTES;
    const uint32_t n_blocks = ((uint32_t)nc + 15u) / 16u;
    if (threadIdx.x == 0) {
#pragma unroll
        for (int s = 0; s < LEAF_STAGES; s++) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(ctl + 8 * s), "r"(1) : "memory");
            asm volatile("st.shared.u32 [%0], %1;" ::"r"(ctl + 32 + 4 * s), "r"(0) : "memory");
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
#pragma unroll
        for (uint32_t s = 0; s < LEAF_STAGES; s++)
            if (s < n_blocks) leaf_fetch(base + s * LEAF_STAGE_BYTES, &map, blockIdx.x * LEAF_ROWS, 16 * s, ctl + 8 * s);
    }
    __syncthreads();
    const uint64_t i = (uint64_t)blockIdx.x * LEAF_ROWS + threadIdx.x;
    uint32_t h[8], m[16];
    b2s_init(h);
    if (c0 != 0) {
        const uint4 a = out[2 * i], b = out[2 * i + 1];
        h[0] = a.x; h[1] = a.y; h[2] = a.z; h[3] = a.w; h[4] = b.x; h[5] = b.y; h[6] = b.z; h[7] = b.w;
    }
    const uint32_t t0 = 4u * (uint32_t)c0;
    const bool closes = c0 + nc == total_cols;
    uint32_t s = 0, ph = 0;
#pragma unroll 1
    for (uint32_t b = 0; b < n_blocks; b++) {
        leaf_mbar_wait(ctl + 8 * s, ph);
#pragma unroll
        for (int k = 0; k < 16; k++) m[k] = leaf_lds(base + s * LEAF_STAGE_BYTES + 4 * threadIdx.x + 4 * LEAF_ROWS * k);
        __syncwarp();
        if ((threadIdx.x & 31) == 0) {
            uint32_t old;
            asm volatile("atom.shared.add.u32 %0, [%1], 1;" : "=r"(old) : "r"(ctl + 32 + 4 * s) : "memory");
            if ((old & (LEAF_ROWS / 32 - 1)) == LEAF_ROWS / 32 - 1 && b + LEAF_STAGES < n_blocks)
                leaf_fetch(base + s * LEAF_STAGE_BYTES, &map, blockIdx.x * LEAF_ROWS, 16 * (b + LEAF_STAGES), ctl + 8 * s);
        }
        const bool last = closes && b + 1 == n_blocks;
        b2s_compress(h, m, last ? (uint32_t)total_cols * 4u : t0 + (b + 1) * 64u, last, one);
        if (++s == LEAF_STAGES) { s = 0; ph ^= 1; }
    }
    out[2 * i] = make_uint4(h[0], h[1], h[2], h[3]);
    out[2 * i + 1] = make_uint4(h[4], h[5], h[6], h[7]);
}

// ---- the bottom of a tree in ONE launch: node layers start_log .. 0 (no column values at those sizes), one CTA.  A layer of a
// few hundred nodes is one compression latency of work, but as its own launch it costs ~3 us: 10 such launches per tree, 23 trees
// per FRI commitment of config 4 (FriProver.commit 5.4 -> 5.0 ms, profiles/r02_fri_c4.txt).  Larger layers stay separate launches:
// a single SM hashes 128 nodes per ~0.7 us, so 4096 nodes alone would take 22 us.  `tree` is the layers buffer, root first.
constexpr int TAIL_LOG = 9;    // 512 + 256 + ... + 1 compressions: beyond that one SM is compute-bound (a 4096-node layer alone is 22 us)
__global__ void __launch_bounds__(1024) merkle_tail_kernel(uint4 *__restrict__ tree, int start_log, uint32_t one)
{
    // a layer's hashes also stay in shared memory for the layer above it (ping-pong): the chain of TAIL_LOG + 1 dependent layers
    // then pays one compression + one barrier per layer instead of a round trip through L2 as well
    __shared__ uint4 keep[2][2 << TAIL_LOG];
    for (int log = start_log; log >= 0; log--) {
        const uint4 *prev = log == start_log ? tree + 2ull * ((2ull << log) - 1) : keep[(log + 1) & 1];
        uint4 *dst = tree + 2ull * ((1ull << log) - 1), *mine = keep[log & 1];
        for (uint32_t i = threadIdx.x; i < (1u << log); i += blockDim.x) {
            uint32_t h[8], m[16];
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const uint4 v = prev[4 * i + q];
                m[4 * q] = v.x; m[4 * q + 1] = v.y; m[4 * q + 2] = v.z; m[4 * q + 3] = v.w;
            }
            b2s_init(h);
            b2s_compress(h, m, 64u, true, one);
            const uint4 lo = make_uint4(h[0], h[1], h[2], h[3]), hi = make_uint4(h[4], h[5], h[6], h[7]);
            dst[2 * i] = lo; dst[2 * i + 1] = hi;
            mine[2 * i] = lo; mine[2 * i + 1] = hi;
        }
        __syncthreads();
    }
}

// ---- proof-of-work grind (backend/cpu/grind.ts:31-42): the smallest nonce for which the channel digest after
// mix_u64(nonce) -- BLAKE2s(digest || LE32(lo) || LE32(hi)), one 40-byte block (channel/blake2.ts:120-151) -- has at least
// pow_bits trailing zero bits in its low 128 bits (:94-111).  Every thread tries nonces base + k * stride; a batch covers a
// contiguous range, so the minimum over the first batch with a hit is the nonce the sequential search returns.
__global__ void __launch_bounds__(256) grind_kernel(const uint32_t *__restrict__ digest, uint32_t pow_bits, uint64_t base, int iters,
                                                    unsigned long long *__restrict__ best, uint32_t one)
{
    uint32_t m[16];
#pragma unroll
    for (int k = 0; k < 8; k++) m[k] = digest[k];
#pragma unroll
    for (int k = 10; k < 16; k++) m[k] = 0;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    uint64_t nonce = base + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (int it = 0; it < iters; it++, nonce += stride) {
        uint32_t h[8];
        b2s_init(h);
        m[8] = (uint32_t)nonce; m[9] = (uint32_t)(nonce >> 32);
        b2s_compress(h, m, 40u, true, one);
        const uint32_t tz = h[0] ? __ffs(h[0]) - 1 : h[1] ? 32 + __ffs(h[1]) - 1 : h[2] ? 64 + __ffs(h[2]) - 1 : h[3] ? 96 + __ffs(h[3]) - 1 : 128;
        if (tz >= pow_bits) atomicMin(best, (unsigned long long)nonce);
    }
}

static int g_tail_mode = 1;    // 1: the bottom layers of tstwo_merkle_commit in one launch (merkle_tail_kernel); 0: one launch per layer
static int g_leaf_mode = 1;    // 1: bulk-copy-fed leaf kernel where it applies; 0: merkle_layer_kernel always (tstwo_debug_set_leaf_mode)
static int g_last_leaf = -1;

typedef CUresult (*LeafEncodeFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                 const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static bool leaf_tma_applies(const ColSet &cs, int n_cols, uint64_t n)
{
    return g_leaf_mode != 0 && n_cols >= 1 && !cs.table && !cs.shard_bases && cs.shard_log == 31 && n >= LEAF_ROWS && n < (1ull << 32) &&
           (reinterpret_cast<uintptr_t>(cs.base) & 15) == 0 && (n_cols == 1 || ((cs.stride * 4) % 16 == 0 && cs.stride >= n && cs.stride * 4 < (1ull << 40)));
}

static int merkle_leaf_tma(const ColSet &cs, int n_cols, uint64_t n, uint8_t *out)
{
    LeafEncodeFn enc = reinterpret_cast<LeafEncodeFn>(tma_encode_tiled_fn());
    if (!enc) return TSTWO_ERR_CUDA;
    if (!ctx().attr_leaf) {   // per device
        TS_CUDA(cudaFuncSetAttribute(merkle_leaf_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, LEAF_SMEM));
        TS_CUDA(cudaFuncSetAttribute(merkle_leaf_tma_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
        ctx().attr_leaf = true;
    }
    CUtensorMap map;
    const cuuint64_t dims[2] = { n, (cuuint64_t)n_cols };
    const cuuint64_t strides[1] = { n_cols > 1 ? cs.stride * 4 : n * 4 };
    const cuuint32_t box[2] = { LEAF_ROWS, 16 }, estr[2] = { 1, 1 };
    CUresult r = enc(&map, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, const_cast<uint32_t *>(cs.base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(TSTWO_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d) for %d leaf columns of %llu rows", (int)r, n_cols, (unsigned long long)n);
    TS_LAUNCH(merkle_leaf_tma_kernel, (unsigned)(n / LEAF_ROWS), LEAF_ROWS, LEAF_SMEM, map, n_cols, (uint4 *)out, 1u);
    return 0;
}

// Leaf layer in column chunks (host-fed pipeline: columns arrive over PCIe 16 at a time).  The BLAKE2s chaining value of
// every row lives in the leaf layer itself; a launch absorbs columns [c0, c0 + nc) of a leaf of `total_cols` words
// (c0 a multiple of 16; nc a multiple of 16 unless the chunk is the last) and the chunk holding the last word finalises.
// Same digests as merkle_layer_kernel<false>: the block boundaries and counters are those of the one-shot hash.
__global__ void __launch_bounds__(256) merkle_leaf_absorb_kernel(ColSet cols, int c0, int nc, int total_cols, uint64_t n_rows,
                                                                 uint4 *__restrict__ state, uint32_t one)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_rows) return;
    uint32_t h[8], m[16];
    if (c0 == 0) b2s_init(h);
    else {
        const uint4 a = state[2 * i], b = state[2 * i + 1];
        h[0] = a.x; h[1] = a.y; h[2] = a.z; h[3] = a.w; h[4] = b.x; h[5] = b.y; h[6] = b.z; h[7] = b.w;
    }
    for (int b0 = 0; b0 < nc; b0 += 16) {
#pragma unroll
        for (int k = 0; k < 16; k++) m[k] = (b0 + k < nc) ? __ldcs(cols.col(b0 + k) + i) : 0u;
        const bool last = c0 + b0 + 16 >= total_cols;
        b2s_compress(h, m, last ? (uint32_t)total_cols * 4u : (uint32_t)(c0 + b0 + 16) * 4u, last, one);
    }
    state[2 * i] = make_uint4(h[0], h[1], h[2], h[3]);
    state[2 * i + 1] = make_uint4(h[4], h[5], h[6], h[7]);
}

int merkle_leaf_absorb(int log_size, const ColSet &cols, int c0, int nc, int total_cols, uint8_t *leaf_layer)
{
    const uint64_t n = 1ull << log_size;
    if (c0 % 16 || nc <= 0 || (nc % 16 && c0 + nc != total_cols)) return fail(TSTWO_ERR_BAD_ARG, "leaf chunk [%d, %d) of %d is not block aligned", c0, c0 + nc, total_cols);
    TS_LAUNCH(merkle_leaf_absorb_kernel, (unsigned)((n + 255) / 256), 256, 0, cols, c0, nc, total_cols, n, (uint4 *)leaf_layer, 1u);
    return 0;
}

// The same chunk through the bulk-copy kernel where its tensor map can describe the columns (evenly spaced, 16-byte aligned,
// >= 256 rows); otherwise the per-thread-load kernel above.  Same digests either way.
int merkle_leaf_absorb_fast(int log_size, const ColSet &cols, int c0, int nc, int total_cols, uint8_t *leaf_layer)
{
    const uint64_t n = 1ull << log_size;
    if (c0 < 0 || c0 % 16 || nc <= 0 || c0 + nc > total_cols || (nc % 16 && c0 + nc != total_cols))
        return fail(TSTWO_ERR_BAD_ARG, "leaf chunk [%d, %d) of %d is not block aligned", c0, c0 + nc, total_cols);
    if (!leaf_tma_applies(cols, nc, n)) return merkle_leaf_absorb(log_size, cols, c0, nc, total_cols, leaf_layer);
    LeafEncodeFn enc = reinterpret_cast<LeafEncodeFn>(tma_encode_tiled_fn());
    if (!enc) return TSTWO_ERR_CUDA;
    if (!ctx().attr_leaf_chunk) {   // per device
        TS_CUDA(cudaFuncSetAttribute(merkle_leaf_chunk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, LEAF_SMEM));
        TS_CUDA(cudaFuncSetAttribute(merkle_leaf_chunk_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
        ctx().attr_leaf_chunk = true;
    }
    CUtensorMap map;
    const cuuint64_t dims[2] = { n, (cuuint64_t)nc };
    const cuuint64_t strides[1] = { nc > 1 ? cols.stride * 4 : n * 4 };
    const cuuint32_t box[2] = { LEAF_ROWS, 16 }, estr[2] = { 1, 1 };
    CUresult r = enc(&map, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, const_cast<uint32_t *>(cols.base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(TSTWO_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d) for %d leaf columns of %llu rows", (int)r, nc, (unsigned long long)n);
    TS_LAUNCH(merkle_leaf_chunk_kernel, (unsigned)(n / LEAF_ROWS), LEAF_ROWS, LEAF_SMEM, map, nc, c0, total_cols, (uint4 *)leaf_layer, 1u);
    return 0;
}

int merkle_commit_layer(int log_size, const uint8_t *prev, const ColSet &cols, int n_cols, uint8_t *out)
{
    const uint64_t n = 1ull << log_size;
    const unsigned block = 256;
    const unsigned grid = (unsigned)((n + block - 1) / block);
    g_last_leaf = 0;
    if (!prev && leaf_tma_applies(cols, n_cols, n)) {
        g_last_leaf = g_leaf_mode;
        return merkle_leaf_tma(cols, n_cols, n, out);
    }
    if (prev && n_cols == 0 && g_leaf_mode) TS_LAUNCH(merkle_node_kernel, grid, block, 0, (const uint4 *)prev, n, (uint4 *)out, 1u, 1u);
    else if (prev && n_cols >= 1 && n_cols <= 16 && g_leaf_mode && !cols.table && !cols.shard_bases && cols.shard_log == 31)
        TS_LAUNCH(merkle_node_cols_kernel, grid, block, 0, (const uint4 *)prev, cols, n_cols, n, (uint4 *)out, 2u, 1u);
    else if (prev) TS_LAUNCH((merkle_layer_kernel<true>), grid, block, 0, (const uint4 *)prev, cols, n_cols, n, (uint4 *)out, 1u);
    else TS_LAUNCH((merkle_layer_kernel<false>), grid, block, 0, (const uint4 *)nullptr, cols, n_cols, n, (uint4 *)out, 1u);
    return 0;
}

}  // namespace tstwo

using namespace tstwo;

namespace tstwo {
int merkle_tail(uint8_t *layers, int start_log)
{
    if (!g_tail_mode || start_log < 0) return -1;
    if (start_log > TAIL_LOG) start_log = TAIL_LOG;
    merkle_tail_kernel<<<1, 1024, 0, ctx().stream>>>((uint4 *)layers, start_log, 1u);
    ctx().launches++;
    return cudaGetLastError() == cudaSuccess ? start_log : -1;
}
}  // namespace tstwo

extern "C" {

int tstwo_debug_set_leaf_mode(int mode) { g_leaf_mode = mode & 1; g_tail_mode = !(mode & 2); return 0; }
int tstwo_debug_last_leaf_mode(void) { return g_last_leaf; }

int tstwo_grind(const uint8_t digest[32], uint32_t pow_bits, uint64_t *nonce)
{
    TS_REQUIRE_INIT();
    Span span("Grind");
    if (!digest || !nonce || pow_bits > 64) return fail(TSTWO_ERR_BAD_ARG, "bad grind arguments (pow_bits <= 64)");
    Context &c = ctx();
    uint32_t *d_digest = nullptr;
    unsigned long long *d_best = nullptr, best = ~0ull;
    TS_CUDA(cudaMallocAsync(&d_digest, 32, c.stream));
    if (cudaMallocAsync(&d_best, 8, c.stream) != cudaSuccess) { cudaFreeAsync(d_digest, c.stream); return fail(TSTWO_ERR_CUDA, "grind: cudaMallocAsync failed"); }
    // every exit below releases the two buffers
    auto search = [&]() -> int {
        TS_CUDA(cudaMemcpyAsync(d_digest, digest, 32, cudaMemcpyHostToDevice, c.stream));
        TS_CUDA(cudaMemcpyAsync(d_best, &best, 8, cudaMemcpyHostToDevice, c.stream));
        const unsigned grid = (unsigned)(c.sm_count > 0 ? c.sm_count : 148) * 8;
        uint64_t base = 0;
        // first batch small (most proofs use few bits), then 2^25 nonces per launch (~1.3 ms)
        for (int iters = 1;; iters = 64) {
            TS_LAUNCH(grind_kernel, grid, 256, 0, d_digest, pow_bits, base, iters, d_best, 1u);
            TS_CUDA(cudaMemcpyAsync(&best, d_best, 8, cudaMemcpyDeviceToHost, c.stream));
            TS_CUDA(cudaStreamSynchronize(c.stream));
            if (best != ~0ull) break;
            const uint64_t step = (uint64_t)grid * 256 * iters;
            if (base + step < base) { best = ~0ull; break; }
            base += step;
        }
        return 0;
    };
    const int rc = search();
    cudaFreeAsync(d_digest, c.stream);
    cudaFreeAsync(d_best, c.stream);
    if (rc) return rc;
    if (best == ~0ull) return fail(TSTWO_ERR_BAD_ARG, "no nonce with %u trailing zero bits", pow_bits);
    *nonce = best;
    return 0;
}

int tstwo_merkle_commit_layer(int log_size, tstwo_dptr prev_layer, const tstwo_dptr *cols, int n_cols, tstwo_dptr out_layer)
{
    TS_REQUIRE_INIT();
    if (log_size < 0 || log_size > 31 || n_cols < 0 || !out_layer) return fail(TSTWO_ERR_BAD_ARG, "bad commit_layer arguments");
    ColSet cs;
    int rc = make_colset(cols, n_cols, 0, &cs);
    if (rc) return rc;
    return merkle_commit_layer(log_size, (const uint8_t *)prev_layer, cs, n_cols, (uint8_t *)out_layer);
}

int tstwo_merkle_leaf_absorb(tstwo_dptr cols, uint64_t col_stride_words, int c0, int nc, int total_cols, int log_size, tstwo_dptr layers)
{
    TS_REQUIRE_INIT();
    if (!cols || !layers || log_size < 0 || log_size > 30 || nc < 1 || total_cols < 1 || (cols & 3))
        return fail(TSTWO_ERR_BAD_ARG, "bad leaf_absorb arguments");
    if (nc > 1 && col_stride_words < (1ull << log_size)) return fail(TSTWO_ERR_BAD_ARG, "column stride shorter than a column");
    ColSet cs = { (const uint32_t *)cols, col_stride_words, nullptr };
    cs.aligned16 = !((cols | (col_stride_words * 4)) & 15);
    return merkle_leaf_absorb_fast(log_size, cs, c0, nc, total_cols, (uint8_t *)layers + 32ull * ((1ull << log_size) - 1));
}

int tstwo_merkle_commit_nodes(int log_size, tstwo_dptr layers, uint8_t root[32])
{
    TS_REQUIRE_INIT();
    if (!layers || log_size < 0 || log_size > 30) return fail(TSTWO_ERR_BAD_ARG, "bad commit_nodes arguments");
    uint8_t *base = (uint8_t *)layers;
    const uint8_t *prev = base + 32ull * ((1ull << log_size) - 1);
    ColSet none = { nullptr, 0, nullptr };
    for (int log = log_size - 1; log >= 0; log--) {
        if (log <= TAIL_LOG && merkle_tail(base, log) == log) break;   // the bottom layers in one launch
        uint8_t *dst = base + 32ull * ((1ull << log) - 1);
        int rc = merkle_commit_layer(log, prev, none, 0, dst);
        if (rc) return rc;
        prev = dst;
    }
    if (root) {
        TS_CUDA(cudaMemcpyAsync(root, base, 32, cudaMemcpyDeviceToHost, ctx().stream));
        TS_CUDA(cudaStreamSynchronize(ctx().stream));
    }
    return 0;
}

int tstwo_merkle_commit(const tstwo_dptr *cols, const int *col_logs, int n_cols, tstwo_dptr layers, uint8_t root[32])
{
    TS_REQUIRE_INIT();
    Span span("Merkle");
    if (n_cols < 0 || !layers) return fail(TSTWO_ERR_BAD_ARG, "bad merkle_commit arguments");
    uint8_t *base = (uint8_t *)layers;
    if (n_cols == 0) {
        ColSet none = { nullptr, 0, nullptr };
        int rc = merkle_commit_layer(0, nullptr, none, 0, base);   // vcs/prover.ts:17-19
        if (rc) return rc;
    } else {
        int max_log = 0;
        for (int c = 0; c < n_cols; c++) {
            if (col_logs[c] < 0 || col_logs[c] > 31) return fail(TSTWO_ERR_NOT_POW2, "column length is not a power of two");
            if (col_logs[c] > max_log) max_log = col_logs[c];
        }
        std::vector<tstwo_dptr> layer_cols((size_t)n_cols);
        const uint8_t *prev = nullptr;
        int min_log = 31;
        for (int c = 0; c < n_cols; c++) if (col_logs[c] < min_log) min_log = col_logs[c];
        // layers below the smallest column are pure node layers: the last TAIL_LOG + 1 of them go out as one launch
        const int tail_log = min_log - 1 < TAIL_LOG ? min_log - 1 : TAIL_LOG;
        for (int log = max_log; log >= 0; log--) {
            if (log == tail_log && g_tail_mode) {
                TS_LAUNCH(merkle_tail_kernel, 1, 1024, 0, (uint4 *)base, tail_log, 1u);
                break;
            }
            int k = 0;   // stable order among equal lengths (vcs/prover.ts:21,25)
            for (int c = 0; c < n_cols; c++) if (col_logs[c] == log) layer_cols[k++] = cols[c];
            ColSet cs;
            int rc = make_colset(layer_cols.data(), k, (max_log - log) & 1, &cs);
            if (rc) return rc;
            uint8_t *dst = base + 32ull * ((1ull << log) - 1);
            rc = merkle_commit_layer(log, prev, cs, k, dst);
            if (rc) return rc;
            prev = dst;
        }
    }
    if (root) {
        TS_CUDA(cudaMemcpyAsync(root, base, 32, cudaMemcpyDeviceToHost, ctx().stream));
        TS_CUDA(cudaStreamSynchronize(ctx().stream));
    }
    return 0;
}

}  // extern "C"
