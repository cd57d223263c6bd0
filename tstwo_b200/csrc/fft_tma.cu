// Circle FFT / iFFT passes for sm_100a: persistent CTAs fed by the TMA engine (bulk asynchronous copies).
//
// Reference semantics: CpuCirclePoly.interpolate / evaluate (packages/core/src/backend/cpu/circle.ts:84-207) with the
// butterflies of packages/core/src/fft.ts:12-30; same layer / twiddle indexing as fft.cu.
//
// Why this shape.  A 64 KiB tile costs ~650 SM clocks of integer-pipe time per layer (tools/microbench_bf.cu: 10.15 clk
// per warp-butterfly per SM sub-partition, ALU-pipe bound) and ~6000 clocks of the SM's share of HBM bandwidth for
// its read + write (22 B/clk/SM).  The two only overlap if no warp ever waits on a copy: per-thread copies (LDG/STG,
// 16-byte cp.async) block the in-order warps in the load/store unit whenever DRAM is the momentary bottleneck
// (scripts/fft_trace.py timelines, profiles/r01_g_*).  Here every byte moves by cp.async.bulk:
//
//   CONTIGUOUS pass  tile = 16384 consecutive indices of one column, layers 0..13 (the circle layer included):
//                    16 bulk copies of 4 KiB each way, landing 4 KiB + 16 B apart (the pad rotates the banks);
//   STRIDED pass     tile = 256 rows (index bits [lo, lo+8)) x 64 consecutive indices of one column, layers lo..lo+7:
//                    ONE 3-D tensor-map copy (box 64 x 256 x 1) each way, dense in shared memory.
//
//   * 512 threads = two groups of 256; group g transforms tiles g, g+2, ... of the CTA's list, tile i in buffer i % 3;
//   * a unit is 16 bytes (4 consecutive indices); a thread holds 16 units x 4 lanes and applies 4 unit-level layers
//     per stage (+ the 2 lane layers in the lowest contiguous stage): 3 stages (6+4+4) / 2 stages (4+4) per tile,
//     exchanging in place through the tile buffer under a per-group named barrier;
//   * the last stage's results go back to the buffer and leave as bulk stores; the buffer is refilled (tile i+3, for
//     the other group) at the owner group's next stage boundary, once cp.async.bulk.wait_group.read says the stores
//     have been read out.  Arrival of a tile is an mbarrier transaction count.
//
// TOP stage on the tensor cores (TC = true, strided full passes).  The twiddles of a pass's highest four layers depend
// only on index bits above the tile, so that stage is ONE 16 x 16 matrix over M31 for every vector of a run of tiles:
// tc_stage.cuh applies it exactly as a byte-split u8 GEMM (tcgen05.mma kind::i8, accumulators in TMEM); the 1 / 2^n of the
// last inverse pass is folded into the weights.  The lower stage stays on the integer pipes (its matrix changes with the
// upper row nibble: 32 vectors per matrix and tile, an MMA needs 128).
#include <cuda.h>

#include <cstdlib>
#include <cstring>
#include <vector>

#include "fft_params.cuh"
#include "m31.cuh"
#include "tc_stage.cuh"

namespace tstwo {
namespace {

constexpr int THREADS = 512;
constexpr int CHUNK_BYTES = 4096 + 16;                 // contiguous pass: 256 units + one pad unit
// Per pass type: groups per CTA, tile buffers, bytes per buffer.
//   contiguous: 2 groups x 256 threads, 64 KiB tiles (padded chunks), 3 buffers -- integer-bound, the deep layer count wins
//   strided   : 4 groups x 128 threads, 32 KiB tiles, 6 buffers -- the copies dominate: more, smaller tiles in flight and
//               four independent instruction streams keep both the pipes and HBM busier
//   512-row strided (R9): 2 groups x 256 threads, 64 KiB dense tiles (512 rows x 128 bytes, two tensor boxes), 3 buffers, nine
//               layers in three stages (4 + 4 + 1) -- for transforms that leave exactly nine layers above the contiguous pass
template <bool STRIDED, bool R9 = false> struct Cfg {
    static constexpr int G = (STRIDED && !R9) ? 4 : 2, GS = THREADS / G, NBUF = (STRIDED && !R9) ? 6 : 3;
    static constexpr int TILE_BYTES = GS * 16 * 16;
    static constexpr int BUF_BYTES = STRIDED ? TILE_BYTES : 16 * CHUNK_BYTES;   // 65792 = 514 * 128: buffers stay 128-byte aligned
};
// twiddles of a run of tiles (same high index bits): unit-level heap, level l (0 = lowest) at HEAP_BASE(l); levels
// 0..5 and the T_1 slice carry 4 pad words per 256 units so that lanes 256 units apart hit distinct bank groups
constexpr int HEAP_WORDS = 4096 + 6 * 64;
constexpr int T1_WORDS = 4096 + 64;
constexpr int STRIDED_HEAP_WORDS = 8 * 256;            // up to 8 coset heaps of 255 words
// tensor-core stage (TC).  Strided pass: after the tile barriers come [4 MMA mbarriers | TMEM base slot] (128 B) and
// TC_MAX_W weight tiles of 4 KiB (one per coset heap of a zero-extended top pass).  Contiguous pass: its shared memory is
// full (232064 of 232448 bytes), so heap level 1 (the 1024 twiddles of layer 3, 4352 bytes with its pads) is NOT staged --
// every thread fetches its four entries per tile straight from the tree in L2 -- and the weight tile + barriers sit there.
constexpr int TC_MAX_W = 4;
constexpr int TC_EXTRA_BYTES = 128 + TC_MAX_W * (int)tc::B_BYTES;
constexpr int HEAP_L1_WORD = 1024 + 64 * 4;            // first word of heap level 1 of the contiguous pass (1088 words)
template <bool STRIDED, bool R9 = false> constexpr int smem_bytes(bool tc_stage = false)
{
    return Cfg<STRIDED, R9>::NBUF * Cfg<STRIDED, R9>::BUF_BYTES + (STRIDED ? STRIDED_HEAP_WORDS : HEAP_WORDS + T1_WORDS) * 4 + 128 +
           (tc_stage && STRIDED ? TC_EXTRA_BYTES : 0);
}
static_assert(smem_bytes<true, true>() <= 232448, "exceeds the 227 KiB of shared memory a CTA can opt into");
static_assert(smem_bytes<false>() <= 232448 && smem_bytes<true>(true) <= 232448, "exceeds the 227 KiB of shared memory a CTA can opt into");

struct alignas(64) TmaParams {
    CUtensorMap src_map, dst_map;   // strided pass only
    PassParams p;
    uint32_t n_tiles;               // all tiles of the pass
    uint32_t tiles_per_hi;          // consecutive tile ids sharing one twiddle heap
    uint32_t cosets_log;            // strided pass reading zero-extended coefficients (src_mask): the 2^cosets_log tiles that
                                    // read the same source tile are neighbours in the list (the re-reads hit L2) and all
                                    // their heaps are resident; the tile's coset is then its "hi"
};

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void cp_async16(uint32_t dst, const void *src)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async4(uint32_t dst, const void *src)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    uint32_t done, spins = 0;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(bar), "r"(parity) : "memory");
        if (!done && ++spins > (1u << 22)) __trap();   // a lost tile is a bug: fail loudly instead of hanging the GPU
    } while (!done);
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void bulk_s2g(void *dst, uint32_t src, uint32_t bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tensor_g2s(uint32_t dst, const CUtensorMap *map, uint32_t c0, uint32_t c1, uint32_t c2, uint32_t bar)
{
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                 ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(bar) : "memory");
}
__device__ __forceinline__ void tensor_s2g(const CUtensorMap *map, uint32_t c0, uint32_t c1, uint32_t c2, uint32_t src)
{
    asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.tile.bulk_group [%0, {%1, %2, %3}], [%4];"
                 ::"l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(src) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
template <int GS> __device__ __forceinline__ void group_sync(int g) { asm volatile("bar.sync %0, %1;" ::"r"(g + 1), "n"(GS) : "memory"); }

#ifdef TSTWO_FFT_TRACE
// debug timeline: per (CTA 0, group, tile, point) the SM clock; see scripts/fft_trace.py
__device__ long long *g_trace = nullptr;
__device__ int g_trace_cta = 0;
#define TRACE(pt, dep) do { if (g_trace && blockIdx.x == g_trace_cta && tg == 0 && tr_n < 64) { long long c_; \
    asm volatile("mov.u64 %0, %%clock64;" : "=l"(c_) : "r"(dep) : "memory"); g_trace[(g * 64 + tr_n) * 16 + (pt)] = c_; } } while (0)
// second region: start clock of every tile of CTA 0 (slot 2048 + 1024 g + n); third: per-CTA begin / end globaltimer
#define TRACE_TILE() do { if (g_trace && blockIdx.x == g_trace_cta && tg == 0 && tr_all < 1024) { long long c_; \
    asm volatile("mov.u64 %0, %%clock64;" : "=l"(c_) :: "memory"); g_trace[2048 + 1024 * g + tr_all] = c_; tr_all++; } } while (0)
#define TRACE_CTA(which) do { if (g_trace && threadIdx.x == 0) { long long c_; \
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(c_) :: "memory"); g_trace[4096 + 2 * blockIdx.x + (which)] = c_; } } while (0)
#else
#define TRACE(pt, dep) do { } while (0)
#define TRACE_TILE() do { } while (0)
#define TRACE_CTA(which) do { } while (0)
#endif

// ---- stage geometry.  U = unit index inside the tile: contiguous (12 bits) U = index >> 2; strided (11 bits) U =
// row << 3 | (index >> 2 & 7).  A thread owns the 16 units U0 | d << SU.  In every map the lanes of a quarter warp differ in
// bits that move the physical unit (U + PAD * (U >> 8)) across the 8 bank groups: U[0..2], or U[8..10] where the
// thread owns U[0..3] itself (lowest contiguous stage; that is what the chunk pad is for).
//   contiguous: ST0 = lanes + U[0..3] (layers 0..5), ST1 = U[4..7] (6..9), ST2 = U[8..11] (10..13)
//   strided   : ST0 = U[3..6] (window layers 0..3), ST1 = U[7..10] (4..7)
//   512-row  : U (12 bits) = row << 3 | (index >> 2 & 7); ST0 = U[3..6] (window layers 0..3), ST1 = U[7..10] (4..7), ST2 = U[8..11] of
//              which only U[11] (layer 8) is applied -- rows 5..7 ride along so that a thread still owns 16 units
template <bool STRIDED, int ST, bool R9 = false> struct Stg {
    static constexpr int SU = R9 ? (ST == 2 ? 8 : 3 + 4 * ST) : STRIDED ? 3 + 4 * ST : 4 * ST;
    static constexpr int PAD = STRIDED ? 0 : 1;
    __device__ static __forceinline__ int u0(int tg)
    {
        if (R9) return SU == 3 ? ((tg & 7) | ((tg >> 3) << 7)) : SU == 7 ? ((tg & 127) | ((tg >> 7) << 11)) : tg;
        if (STRIDED) return SU == 3 ? ((tg & 7) | ((tg >> 3) << 7)) : tg;
        if (SU == 0) return ((tg & 7) << 8) | (((tg >> 3) & 15) << 4) | ((tg >> 7) << 11);
        if (SU == 4) return (tg & 15) | ((tg >> 4) << 8);
        return tg;
    }
    __device__ static __forceinline__ constexpr int unit_off(int d) { return (d << SU) + PAD * ((d << SU) >> 8); }
    __device__ static __forceinline__ int phys0(int tg) { const int U0 = u0(tg); return U0 + PAD * (U0 >> 8); }
};

template <bool STRIDED, int ST, bool R9 = false> __device__ __forceinline__ void tile_load(const uint4 *buf, int tg, uint32_t (&v)[16][4])
{
    using S = Stg<STRIDED, ST, R9>;
    const uint4 *b = buf + S::phys0(tg);
#pragma unroll
    for (int d = 0; d < 16; d++) {
        const uint4 a = b[S::unit_off(d)];
        v[d][0] = a.x; v[d][1] = a.y; v[d][2] = a.z; v[d][3] = a.w;
    }
}
template <bool STRIDED, int ST, bool R9 = false> __device__ __forceinline__ void tile_store(uint4 *buf, int tg, const uint32_t (&v)[16][4])
{
    using S = Stg<STRIDED, ST, R9>;
    uint4 *b = buf + S::phys0(tg);
#pragma unroll
    for (int d = 0; d < 16; d++) b[S::unit_off(d)] = make_uint4(v[d][0], v[d][1], v[d][2], v[d][3]);
}

// Heap position of entry q of level l.  NLEV levels (12 contiguous / 8 strided): level l holds 2^(NLEV-1-l) entries at
// 2^(NLEV-1-l) + q; in the contiguous pass levels 0..5 are followed by 64 pad words each and entry q of such a level
// moves up by 4 words per 256 units it lies into the tile (q >> (7 - l) == unit >> 8).
template <bool STRIDED> __device__ __forceinline__ constexpr int heap_base(int l)
{
    return STRIDED ? (1 << (7 - l)) : (1 << (11 - l)) + (l <= 5 ? 64 * (5 - l) : 0);
}

// unit-level layers of one stage: digit bit k <-> heap level l = sl + k; the butterfly (d, d | 1 << k) uses entry
// q = unit >> (l + 1); a thread needs 8 >> k consecutive entries per level.  The stage is a RUN-TIME parameter (U0 =
// the thread's base unit, sl = heap level of digit bit 0): one copy of the butterfly code serves all stages, which
// keeps a kernel's loop body within the instruction cache (fully unrolled per stage it is ~64 KiB and the two groups,
// sitting in different stages, evict each other: ncu stall_no_instruction 2.9 per issue, profiles/r01_k_*).
// l1_global != nullptr (contiguous pass with the tensor-core TOP stage): heap level 1 is not in shared memory; entry q of it is
// l1_global[q] (the tree itself, L2-resident), fetched at the top of the stage that needs it so that the latency hides
// behind the other layers' butterflies.
template <bool STRIDED, bool INV, bool FULL, int SNLEV = 8>
__device__ __forceinline__ void unit_layers(uint32_t (&v)[16][4], const uint32_t *heap, int U0, int sl, int j_min,
                                            const uint32_t *l1_global = nullptr)
{
    const int ub = STRIDED ? (U0 >> 3) : U0;           // unit-level coordinate the heap is indexed by (row / unit)
    const int padw = 4 * (U0 >> 8);
    uint4 l1 = make_uint4(0, 0, 0, 0);
    if (!STRIDED && l1_global && sl == 0) l1 = __ldg(reinterpret_cast<const uint4 *>(l1_global + (ub >> 2)));
#pragma unroll
    for (int kk = 0; kk < 4; kk++) {
        const int k = INV ? kk : 3 - kk;
        const int l = sl + k;
        if (!FULL && l < j_min) continue;              // uniform across the grid
        const int cnt = 8 >> k;
        const int base = STRIDED ? (1 << (SNLEV - 1 - l)) : (1 << (11 - l)) + (l <= 5 ? 64 * (5 - l) + padw : 0);
        const uint32_t *q = heap + base + (ub >> (l + 1));
        uint32_t t[8];
        if (cnt == 8) {
            const uint4 a = *reinterpret_cast<const uint4 *>(q), b = *reinterpret_cast<const uint4 *>(q + 4);
            t[0] = a.x; t[1] = a.y; t[2] = a.z; t[3] = a.w; t[4] = b.x; t[5] = b.y; t[6] = b.z; t[7] = b.w;
        } else if (cnt == 4) {
            const uint4 a = (!STRIDED && l1_global && l == 1) ? l1 : *reinterpret_cast<const uint4 *>(q);
            t[0] = a.x; t[1] = a.y; t[2] = a.z; t[3] = a.w;
        } else if (cnt == 2) {
            const uint2 a = *reinterpret_cast<const uint2 *>(q);
            t[0] = a.x; t[1] = a.y;
        } else t[0] = q[0];
#pragma unroll
        for (int d = 0; d < 16; d++) {
            if (d & (1 << k)) continue;
            const uint32_t tw = t[d >> (k + 1)];
#pragma unroll
            for (int e = 0; e < 4; e++) {
                if (INV) m31::ibutterfly(v[d][e], v[d | (1 << k)][e], tw);
                else m31::butterfly(v[d][e], v[d | (1 << k)][e], tw);
            }
        }
    }
}

// layers 1 and 0 of the contiguous pass act inside a unit (lanes = index bits 0..1); both read the T_1 slice:
// layer 1: (lane 0, lane 2), (lane 1, lane 3) with T_1[unit]; circle layer 0: (0,1), (2,3) with, per unit pair
// (x, y) = T_1[2e], T_1[2e+1]:  unit 2e -> [y, -y], unit 2e+1 -> [-x, x]   (circle.ts:270-278)
template <bool INV> __device__ __forceinline__ void lane_layers(uint32_t (&v)[16][4], const uint32_t *tw1, int tg)
{
    const uint32_t P2 = 0xfffffffeu;
    const int U0 = Stg<false, 0>::u0(tg);
    const uint32_t *q = tw1 + U0 + 4 * (U0 >> 8);
#pragma unroll
    for (int h = 0; h < 4; h++) {                      // 4 units at a time keeps the twiddle registers short-lived
        const uint4 a = *reinterpret_cast<const uint4 *>(q + 4 * h);
        const uint32_t t[4] = { a.x, a.y, a.z, a.w };
#pragma unroll
        for (int pass = 0; pass < 2; pass++) {
            const bool circle = INV ? pass == 0 : pass == 1;
            if (circle) {
#pragma unroll
                for (int e = 0; e < 2; e++) {
                    const uint32_t x = t[2 * e], y = t[2 * e + 1];
                    uint32_t(&a0)[4] = v[4 * h + 2 * e], (&a1)[4] = v[4 * h + 2 * e + 1];
                    if (INV) {
                        m31::ibutterfly(a0[0], a0[1], y); m31::ibutterfly(a0[2], a0[3], P2 - y);
                        m31::ibutterfly(a1[0], a1[1], P2 - x); m31::ibutterfly(a1[2], a1[3], x);
                    } else {
                        m31::butterfly(a0[0], a0[1], y); m31::butterfly(a0[2], a0[3], P2 - y);
                        m31::butterfly(a1[0], a1[1], P2 - x); m31::butterfly(a1[2], a1[3], x);
                    }
                }
            } else {
#pragma unroll
                for (int e = 0; e < 4; e++) {
                    uint32_t(&a)[4] = v[4 * h + e];
                    if (INV) { m31::ibutterfly(a[0], a[2], t[e]); m31::ibutterfly(a[1], a[3], t[e]); }
                    else { m31::butterfly(a[0], a[2], t[e]); m31::butterfly(a[1], a[3], t[e]); }
                }
            }
        }
    }
}

// ---- tile coordinates
struct Tile { uint32_t hi, col, mid; };
template <bool STRIDED> __device__ __forceinline__ Tile decode(const TmaParams &sp, uint32_t t, uint32_t hi)
{
    if (STRIDED && sp.cosets_log) {
        const uint32_t r = t >> sp.cosets_log;
        return { t & ((1u << sp.cosets_log) - 1), r >> (sp.p.lo - 5), r & ((1u << (sp.p.lo - 5)) - 1) };
    }
    const uint32_t rem = t - hi * sp.tiles_per_hi;
    if (STRIDED) return { hi, rem >> (sp.p.lo - 5), rem & ((1u << (sp.p.lo - 5)) - 1) };
    return { hi, rem, 0 };
}

// fetch tile `tl` into the buffer at shared address `buf`; its 64 KiB complete one phase of `bar`
template <bool STRIDED, bool R9 = false>
__device__ __forceinline__ void tile_fetch(const TmaParams &sp, const Tile tl, uint32_t buf, uint32_t bar, int tg)
{
    const PassParams &p = sp.p;
    if (R9) {                                          // 512 rows = two boxes of 256 rows
        if (tg == 0) {
            mbar_arrive_expect_tx(bar, Cfg<true, true>::TILE_BYTES);
            const uint32_t row0 = ((tl.hi << (p.lo + 9)) & p.src_mask) >> p.lo;
            tensor_g2s(buf, &sp.src_map, tl.mid << 5, row0, tl.col, bar);
            tensor_g2s(buf + 32768, &sp.src_map, tl.mid << 5, row0 + 256, tl.col, bar);
        }
    } else if (STRIDED) {
        if (tg == 0) {
            mbar_arrive_expect_tx(bar, Cfg<true>::TILE_BYTES);
            tensor_g2s(buf, &sp.src_map, tl.mid << 5, ((tl.hi << (p.lo + 8)) & p.src_mask) >> p.lo, tl.col, bar);
        }
    } else if (tg < 16) {
        if (tg == 0) mbar_arrive_expect_tx(bar, 65536);
        __syncwarp(0xffffu);
        const uint32_t *src = p.src.col(tl.col) + p.src.off((tl.hi << 14) & p.src_mask) + 1024 * tg;
        bulk_g2s(buf + tg * CHUNK_BYTES, src, 4096, bar);
    }
}
// the transformed tile (complete in the buffer, made visible to the async proxy by the caller) leaves as bulk stores
template <bool STRIDED, bool R9 = false>
__device__ __forceinline__ void tile_drain(const TmaParams &sp, const Tile tl, uint32_t buf, int tg)
{
    const PassParams &p = sp.p;
    if (R9) {
        if (tg == 0) {
            tensor_s2g(&sp.dst_map, tl.mid << 5, tl.hi << 9, tl.col, buf);
            tensor_s2g(&sp.dst_map, tl.mid << 5, (tl.hi << 9) + 256, tl.col, buf + 32768);
            bulk_commit();
        }
    } else if (STRIDED) {
        if (tg == 0) { tensor_s2g(&sp.dst_map, tl.mid << 5, tl.hi << 8, tl.col, buf); bulk_commit(); }
    } else if (tg < 16) {
        uint32_t *dst = p.dst.ptr(tl.col, tl.hi << 14) + 1024 * tg;
        bulk_s2g(dst, buf + tg * CHUNK_BYTES, 4096);
        bulk_commit();
    }
}

// twiddles of the run of tiles with high bits `hi` (all 512 threads; one cp.async group)
template <bool STRIDED, bool SKIP_L1 = false, bool R9 = false>
__device__ __forceinline__ void stage_twiddles(uint32_t *heap, const PassParams &p, uint32_t hi, int tid)
{
    constexpr int NLEV = R9 ? 9 : STRIDED ? 8 : 12;
    const int base_layer = STRIDED ? p.lo : 2;         // tree layer of heap level 0
    for (int e = 4 * tid; e < (1 << NLEV); e += 4 * THREADS) {
        if (e == 0) {
            for (int ee = 1; ee < 4; ee++) {           // the top level (1 entry) and the one below (2 entries)
                const int lg = ee >> 1, l = NLEV - 1 - lg, i = base_layer + l;
                if (STRIDED && l < p.j_min) continue;
                cp_async4(smem_u32(heap + ee), p.tw2 + (p.tree_len - (1u << (p.n - i))) + (hi << lg) + (ee - (1 << lg)));
            }
            continue;
        }
        const int lg = 31 - __clz(e), l = NLEV - 1 - lg, i = base_layer + l, q = e - (1 << lg);
        if (STRIDED && l < p.j_min) continue;
        if (SKIP_L1 && l == 1) continue;
        const int dst = STRIDED ? e : e + (l <= 5 ? 64 * (5 - l) + 4 * (q >> (7 - l)) : 0);
        cp_async16(smem_u32(heap + dst), p.tw2 + (p.tree_len - (1u << (p.n - i))) + (hi << lg) + q);
    }
    if (!STRIDED)
        for (int u = 4 * tid; u < 4096; u += 4 * THREADS)
            cp_async16(smem_u32(heap + HEAP_WORDS + u + 4 * (u >> 8)), p.tw2 + (p.tree_len - (1u << (p.n - 1))) + (hi << 12) + u);
    asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
}

// all stages of a tile, in place in its buffer: load -> butterflies -> store back, group barrier between stages
struct TcTile { tc::Group gr; uint64_t bd0, bd1; int lane; const uint32_t *l1_global; };   // what the tensor-core TOP stage of a tile needs
template <bool STRIDED, bool INV, bool FULL, bool TC, bool R9, typename AtFirstBoundary>
__device__ __forceinline__ void transform_tile(uint4 *buf, const uint32_t *heap, int tg, int g, int j_min, uint32_t scale2,
                                               const TcTile &tct, AtFirstBoundary at_first_boundary)
{
    constexpr int NST = (STRIDED && !R9) ? 2 : 3, TOP = NST - 1;
    constexpr int GS = Cfg<STRIDED, R9>::GS;
#pragma unroll 1
    for (int s = 0; s < NST; s++) {
        const int st = INV ? s : TOP - s;
        if (TC && st == TOP) {
            // tensor-core stage, in place in the buffer: vector (rlow, w) = rows c * 16 + rlow (c = 0..15), word w; batch b of
            // the group takes rows rlow = 4 b + warp, lanes = words.  scale2 is folded into the weights.
            // Contiguous pass: vector w (word of a chunk, 0..1023) = words c * 1028 + w of the padded chunks; the group's two MMA
            // groups take the batches w / 128 = 2 b + (tg >> 7).
            if (STRIDED) tc::stage16_smem<4>(reinterpret_cast<uint32_t *>(buf) + (tg >> 5) * 32 + tct.lane, 512, 128, tct.gr, tct.bd0, tct.bd1, tct.lane);
            else tc::stage16_smem<4>(reinterpret_cast<uint32_t *>(buf) + tg, CHUNK_BYTES / 4, 256, tct.gr, tct.bd0, tct.bd1, tct.lane);
        } else {
            uint32_t v[16][4];
            int U0;
            if (st == 0) { tile_load<STRIDED, 0, R9>(buf, tg, v); U0 = Stg<STRIDED, 0, R9>::u0(tg); }
            else if (st == TOP) { tile_load<STRIDED, TOP, R9>(buf, tg, v); U0 = Stg<STRIDED, TOP, R9>::u0(tg); }
            else { tile_load<STRIDED, 1, R9>(buf, tg, v); U0 = Stg<STRIDED, 1, R9>::u0(tg); }
            if (!STRIDED && INV && st == 0) lane_layers<INV>(v, heap + HEAP_WORDS, tg);
            // 512-row pass: stage 2 owns unit bits 8..11 = rows 5..8, of which rows 5..7 (levels 5..7) were done in stage 1: level 8 only
            if (R9) unit_layers<true, INV, false, 9>(v, heap, U0, st == 2 ? 5 : 4 * st, st == 2 ? 8 : 0);
            else unit_layers<STRIDED, INV, FULL>(v, heap, U0, 4 * st, j_min, TC ? tct.l1_global : nullptr);
            if (!STRIDED && !INV && st == 0) lane_layers<INV>(v, heap + HEAP_WORDS, tg);
            if (INV && s == TOP && scale2) {
#pragma unroll
                for (int d = 0; d < 16; d++)
#pragma unroll
                    for (int e = 0; e < 4; e++) v[d][e] = m31::mul2(v[d][e], scale2);
            }
            if (st == 0) tile_store<STRIDED, 0, R9>(buf, tg, v);
            else if (st == TOP) tile_store<STRIDED, TOP, R9>(buf, tg, v);
            else tile_store<STRIDED, 1, R9>(buf, tg, v);
        }
        if (s == 0) at_first_boundary();
        if (s < TOP) group_sync<GS>(g);
    }
}

template <bool INV, bool STRIDED, bool FULL, bool TC = false, bool R9 = false>
__global__ void __launch_bounds__(THREADS, 1) fft_tma_kernel(const __grid_constant__ TmaParams sp)
{
    static_assert(!TC || FULL, "the tensor-core TOP stage needs all layers of its window");
    static_assert(!R9 || (STRIDED && FULL && !TC), "the 512-row pass is a full strided window on the integer pipes");
    extern __shared__ __align__(1024) uint8_t smem[];
    using C = Cfg<STRIDED, R9>;
    constexpr int NBUF = C::NBUF, BUF_BYTES = C::BUF_BYTES, G = C::G;
    uint32_t *heap = reinterpret_cast<uint32_t *>(smem + NBUF * BUF_BYTES);
    const uint32_t bar0 = smem_u32(heap + (STRIDED ? STRIDED_HEAP_WORDS : HEAP_WORDS + T1_WORDS));
    const uint32_t buf0 = smem_u32(smem);
    const PassParams &p = sp.p;
    const int tid = threadIdx.x, g = tid / C::GS, tg = tid % C::GS;

    // this CTA's tile list: t(k) = start + k * step.  Strided passes interleave CTAs (neighbouring tiles, i.e.
    // neighbouring 256-byte pieces of the same rows, are in flight together); the contiguous pass walks consecutive
    // tiles of its own block.  Either way a twiddle heap serves a long run of tiles.
    uint32_t start, step, count;
    if (STRIDED) {
        start = blockIdx.x; step = gridDim.x;
        count = start < sp.n_tiles ? (sp.n_tiles - start + step - 1) / step : 0;
    } else {
        const uint32_t a = (uint32_t)((uint64_t)sp.n_tiles * blockIdx.x / gridDim.x);
        const uint32_t b = (uint32_t)((uint64_t)sp.n_tiles * (blockIdx.x + 1) / gridDim.x);
        start = a; step = 1; count = b - a;
    }
    if (tid == 0) {
        for (int b = 0; b < 2 * NBUF; b++) mbar_init(bar0 + 8 * b, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    // Buffer b has TWO barriers, used by its even and odd rounds.  Consecutive tiles of a buffer belong to different
    // groups (NBUF is not a multiple of G), and a group that finishes its tiles quickly (the 2-layer partial pass) could
    // otherwise wait for round r + 1 while round r is still in flight -- a parity wait cannot tell those apart and
    // would fall through onto stale data.  Tiles i and i + 2 NBUF always belong to the same group, which takes them in order.
    auto bar_of = [&](uint32_t i) { return bar0 + 8 * (i % NBUF + NBUF * ((i / NBUF) & 1)); };
    // tensor-core stage: TMEM (all 512 columns: 128 per 128-thread MMA group), one mbarrier per group, weight tiles
    // strided: [barriers | slot] (128 B) then the weight tiles, behind the tile barriers; contiguous: weight tile then
    // [barriers | slot] in the place of heap level 1
    uint8_t *tc_weights = STRIDED ? reinterpret_cast<uint8_t *>(heap + STRIDED_HEAP_WORDS) + 128 + 128 : reinterpret_cast<uint8_t *>(heap + HEAP_L1_WORD);
    uint8_t *tc_base = STRIDED ? reinterpret_cast<uint8_t *>(heap + STRIDED_HEAP_WORDS) + 128 : tc_weights + tc::B_BYTES;
    TcTile tct;
    tct.l1_global = nullptr;
    if (TC) {
        if (tid == 0) {
            for (int b = 0; b < 4; b++) mbar_init(smem_u32(tc_base) + 8 * b, 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        if (tid < 32) tc::tmem_alloc(smem_u32(tc_base) + 32, 512);
        tc::fence_before();
        __syncthreads();
        tc::fence_after();
        const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0), mg = warp >> 2;
        tct.gr.tmem = __shfl_sync(0xffffffffu, *reinterpret_cast<volatile uint32_t *>(tc_base + 32), 0) + (uint32_t)mg * 128;
        tct.gr.tmem_lane = tct.gr.tmem + ((uint32_t)((warp & 3) * 32) << 16);
        tct.gr.bar = smem_u32(tc_base) + 8 * mg;
        tct.gr.bar_a = 8 + mg;                         // named barriers 8..11: the 128 threads of an MMA group
        tct.gr.bar_d = 12 + mg;
        tct.gr.issuer_warp = (warp & 3) == 0;
        tct.lane = tid & 31;
        tct.bd0 = tct.bd1 = 0;
    }
    uint32_t par = 0;                                  // bit b + NBUF x: parity of the completed phases of barrier x of buffer b
#ifdef TSTWO_FFT_TRACE
    int tr_n = 0, tr_all = 0;
#endif
    TRACE_CTA(0);

    uint32_t k = 0;
    while (k < count) {
        const uint32_t t0 = start + k * step;
        const uint32_t hi = t0 / sp.tiles_per_hi;
        uint32_t len = ((hi + 1) * sp.tiles_per_hi - t0 + step - 1) / step;   // tiles of this run that are mine
        if (len > count - k) len = count - k;
        __syncthreads();                               // previous run drained: buffers and heap are free
        // the run's first tiles are requested BEFORE its twiddles are staged: both latencies overlap (a run is short when
        // the column set is small -- the column groups of the multi-GPU path, the 16-column chunks of the host-fed path)
#pragma unroll
        for (int i = 0; i < NBUF; i++)
            if ((uint32_t)i < len && (i % G) == g)
                tile_fetch<STRIDED, R9>(sp, decode<STRIDED>(sp, t0 + i * step, hi), buf0 + i * BUF_BYTES, bar_of(i), tg);
        if (STRIDED && sp.cosets_log) {
            for (uint32_t c = 0; c < (1u << sp.cosets_log); c++) stage_twiddles<STRIDED, false, R9>(heap + (R9 ? 512 : 256) * c, p, c, tid);
        } else stage_twiddles<STRIDED, TC && !STRIDED, R9>(heap, p, hi, tid);
        __syncthreads();
        if (TC) {                                      // the run's weight tiles, one per resident heap (ends with a __syncthreads)
            const uint32_t nw = (STRIDED && sp.cosets_log) ? (1u << sp.cosets_log) : 1u;
            for (uint32_t c = 0; c < nw; c++)
                tc::build_weights<INV>(tc_weights + c * tc::B_BYTES, heap + 256 * c, p.scale2, tid, THREADS);
            if (!STRIDED) tct.l1_global = p.tw2 + (p.tree_len - (1u << (p.n - 3))) + (hi << 10);   // tree layer 3 = heap level 1
        }
        uint32_t refill = 0xffffffffu;                 // DEFER: tile of mine whose buffer is still to be refilled

        for (uint32_t i = g; i < len; i += G) {
            const uint32_t b = i % NBUF, round = i / NBUF;
            uint4 *buf = reinterpret_cast<uint4 *>(smem + b * BUF_BYTES);
            const Tile tl = decode<STRIDED>(sp, t0 + i * step, hi);
            TRACE(0, i);
            TRACE_TILE();
            mbar_wait(bar_of(i), ((par >> (b + NBUF * (round & 1))) ^ (round >> 1)) & 1);
            TRACE(1, i);
            // The lanes that issue the stores wait until those have been read out of shared memory and refill the buffer
            // at once (the other seven warps are already on the next tile).  DEFER = refill at the first stage boundary
            // of the group's next tile instead, so that nobody blocks on the read-out.
            constexpr bool DEFER = false;   // measured: deferring costs the inverse contiguous pass 20 % (its first stage is the longest)
            const uint32_t *tile_heap = (STRIDED && sp.cosets_log) ? heap + (R9 ? 512 : 256) * tl.hi : heap;
            if (TC) {
                const uint32_t w = smem_u32(tc_weights) + ((STRIDED && sp.cosets_log) ? tl.hi * tc::B_BYTES : 0u);
                tct.bd0 = tc::make_sdesc(w, tc::B_KCHUNK, tc::B_NGROUP);
                tct.bd1 = tc::make_sdesc(w + 2 * tc::B_KCHUNK, tc::B_KCHUNK, tc::B_NGROUP);
            }
            transform_tile<STRIDED, INV, FULL, TC, R9>(buf, tile_heap, tg, g, p.j_min, p.scale2, tct, [&]() {
                if (DEFER && refill != 0xffffffffu) {
                    if (tg < 16) bulk_wait_read();
                    const uint32_t rb = refill % NBUF;
                    tile_fetch<STRIDED, R9>(sp, decode<STRIDED>(sp, t0 + (refill + NBUF) * step, hi), buf0 + rb * BUF_BYTES, bar_of(refill + NBUF), tg);
                    refill = 0xffffffffu;
                }
            });
            TRACE(2, i);
            fence_proxy_async();                       // generic-proxy stores above -> async-proxy reads below
            group_sync<C::GS>(g);
            tile_drain<STRIDED, R9>(sp, tl, buf0 + b * BUF_BYTES, tg);
            if (i + NBUF < len) {
                if (DEFER) refill = i;
                else {
                    if (tg < 16) bulk_wait_read();
                    tile_fetch<STRIDED, R9>(sp, decode<STRIDED>(sp, t0 + (i + NBUF) * step, hi), buf0 + b * BUF_BYTES, bar_of(i + NBUF), tg);
                }
            }
            TRACE(3, i);
#ifdef TSTWO_FFT_TRACE
            tr_n++;
#endif
        }
        if (tg < 16) bulk_wait_read();                 // the buffers are reused by the next run and must outlive the copies
#pragma unroll
        for (int b = 0; b < NBUF; b++) {
            const uint32_t uses = (len + NBUF - 1 - b) / NBUF;      // rounds 0 .. uses - 1 of buffer b in this run
            par ^= ((((uses + 1) >> 1) & 1) << b) ^ (((uses >> 1) & 1) << (b + NBUF));
        }
        k += len;
    }
    if (TC) {
        tc::fence_before();
        __syncthreads();
        if (tid < 32) tc::tmem_dealloc(*reinterpret_cast<volatile uint32_t *>(tc_base + 32), 512);
    }
    TRACE_CTA(1);
}

template <bool INV, bool STRIDED, bool FULL, bool TC = false, bool R9 = false> cudaError_t set_attr()
{
    return cudaFuncSetAttribute(fft_tma_kernel<INV, STRIDED, FULL, TC, R9>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes<STRIDED, R9>(TC));
}
// tstwo_debug_set_fft_tc: bit 0 = forward strided passes, bit 1 = contiguous passes, bit 2 = inverse strided passes.
// Default = forward strided only, the one pass where it wins (scripts/tc_ab.py, profiles/r02_tc_ab.txt: 2.98 -> 2.52 ms;
// the inverse strided pass is already at 80 % of the HBM peak and loses 0.09 ms to the MMA round trips, the contiguous
// passes with two 256-thread groups cannot hide them: +0.0 / +0.07 ms)
int g_use_tc = 1;
// tstwo_debug_pass_timing: CUDA events around every pass launch, so that bench.py reports each launch's duration live
struct PassStamp { int tag; cudaEvent_t e0, e1; };
bool g_timing = false;
std::vector<PassStamp> g_stamps;
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn g_encode = nullptr;

// 3-D view of a column set for the strided pass: (index & (2^lo - 1), index >> lo, column), box 32 x 256 x 1
int make_map(CUtensorMap *map, const ColSet &cs, int n_cols, int log_len, int lo)
{
    const cuuint64_t dims[3] = { 1ull << lo, 1ull << (log_len - lo), (cuuint64_t)n_cols };
    const cuuint64_t strides[2] = { 4ull << lo, n_cols > 1 ? cs.stride * 4 : (4ull << log_len) };
    const cuuint32_t box[3] = { 32, 256, 1 }, estr[3] = { 1, 1, 1 };
    CUresult r = g_encode(map, CU_TENSOR_MAP_DATA_TYPE_UINT32, 3, const_cast<uint32_t *>(cs.base), dims, strides, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(TSTWO_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d) for lo=%d len=2^%d cols=%d", (int)r, lo, log_len, n_cols);
    return 0;
}

}  // namespace

#ifdef TSTWO_FFT_TRACE
int tma_trace_set(long long *buf)
{
    TS_CUDA(cudaMemcpyToSymbol(g_trace, &buf, sizeof buf));
    return 0;
}
int tma_trace_cta(int cta)
{
    TS_CUDA(cudaMemcpyToSymbol(g_trace_cta, &cta, sizeof cta));
    return 0;
}
#endif

int tma_pass_init()
{
    if (ctx().attr_tma) return 0;      // per device: cudaFuncSetAttribute applies to the current device
    TS_CUDA((set_attr<false, false, true>())); TS_CUDA((set_attr<true, false, true>()));
    TS_CUDA((set_attr<false, true, true>()));  TS_CUDA((set_attr<true, true, true>()));
    TS_CUDA((set_attr<false, true, false>())); TS_CUDA((set_attr<true, true, false>()));
    TS_CUDA((set_attr<false, true, true, true>())); TS_CUDA((set_attr<true, true, true, true>()));
    TS_CUDA((set_attr<false, false, true, true>())); TS_CUDA((set_attr<true, false, true, true>()));
    TS_CUDA((set_attr<false, true, true, false, true>())); TS_CUDA((set_attr<true, true, true, false, true>()));
    void *fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    TS_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
    if (!fn || qres != cudaDriverEntryPointSuccess) return fail(TSTWO_ERR_CUDA, "driver has no cuTensorMapEncodeTiled");
    g_encode = reinterpret_cast<EncodeTiledFn>(fn);
    if (const char *e = getenv("TSTWO_FFT_TC")) g_use_tc = atoi(e);   // A/B without a rebuild (see tstwo_debug_set_fft_tc)
    ctx().attr_tma = true;
    return 0;
}

void *tma_encode_tiled_fn() { return tma_pass_init() ? nullptr : reinterpret_cast<void *>(g_encode); }

// What a pass needs of its column sets: 16-byte aligned columns; the strided pass, whose tensor maps describe a whole
// column set, also uniformly spaced columns (the contiguous pass copies column by column and takes any table).
bool tma_pass_supported(const PassParams &p)
{
    if (!p.src.aligned16 || !p.dst.aligned16 || p.src.shard_bases) return false;
    if (p.dst.shard_bases && p.lo != 0) return false;
    if (p.lo == 0) return p.src.shard_log >= TMA_W0 && p.dst.shard_log >= TMA_W0;
    if (p.src.table || p.dst.table || p.src.shard_log != 31 || p.dst.shard_log != 31) return false;
    if ((reinterpret_cast<uintptr_t>(p.src.base) | reinterpret_cast<uintptr_t>(p.dst.base)) & 15) return false;
    if (p.n_cols > 1 && ((p.src.stride | p.dst.stride) & 3)) return false;
    return true;
}

int tma_pass(bool inverse, const PassParams &p)
{
    int rc = tma_pass_init();
    if (rc) return rc;
    TmaParams sp;
    memset(&sp, 0, sizeof sp);
    sp.p = p;
    const bool strided = p.lo != 0;
    const bool rows9 = strided && p.win == TMA_W9;
    const int width = rows9 ? TMA_W9 : TMA_W1;
    uint64_t n_hi, per_hi;
    if (strided) {
        if (p.src.shard_log != 31 || p.dst.shard_log != 31) return fail(TSTWO_ERR_BAD_ARG, "strided FFT pass on a row-sharded column set");
        if (p.lo < 5 || p.lo + width > p.n || (rows9 && p.j_min != 0)) return fail(TSTWO_ERR_BAD_ARG, "bad strided window lo=%d n=%d", p.lo, p.n);
        int src_log = 0;
        while ((1ull << src_log) <= p.src_mask) src_log++;          // src holds 2^src_log words per column
        if ((rc = make_map(&sp.src_map, p.src, p.n_cols, src_log, p.lo))) return rc;
        if ((rc = make_map(&sp.dst_map, p.dst, p.n_cols, p.n, p.lo))) return rc;
        if (rows9 && src_log < p.lo + TMA_W9) return fail(TSTWO_ERR_BAD_ARG, "512-row window above the source (lo=%d, source 2^%d)", p.lo, src_log);
        n_hi = 1ull << (p.n - p.lo - width);
        per_hi = ((uint64_t)p.n_cols) << (p.lo - 5);
        if (src_log < p.n && p.lo + width == src_log && p.n - src_log <= (rows9 ? 2 : 3)) {   // top window of a blown-up evaluation
            sp.cosets_log = (uint32_t)(p.n - src_log);
            per_hi *= n_hi;                                                      // one run: every coset's heap is staged
            n_hi = 1;
        }
    } else {
        n_hi = 1ull << (p.n - TMA_W0);
        per_hi = (uint64_t)p.n_cols;
    }
    if (n_hi * per_hi >= (1ull << 31)) return fail(TSTWO_ERR_BAD_ARG, "FFT pass too large (%llu tiles)", (unsigned long long)(n_hi * per_hi));
    sp.n_tiles = (uint32_t)(n_hi * per_hi);
    sp.tiles_per_hi = (uint32_t)per_hi;
    unsigned grid = (unsigned)ctx().sm_count;
    if (ctx().grid_limit > 0 && grid > (unsigned)ctx().grid_limit) grid = (unsigned)ctx().grid_limit;
    if (grid > sp.n_tiles) grid = sp.n_tiles;
    const bool use_tc = rows9 ? false : strided ? ((g_use_tc & (inverse ? 4 : 1)) && p.j_min == 0 && (1u << sp.cosets_log) <= (unsigned)TC_MAX_W) : (g_use_tc & 2) != 0;
    PassStamp st{ (inverse ? 1 : 0) | (strided ? 2 : 0) | (use_tc ? 4 : 0) | (rows9 ? 8 : 0), nullptr, nullptr };
    if (g_timing) {
        TS_CUDA(cudaEventCreate(&st.e0)); TS_CUDA(cudaEventCreate(&st.e1));
        TS_CUDA(cudaEventRecord(st.e0, ctx().stream));
    }
    if (rows9) {
        if (inverse) TS_LAUNCH((fft_tma_kernel<true, true, true, false, true>), grid, THREADS, (smem_bytes<true, true>()), sp);
        else TS_LAUNCH((fft_tma_kernel<false, true, true, false, true>), grid, THREADS, (smem_bytes<true, true>()), sp);
    } else if (!strided && use_tc) {
        if (inverse) TS_LAUNCH((fft_tma_kernel<true, false, true, true>), grid, THREADS, smem_bytes<false>(true), sp);
        else TS_LAUNCH((fft_tma_kernel<false, false, true, true>), grid, THREADS, smem_bytes<false>(true), sp);
    } else if (!strided) {
        if (inverse) TS_LAUNCH((fft_tma_kernel<true, false, true>), grid, THREADS, smem_bytes<false>(), sp);
        else TS_LAUNCH((fft_tma_kernel<false, false, true>), grid, THREADS, smem_bytes<false>(), sp);
    } else if (use_tc) {
        if (inverse) TS_LAUNCH((fft_tma_kernel<true, true, true, true>), grid, THREADS, smem_bytes<true>(true), sp);
        else TS_LAUNCH((fft_tma_kernel<false, true, true, true>), grid, THREADS, smem_bytes<true>(true), sp);
    } else if (p.j_min == 0) {
        if (inverse) TS_LAUNCH((fft_tma_kernel<true, true, true>), grid, THREADS, smem_bytes<true>(), sp);
        else TS_LAUNCH((fft_tma_kernel<false, true, true>), grid, THREADS, smem_bytes<true>(), sp);
    } else {
        if (inverse) TS_LAUNCH((fft_tma_kernel<true, true, false>), grid, THREADS, smem_bytes<true>(), sp);
        else TS_LAUNCH((fft_tma_kernel<false, true, false>), grid, THREADS, smem_bytes<true>(), sp);
    }
    if (g_timing) {
        TS_CUDA(cudaEventRecord(st.e1, ctx().stream));
        g_stamps.push_back(st);
    }
    return 0;
}

}  // namespace tstwo

extern "C" int tstwo_debug_set_fft_tc(int mask) { tstwo::g_use_tc = mask; return 0; }
extern "C" int tstwo_debug_pass_timing(int on)
{
    for (auto &s : tstwo::g_stamps) { cudaEventDestroy(s.e0); cudaEventDestroy(s.e1); }
    tstwo::g_stamps.clear();
    tstwo::g_timing = on != 0;
    return 0;
}
extern "C" int tstwo_debug_pass_times(int max, int *tags, float *ms)
{
    int n = 0;
    for (auto &s : tstwo::g_stamps) {
        if (n >= max) break;
        if (cudaEventSynchronize(s.e1) != cudaSuccess || cudaEventElapsedTime(&ms[n], s.e0, s.e1) != cudaSuccess)
            return tstwo::fail(TSTWO_ERR_CUDA, "pass timing events failed");
        tags[n++] = s.tag;
    }
    return n;
}
#ifdef TSTWO_FFT_TRACE
extern "C" int tstwo_debug_fft_trace(tstwo_dptr buf) { return tstwo::tma_trace_set((long long *)buf); }
extern "C" int tstwo_debug_fft_trace_cta(int cta) { return tstwo::tma_trace_cta(cta); }
#endif
