// Circle FFT / iFFT passes as a persistent, software-pipelined kernel for sm_100a (one 512-thread CTA per SM).
//
// Same pass semantics as fft.cu (reference: CpuCirclePoly.interpolate / evaluate, packages/core/src/backend/cpu/
// circle.ts:84-207; butterflies of packages/core/src/fft.ts:12-30), different execution model.  fft.cu runs one CTA
// per tile: load -> transform -> store, so all CTAs of a wave sit in the same phase and the integer pipes idle while
// the loads are in flight (ncu: issue slots ~50 % busy, profiles/r01_c_*).  Here a CTA walks a list of tiles:
//
//   * a tile is 4096 units of 16 bytes (64 KiB): CONTIGUOUS pass = 4 columns x 4096 consecutive indices (a unit is 4
//     consecutive indices of one column, layers 0..11, the circle layer included), STRIDED pass = 1024 rows (index
//     bits [lo, lo+10)) x 16 consecutive indices of one column (layers lo..lo+9);
//   * three tile buffers live in shared memory.  Tiles are fetched with 16-byte cp.async straight into the (padded)
//     exchange layout and their arrival is tracked by one mbarrier per buffer (cp.async.mbarrier.arrive.noinc);
//   * the CTA is two groups of 256 threads.  Group g transforms tiles g, g+2, ... of the list; tile i lives in buffer
//     i % 3.  When a group has read its last stage out of a buffer it immediately issues the loads of tile i+3 (which
//     the OTHER group will transform), so a fetch always has a whole tile time of lead and the two groups run half a
//     tile out of phase: one is in butterflies while the other exchanges / stores;
//   * a thread holds 16 units x 4 lanes in registers and applies 4 (+2 lane) layers per stage, three stages per tile,
//     exchanging in place through the tile buffer (named barriers per group); results leave from registers.
//
// Shared-memory unit index U (12 bits): contiguous U = column << 10 | (index >> 2); strided U = row << 2 | (index>>2&3);
// physical unit = U + (U >> 4) (one pad unit per 16): every quarter warp of every access below touches 8 distinct
// 16-byte bank groups.  Twiddles of a tile depend only on its high index bits ("hi"): they are staged once per run of
// tiles with the same hi (unit-level heap of 1023 words; the contiguous pass adds the T_1 slice for its lane layers).
#include "fft_params.cuh"
#include "m31.cuh"

namespace tstwo {
namespace {

constexpr int THREADS = 512, GROUP = 256, NBUF = 3;
constexpr int UNITS = 4096;
constexpr int BUF_UNITS = UNITS + UNITS / 16;
constexpr int TW_WORDS = 1024 + 1280;   // [0,1024): unit-level heap; then the T_1 slice of the tile (contiguous pass),
                                        // 16 words per 20 so that threads 80 bytes apart hit distinct bank groups
constexpr int SMEM_BYTES = NBUF * BUF_UNITS * 16 + TW_WORDS * 4 + 64;

struct StreamParams {
    PassParams p;
    uint32_t n_tiles;        // all tiles of the pass
    uint32_t tiles_per_hi;   // consecutive tile ids sharing one twiddle heap
};

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void cp_async16(uint32_t dst, const void *src)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async16_zfill(uint32_t dst, const void *src, uint32_t bytes)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void cp_async4(uint32_t dst, const void *src)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
// the barrier receives one arrival from this thread once all its earlier cp.async have landed
__device__ __forceinline__ void cp_async_arrive(uint32_t bar)
{
    asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
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
__device__ __forceinline__ void group_sync(int g) { asm volatile("bar.sync %0, 256;" ::"r"(g + 1) : "memory"); }

// ---- stage geometry.  ST: 0 = low, 1 = mid, 2 = top.  SU = position of the 4 digit bits in U, SL = unit-level
// (heap) level of digit bit 0, layers are applied on digit bits [K0, K0+NL).
template <bool STRIDED, int ST> struct Stg {
    static constexpr int SU = STRIDED ? (ST == 0 ? 2 : ST == 1 ? 4 : 8) : (ST == 0 ? 0 : ST == 1 ? 4 : 6);
    static constexpr int SL = STRIDED ? SU - 2 : SU;
    static constexpr int K0 = (!STRIDED && ST == 2) ? 2 : 0;
    static constexpr int NL = STRIDED ? (ST == 0 ? 2 : 4) : (ST == 2 ? 2 : 4);
    __device__ static __forceinline__ int u0(int tg)
    {
        if (STRIDED) {
            const int g = tg & 3, rr = tg >> 2;
            return ST == 0 ? (g | (rr << 6)) : ST == 1 ? (g | ((rr & 3) << 2) | ((rr >> 2) << 8)) : tg;
        }
        const int c = tg >> 6, m = tg & 63;
        return ST == 0 ? (tg << 4) : ST == 1 ? ((c << 10) | (m & 15) | ((m >> 4) << 8)) : ((c << 10) | m);
    }
    __device__ static __forceinline__ constexpr int unit_off(int d) { return (d << SU) + ((d << SU) >> 4); }
};

template <bool STRIDED, int ST> __device__ __forceinline__ void tile_load(const uint4 *buf, int tg, uint32_t (&v)[16][4])
{
    using S = Stg<STRIDED, ST>;
    const int U0 = S::u0(tg);
    const uint4 *b = buf + U0 + (U0 >> 4);
#pragma unroll
    for (int d = 0; d < 16; d++) {
        const uint4 a = b[S::unit_off(d)];
        v[d][0] = a.x; v[d][1] = a.y; v[d][2] = a.z; v[d][3] = a.w;
    }
}
template <bool STRIDED, int ST> __device__ __forceinline__ void tile_store(uint4 *buf, int tg, const uint32_t (&v)[16][4])
{
    using S = Stg<STRIDED, ST>;
    const int U0 = S::u0(tg);
    uint4 *b = buf + U0 + (U0 >> 4);
#pragma unroll
    for (int d = 0; d < 16; d++) b[S::unit_off(d)] = make_uint4(v[d][0], v[d][1], v[d][2], v[d][3]);
}

// unit-level layers of one stage: digit bit k <-> heap level l = SL + k; the butterfly (d, d | 1<<k) uses
// heap[2^(9-l) + (unit >> (l+1))]; a thread needs 8 >> k consecutive heap words per level.
template <bool STRIDED, int ST, bool INV, bool FULL>
__device__ __forceinline__ void unit_layers(uint32_t (&v)[16][4], const uint32_t *tws, int tg, int j_min)
{
    using S = Stg<STRIDED, ST>;
    const int U0 = S::u0(tg);
    const int ub = STRIDED ? (U0 >> 2) : (U0 & 1023);
#pragma unroll
    for (int kk = 0; kk < S::NL; kk++) {
        const int k = INV ? S::K0 + kk : S::K0 + S::NL - 1 - kk;
        const int l = S::SL + k;
        if (!FULL && l < j_min) continue;              // uniform across the grid
        const int cnt = 8 >> k;
        const uint32_t *q = tws + (1 << (9 - l)) + (ub >> (l + 1));
        uint32_t t[8];
        if (cnt == 8) {
            const uint4 a = *reinterpret_cast<const uint4 *>(q), b = *reinterpret_cast<const uint4 *>(q + 4);
            t[0] = a.x; t[1] = a.y; t[2] = a.z; t[3] = a.w; t[4] = b.x; t[5] = b.y; t[6] = b.z; t[7] = b.w;
        } else if (cnt == 4) {
            const uint4 a = *reinterpret_cast<const uint4 *>(q);
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
    const uint32_t *q = tw1 + 20 * (tg & 63);
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

template <bool STRIDED, int ST, bool INV, bool FULL>
__device__ __forceinline__ void stage_math(uint32_t (&v)[16][4], const uint32_t *tws, int tg, int j_min)
{
    if (!STRIDED && ST == 0 && INV) lane_layers<INV>(v, tws + 1024, tg);
    unit_layers<STRIDED, ST, INV, FULL>(v, tws, tg, j_min);
    if (!STRIDED && ST == 0 && !INV) lane_layers<INV>(v, tws + 1024, tg);
}

// ---- tile coordinates
struct Tile { uint32_t hi, a, b; };   // contiguous: a = first column; strided: a = column, b = mid
template <bool STRIDED> __device__ __forceinline__ Tile decode(const StreamParams &sp, uint32_t t, uint32_t hi)
{
    const uint32_t rem = t - hi * sp.tiles_per_hi;
    if (STRIDED) return { hi, rem >> (sp.p.lo - 4), rem & ((1u << (sp.p.lo - 4)) - 1) };
    return { hi, rem * 4, 0 };
}

// all 256 threads of a group: fetch tile `tl` into buffer `buf` (shared-space address), arrival on `bar`
template <bool STRIDED>
__device__ __forceinline__ void tile_fetch(const PassParams &p, const Tile tl, uint32_t buf, uint32_t bar, int tg)
{
    const uint32_t dst0 = buf + 16u * (tg + (tg >> 4));
    if (STRIDED) {
        const uint32_t base = (((tl.hi << (p.lo + 10)) | (tl.b << 4)) & p.src_mask) | ((uint32_t)(tg >> 2) << p.lo) | ((tg & 3) << 2);
        const uint4 *ptr = reinterpret_cast<const uint4 *>(p.src.col(tl.a) + base);
        const size_t st = (size_t)16 << p.lo;          // 64 rows further, in uint4 units
#pragma unroll
        for (int i = 0; i < 16; i++) cp_async16(dst0 + 16u * (i * 272), ptr + i * st);
    } else {
        const uint64_t off = p.src.off((tl.hi << 12) & p.src_mask) + 4u * tg;
#pragma unroll
        for (int c = 0; c < 4; c++) {
            const bool ok = (int)(tl.a + c) < p.n_cols;
            const uint4 *ptr = reinterpret_cast<const uint4 *>(p.src.col(ok ? tl.a + c : tl.a) + off);
#pragma unroll
            for (int i = 0; i < 4; i++) cp_async16_zfill(dst0 + 16u * ((4 * c + i) * 272), ptr + 256 * i, ok ? 16u : 0u);
        }
    }
    cp_async_arrive(bar);
}

// results leave from the registers of the last stage
template <bool STRIDED, int ST>
__device__ __forceinline__ void tile_write(const PassParams &p, const Tile tl, int tg, const uint32_t (&v)[16][4])
{
    using S = Stg<STRIDED, ST>;
    const int U0 = S::u0(tg);
    if (STRIDED) {
        const uint32_t base = (tl.hi << (p.lo + 10)) | (tl.b << 4) | ((uint32_t)(U0 >> 2) << p.lo) | ((U0 & 3) << 2);
        uint4 *ptr = reinterpret_cast<uint4 *>(p.dst.col(tl.a) + base);
        const size_t st = (size_t)1 << (p.lo + S::SU - 4);
#pragma unroll
        for (int d = 0; d < 16; d++) __stcg(ptr + d * st, make_uint4(v[d][0], v[d][1], v[d][2], v[d][3]));
    } else {
        const int c = U0 >> 10;
        if ((int)(tl.a + c) >= p.n_cols) return;
        uint4 *ptr = reinterpret_cast<uint4 *>(p.dst.col(tl.a + c) + p.dst.off(tl.hi << 12)) + (U0 & 1023);
#pragma unroll
        for (int d = 0; d < 16; d++) __stcg(ptr + (d << S::SU), make_uint4(v[d][0], v[d][1], v[d][2], v[d][3]));
    }
}

// Forward contiguous pass: the last stage leaves a thread with 16 consecutive units (256 bytes), i.e. a warp store
// would touch 32 lines for 512 bytes.  The results go back to the tile buffer instead and leave in tile order:
// thread tg writes units i*256 + tg, a warp 512 contiguous bytes.
__device__ __forceinline__ void tile_write_linear(const PassParams &p, const Tile tl, const uint4 *buf, int tg)
{
    const uint4 *b = buf + tg + (tg >> 4);
    const uint64_t off = p.dst.off(tl.hi << 12) + 4u * tg;
#pragma unroll
    for (int c = 0; c < 4; c++) {
        if ((int)(tl.a + c) >= p.n_cols) break;
        uint4 *ptr = reinterpret_cast<uint4 *>(p.dst.col(tl.a + c) + off);
#pragma unroll
        for (int i = 0; i < 4; i++) __stcg(ptr + 256 * i, b[(4 * c + i) * 272]);
    }
}

// twiddles of the run of tiles with high bits `hi` (all 512 threads; one cp.async group)
template <bool STRIDED> __device__ __forceinline__ void stage_twiddles(uint32_t *tws, const PassParams &p, uint32_t hi, int tid)
{
    const int base = STRIDED ? p.lo : 2;               // tree layer of heap level 0
    if (tid < 256) {
        const int e = 4 * tid;
        if (e == 0) {
            for (int ee = 1; ee < 4; ee++) {           // heap[1] = level 9, heap[2..3] = level 8
                const int lg = ee >> 1, l = 9 - lg, i = base + l;
                if (STRIDED && l < p.j_min) continue;
                cp_async4(smem_u32(tws + ee), p.tw2 + (p.tree_len - (1u << (p.n - i))) + (hi << lg) + (ee - (1 << lg)));
            }
        } else {
            const int lg = 31 - __clz(e), l = 9 - lg, i = base + l;
            if (!(STRIDED && l < p.j_min))
                cp_async16(smem_u32(tws + e), p.tw2 + (p.tree_len - (1u << (p.n - i))) + (hi << lg) + (e - (1 << lg)));
        }
    } else if (!STRIDED) {
        const int e = 4 * (tid - 256);
        cp_async16(smem_u32(tws + 1024 + e + (e >> 4) * 4), p.tw2 + (p.tree_len - (1u << (p.n - 1))) + (hi << 10) + e);
    }
    asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
}

template <bool INV, bool STRIDED, bool FULL>
__global__ void __launch_bounds__(THREADS, 1) fft_stream_kernel(const StreamParams sp)
{
    extern __shared__ uint4 sm[];
    uint32_t *tws = reinterpret_cast<uint32_t *>(sm + NBUF * BUF_UNITS);
    const uint32_t bar0 = smem_u32(tws + TW_WORDS);
    const PassParams &p = sp.p;
    const int tid = threadIdx.x, g = tid >> 8, tg = tid & 255;

    // this CTA's tile list: t(k) = start + k * step.  Strided passes interleave CTAs (neighbouring tiles, i.e.
    // neighbouring 64-byte pieces of the same rows, are in flight together); the contiguous pass takes a block of
    // consecutive tiles so that a twiddle heap serves a long run.
    uint32_t start, step, count;
    if (STRIDED) {
        start = blockIdx.x; step = gridDim.x;
        count = start < sp.n_tiles ? (sp.n_tiles - start + step - 1) / step : 0;
    } else {
        const uint32_t a = (uint32_t)((uint64_t)sp.n_tiles * blockIdx.x / gridDim.x);
        const uint32_t b = (uint32_t)((uint64_t)sp.n_tiles * (blockIdx.x + 1) / gridDim.x);
        start = a; step = 1; count = b - a;
    }
    if (tid == 0)
        for (int b = 0; b < 2 * NBUF; b++) mbar_init(bar0 + 8 * b, GROUP);
    // Two barriers per buffer, for its even and its odd rounds: consecutive tiles of a buffer belong to different groups,
    // and a group that is quick (the partial pass) could otherwise wait for round r + 1 while round r is still in flight,
    // which a parity wait cannot tell apart.  Tiles i and i + 6 always belong to the same group.
    auto bar_of = [&](uint32_t i) { return bar0 + 8 * (i % 3 + 3 * ((i / 3) & 1)); };
    uint32_t par = 0;                                  // bit b + 3 x: parity of the completed phases of barrier x of buffer b

    uint32_t k = 0;
    while (k < count) {
        const uint32_t t0 = start + k * step;
        const uint32_t hi = t0 / sp.tiles_per_hi;
        uint32_t len = ((hi + 1) * sp.tiles_per_hi - t0 + step - 1) / step;   // tiles of this run that are mine
        if (len > count - k) len = count - k;
        __syncthreads();                               // previous run drained: buffers and heap are free
        stage_twiddles<STRIDED>(tws, p, hi, tid);
        __syncthreads();
#pragma unroll
        for (int i = 0; i < NBUF; i++)
            if ((uint32_t)i < len && (i & 1) == g)
                tile_fetch<STRIDED>(p, decode<STRIDED>(sp, t0 + i * step, hi), smem_u32(sm + i * BUF_UNITS), bar_of(i), tg);

        for (uint32_t i = g; i < len; i += 2) {
            const uint32_t b = i % 3, round = i / 3;
            uint4 *buf = sm + b * BUF_UNITS;
            const Tile tl = decode<STRIDED>(sp, t0 + i * step, hi);
            mbar_wait(bar_of(i), ((par >> (b + 3 * (round & 1))) ^ (round >> 1)) & 1);
            uint32_t v[16][4];
            constexpr int FIRST = INV ? 0 : 2, LAST = INV ? 2 : 0;
            tile_load<STRIDED, FIRST>(buf, tg, v);
            stage_math<STRIDED, FIRST, INV, FULL>(v, tws, tg, p.j_min);
            tile_store<STRIDED, FIRST>(buf, tg, v);
            group_sync(g);
            tile_load<STRIDED, 1>(buf, tg, v);
            stage_math<STRIDED, 1, INV, FULL>(v, tws, tg, p.j_min);
            tile_store<STRIDED, 1>(buf, tg, v);
            group_sync(g);
            tile_load<STRIDED, LAST>(buf, tg, v);
            constexpr bool LINEAR_OUT = !STRIDED && !INV;
            if (!LINEAR_OUT) {
                group_sync(g);                         // the whole group has read: the buffer can be refilled
                if (i + 3 < len)
                    tile_fetch<STRIDED>(p, decode<STRIDED>(sp, t0 + (i + 3) * step, hi), smem_u32(buf), bar_of(i + 3), tg);
            }
            stage_math<STRIDED, LAST, INV, FULL>(v, tws, tg, p.j_min);
            if (INV && p.scale2) {
#pragma unroll
                for (int d = 0; d < 16; d++)
#pragma unroll
                    for (int e = 0; e < 4; e++) v[d][e] = m31::mul2(v[d][e], p.scale2);
            }
            if (LINEAR_OUT) {
                tile_store<STRIDED, LAST>(buf, tg, v);
                group_sync(g);
                tile_write_linear(p, tl, buf, tg);
                group_sync(g);
                if (i + 3 < len)
                    tile_fetch<STRIDED>(p, decode<STRIDED>(sp, t0 + (i + 3) * step, hi), smem_u32(buf), bar_of(i + 3), tg);
            } else tile_write<STRIDED, LAST>(p, tl, tg, v);
        }
#pragma unroll
        for (int b = 0; b < 3; b++) {
            const uint32_t uses = (len + 2 - b) / 3;               // rounds 0 .. uses - 1 of buffer b in this run
            par ^= ((((uses + 1) >> 1) & 1) << b) ^ (((uses >> 1) & 1) << (b + 3));
        }
        k += len;
    }
}

template <bool INV, bool STRIDED, bool FULL> cudaError_t set_attr()
{
    return cudaFuncSetAttribute(fft_stream_kernel<INV, STRIDED, FULL>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
}

}  // namespace

int stream_pass_init()
{
    if (ctx().attr_stream) return 0;   // per device
    TS_CUDA((set_attr<false, false, true>())); TS_CUDA((set_attr<true, false, true>()));
    TS_CUDA((set_attr<false, true, true>()));  TS_CUDA((set_attr<true, true, true>()));
    TS_CUDA((set_attr<false, true, false>())); TS_CUDA((set_attr<true, true, false>()));
    ctx().attr_stream = true;
    return 0;
}

int stream_pass(bool inverse, const PassParams &p)
{
    int rc = stream_pass_init();
    if (rc) return rc;
    StreamParams sp;
    sp.p = p;
    const bool strided = p.lo != 0;
    uint64_t n_hi, per_hi;
    if (strided) {
        if (p.src.shard_log != 31 || p.dst.shard_log != 31) return fail(TSTWO_ERR_BAD_ARG, "strided FFT pass on a row-sharded column set");
        n_hi = 1ull << (p.n - p.lo - 10);
        per_hi = ((uint64_t)p.n_cols) << (p.lo - 4);
    } else {
        n_hi = 1ull << (p.n - 12);
        per_hi = (uint64_t)(p.n_cols + 3) / 4;
    }
    if (n_hi * per_hi >= (1ull << 31)) return fail(TSTWO_ERR_BAD_ARG, "FFT pass too large (%llu tiles)", (unsigned long long)(n_hi * per_hi));
    sp.n_tiles = (uint32_t)(n_hi * per_hi);
    sp.tiles_per_hi = (uint32_t)per_hi;
    unsigned grid = (unsigned)ctx().sm_count;
    if (grid > sp.n_tiles) grid = sp.n_tiles;
    if (!strided) {
        if (inverse) TS_LAUNCH((fft_stream_kernel<true, false, true>), grid, THREADS, SMEM_BYTES, sp);
        else TS_LAUNCH((fft_stream_kernel<false, false, true>), grid, THREADS, SMEM_BYTES, sp);
    } else if (p.j_min == 0) {
        if (inverse) TS_LAUNCH((fft_stream_kernel<true, true, true>), grid, THREADS, SMEM_BYTES, sp);
        else TS_LAUNCH((fft_stream_kernel<false, true, true>), grid, THREADS, SMEM_BYTES, sp);
    } else {
        if (inverse) TS_LAUNCH((fft_stream_kernel<true, true, false>), grid, THREADS, SMEM_BYTES, sp);
        else TS_LAUNCH((fft_stream_kernel<false, true, false>), grid, THREADS, SMEM_BYTES, sp);
    }
    return 0;
}

}  // namespace tstwo
