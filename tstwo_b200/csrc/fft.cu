// Circle-domain FFT / iFFT over M31 for sm_100a: CpuCirclePoly.interpolate / evaluate of the reference
// (packages/core/src/backend/cpu/circle.ts:84-207) as radix-2^11 tiled passes.
//
// Layer i of a domain of log size n pairs indices differing in bit i; block h = idx >> (i+1) uses twiddle
//   T_i[h], T_i = tree[2^root_log - 2^(n-i) ..)  for i >= 1   (poly/utils.ts:78-99 slices from the END of the tree)
//   circle layer 0: +-T_1[(h>>1)^1]  with sign = bit0(h ^ (h>>1))  (circle.ts:270-278: [y,-y,-x,x])
// A pass loads a tile of 2^11 "rows" (index bits [lo, lo+11)) x 8 "lanes" into registers/shared memory and
// runs up to 11 layers on it in three register-resident radix-16 stages (row bits 10..7, 6..3, 3..0), exchanging
// through a padded shared-memory tile of uint4 units between stages.  Lanes are either 8 different columns
// (contiguous pass, lo = 0) or the 8 lowest index bits of one column (strided pass, lo >= 3), so every twiddle a
// thread loads is reused by 4 butterflies.  All twiddles are read pre-doubled (m31::mul2).
#include "common.cuh"
#include "fft_params.cuh"
#include "host_m31.h"
#include "m31.cuh"

namespace tstwo {

constexpr int FAST_MIN_LOG = 14;

// Tile shapes.  A tile holds 4096 uint4 units; a thread owns 2^DIG rows x 4 lanes and applies up to DIG layers per
// stage in registers, exchanging through shared memory between stages.
//   CONTIG : 4096 rows = index bits [0,12) of 4 different columns (the 4 lanes)
//   STRIDED: 1024 rows = index bits [lo,lo+10) x 16 consecutive indices (4 thread groups x 4 lanes) of one column
//   DIG = 4: 256 threads, 16 rows per thread, 3 stages;  DIG = 3: 512 threads, 8 rows per thread, 4 stages
template <bool STRIDED, int DIG> struct Shape {
    static constexpr int RHO = STRIDED ? 10 : 12;            // row bits per tile
    static constexpr int G = STRIDED ? 4 : 1;                // thread groups side by side in a row
    static constexpr int NE = 1 << DIG;
    static constexpr int THREADS = 4096 / NE;
    static constexpr int NST = DIG == 4 ? 3 : 4;             // stages, listed from the top row bits down
    __host__ __device__ static constexpr int S(int k) { return DIG == 4 ? (STRIDED ? (k == 0 ? 6 : k == 1 ? 2 : 0) : (k == 0 ? 8 : k == 1 ? 4 : 0))
                                                    : (STRIDED ? (k == 0 ? 7 : k == 1 ? 4 : k == 2 ? 2 : 0) : (k == 0 ? 9 : k == 1 ? 6 : k == 2 ? 3 : 0)); }
    __host__ __device__ static constexpr int N(int k) { return DIG == 4 ? (STRIDED ? (k == 2 ? 2 : 4) : 4) : (STRIDED ? (k >= 2 ? 2 : 3) : 3); }
    static constexpr int TW_WORDS = 1 << RHO;
    static constexpr int TILE_UNITS = ((1 << RHO) + ((1 << RHO) >> DIG)) * G;   // one pad row per 2^DIG rows
    static constexpr int SMEM = TILE_UNITS * 16 + TW_WORDS * 4;
};


__device__ __forceinline__ void cp_async16(void *smem, const void *gmem)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem));
}
__device__ __forceinline__ void cp_async4(void *smem, const void *gmem)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem));
}

// Stage all twiddles of the tile in shared memory as a heap: relative layer j (absolute i = lo + j) needs
// T_i[(tile_hi << (RHO-1-j)) + r], r < 2^(RHO-1-j); it is stored at tws[2^(RHO-1-j) + r].  Layers of >= 4 entries are
// fetched with 16-byte cp.async (source ranges are aligned to their size), the three small ones word by word.
template <int RHO, int THREADS>
__device__ __forceinline__ void tile_twiddles_async(uint32_t *tws, const PassParams &p, uint32_t tile_hi, int tid)
{
#pragma unroll
    for (int e = 4 * tid; e < (1 << RHO); e += 4 * THREADS) {
        if (e == 0) {
            for (int ee = 1; ee < 4; ee++) {          // heap[1] = top layer, heap[2..3] = the next one
                const int lg = ee >> 1;
                const int j = (RHO - 1) - lg, i = p.lo + j;
                if (j < p.j_min || i < 1) continue;
                cp_async4(tws + ee, p.tw2 + (p.tree_len - (1u << (p.n - i))) + (tile_hi << lg) + (ee - (1 << lg)));
            }
            continue;
        }
        const int lg = 31 - __clz(e);
        const int j = (RHO - 1) - lg, i = p.lo + j;
        if (j < p.j_min || i < 1) continue;           // the circle layer derives its twiddles from T_1
        cp_async16(tws + e, p.tw2 + (p.tree_len - (1u << (p.n - i))) + (tile_hi << lg) + (e - (1 << lg)));
    }
    asm volatile("cp.async.commit_group;");
}

// One radix-2^DIG stage on v[NE][4]: digit bit k <-> row bit S+k, layers j = S+k for k in [0, NL).
// FULL: every layer of the window is applied (no run-time j_min test, hence no control flow between layers);
// CIRCLE: the window starts at index bit 0, so relative layer 0 is the circle layer.
template <int NE, int RHO, int S, int NL, bool INV, bool FULL, bool CIRCLE>
__device__ __forceinline__ void stage(uint32_t (&v)[NE][4], const PassParams &p, const uint32_t *tws, int rbase)
{
#pragma unroll
    for (int kk = 0; kk < NL; kk++) {
        const int k = INV ? kk : NL - 1 - kk;
        const int j = S + k;
        if (!FULL && j < p.j_min) continue;            // uniform across the grid
        const int cnt = (NE / 2) >> k;
        uint32_t t[NE / 2];
        if (!(CIRCLE && j == 0)) {
            const uint32_t *q = tws + (1 << (RHO - 1 - j)) + (rbase >> (j + 1));
            if (cnt >= 4) {
#pragma unroll
                for (int c = 0; c < cnt / 4; c++) {
                    uint4 a = *reinterpret_cast<const uint4 *>(q + 4 * c);
                    t[4 * c] = a.x; t[4 * c + 1] = a.y; t[4 * c + 2] = a.z; t[4 * c + 3] = a.w;
                }
            } else if (cnt == 2) {
                uint2 a = *reinterpret_cast<const uint2 *>(q);
                t[0] = a.x; t[1] = a.y;
            } else t[0] = q[0];
        } else {
            // circle layer (j == 0, k == 0): T_1 entries (x, y) pairs of this thread's rows -> [y, -y, -x, x] each
            const uint32_t P2 = 0xfffffffeu;
#pragma unroll
            for (int c = 0; c < NE / 8; c++) {
                uint2 a = *reinterpret_cast<const uint2 *>(tws + (1 << (RHO - 2)) + (rbase >> 2) + 2 * c);
                t[4 * c] = a.y; t[4 * c + 1] = P2 - a.y; t[4 * c + 2] = P2 - a.x; t[4 * c + 3] = a.x;
            }
        }
#pragma unroll
        for (int d = 0; d < NE; d++) {
            if (d & (1 << k)) continue;
            const uint32_t tw = t[d >> (k + 1)];
#pragma unroll
            for (int q = 0; q < 4; q++) {
                if (INV) m31::ibutterfly(v[d][q], v[d | (1 << k)][q], tw);
                else m31::butterfly(v[d][q], v[d | (1 << k)][q], tw);
            }
        }
    }
}

// Thread -> rows map of the stage whose digit is row bits [S, S+DIG): rows rbase + (d << S), d = 0..2^DIG-1, with
// rbase = ((rg >> S) << (S+DIG)) | (rg & (2^S - 1)).
template <int DIG, int S> __device__ __forceinline__ int map_rbase(int rg) { return ((rg >> S) << (S + DIG)) | (rg & ((1 << S) - 1)); }
// Shared-memory unit (16 B) of (row r, group g): G*(r + (r >> DIG)) + g -- one pad row per 2^DIG rows.  Every quarter
// warp (8 lanes) then touches 8 distinct 16-byte bank groups in every map of both shapes (ncu: wavefronts == ideal).
template <bool STRIDED, int DIG, int S> __device__ __forceinline__ int map_unit0(int rg, int g)
{
    const int r = map_rbase<DIG, S>(rg);
    return Shape<STRIDED, DIG>::G * (r + (r >> DIG)) + g;
}
template <bool STRIDED, int DIG, int S> __device__ __forceinline__ constexpr int map_unit_off(int d)
{
    return Shape<STRIDED, DIG>::G * ((d << S) + ((d << S) >> DIG));
}
template <bool STRIDED, int DIG, int S, int NE> __device__ __forceinline__ void smem_store(uint4 *sm, int rg, int g, const uint32_t (&v)[NE][4])
{
    const int u0 = map_unit0<STRIDED, DIG, S>(rg, g);
#pragma unroll
    for (int d = 0; d < NE; d++) sm[u0 + map_unit_off<STRIDED, DIG, S>(d)] = make_uint4(v[d][0], v[d][1], v[d][2], v[d][3]);
}
template <bool STRIDED, int DIG, int S, int NE> __device__ __forceinline__ void smem_load(const uint4 *sm, int rg, int g, uint32_t (&v)[NE][4])
{
    const int u0 = map_unit0<STRIDED, DIG, S>(rg, g);
#pragma unroll
    for (int d = 0; d < NE; d++) {
        uint4 a = sm[u0 + map_unit_off<STRIDED, DIG, S>(d)];
        v[d][0] = a.x; v[d][1] = a.y; v[d][2] = a.z; v[d][3] = a.w;
    }
}

// global I/O.  STRIDED: the tile is rows x 16 consecutive indices of column blockIdx.y; CONTIG: the 4 lanes are
// columns 4*blockIdx.y .. +3.  All 64-bit address arithmetic is done once per thread (tile base); the digit offsets
// are compile-time (CONTIG) or a single runtime stride (STRIDED).  Only the contiguous pass honours the row-sharded
// layout of ColSet (a 4096-row tile never straddles a shard).
template <int S, bool STRIDED, int DIG, int NE>
__device__ __forceinline__ void gmem_load(const PassParams &p, uint32_t tile_hi, uint32_t mid, int rg, int g, uint32_t (&v)[NE][4])
{
    constexpr int RHO = Shape<STRIDED, DIG>::RHO;
    const int rbase = map_rbase<DIG, S>(rg);
    if (STRIDED) {
        const uint32_t base = (((tile_hi << (p.lo + RHO)) | (mid << 4) | (g << 2)) & p.src_mask) | ((uint32_t)rbase << p.lo);
        const uint4 *ptr = reinterpret_cast<const uint4 *>(p.src.col(blockIdx.y) + base);
        const uint32_t st = 1u << (p.lo + S - 2);    // uint4 units between digits
#pragma unroll
        for (int d = 0; d < NE; d++) {
            uint4 a = __ldcg(ptr + (size_t)d * st);
            v[d][0] = a.x; v[d][1] = a.y; v[d][2] = a.z; v[d][3] = a.w;
        }
    } else {
        const int c0 = blockIdx.y * 4;
        const uint64_t tile_off = p.src.off((tile_hi << RHO) & p.src_mask) + (uint32_t)rbase;
#pragma unroll
        for (int q = 0; q < 4; q++) {
            if (c0 + q < p.n_cols) {
                const uint32_t *ptr = p.src.col(c0 + q) + tile_off;
                if (S == 0) {
#pragma unroll
                    for (int dd = 0; dd < NE / 4; dd++) {
                        uint4 a = __ldg(reinterpret_cast<const uint4 *>(ptr) + dd);
                        v[4 * dd][q] = a.x; v[4 * dd + 1][q] = a.y; v[4 * dd + 2][q] = a.z; v[4 * dd + 3][q] = a.w;
                    }
                } else {
#pragma unroll
                    for (int d = 0; d < NE; d++) v[d][q] = __ldcg(ptr + (d << S));
                }
            } else {
#pragma unroll
                for (int d = 0; d < NE; d++) v[d][q] = 0;
            }
        }
    }
}
template <int S, bool STRIDED, int DIG, int NE>
__device__ __forceinline__ void gmem_store(const PassParams &p, uint32_t tile_hi, uint32_t mid, int rg, int g, const uint32_t (&v)[NE][4])
{
    constexpr int RHO = Shape<STRIDED, DIG>::RHO;
    const int rbase = map_rbase<DIG, S>(rg);
    if (STRIDED) {
        const uint32_t base = (tile_hi << (p.lo + RHO)) | (mid << 4) | (g << 2) | ((uint32_t)rbase << p.lo);
        uint4 *ptr = reinterpret_cast<uint4 *>(p.dst.col(blockIdx.y) + base);
        const uint32_t st = 1u << (p.lo + S - 2);
#pragma unroll
        for (int d = 0; d < NE; d++) __stcg(ptr + (size_t)d * st, make_uint4(v[d][0], v[d][1], v[d][2], v[d][3]));
    } else {
        const int c0 = blockIdx.y * 4;
        const uint64_t tile_off = p.dst.off(tile_hi << RHO) + (uint32_t)rbase;
#pragma unroll
        for (int q = 0; q < 4; q++) {
            if (c0 + q < p.n_cols) {
                uint32_t *ptr = p.dst.col(c0 + q) + tile_off;
                if (S == 0) {
#pragma unroll
                    for (int dd = 0; dd < NE / 4; dd++)
                        __stcg(reinterpret_cast<uint4 *>(ptr) + dd,
                               make_uint4(v[4 * dd][q], v[4 * dd + 1][q], v[4 * dd + 2][q], v[4 * dd + 3][q]));
                } else {
#pragma unroll
                    for (int d = 0; d < NE; d++) __stcg(ptr + (d << S), v[d][q]);
                }
            }
        }
    }
}

// one stage k of the shape: [exchange in] -> butterflies -> [exchange out]
template <bool INV, bool STRIDED, int DIG, bool FULL, int K>
__device__ __forceinline__ void run_stage(uint32_t (&v)[1 << DIG][4], const PassParams &p, uint4 *sm, const uint32_t *tws, int rg, int g,
                                          bool load_smem, bool store_smem)
{
    using SH = Shape<STRIDED, DIG>;
    constexpr int S = SH::S(K), N = SH::N(K);
    if (load_smem) smem_load<STRIDED, DIG, S, SH::NE>(sm, rg, g, v);
    stage<SH::NE, SH::RHO, S, N, INV, FULL, !STRIDED>(v, p, tws, map_rbase<DIG, S>(rg));
    if (store_smem) {
        smem_store<STRIDED, DIG, S, SH::NE>(sm, rg, g, v);
        __syncthreads();
    }
}

template <bool INV, bool STRIDED, int DIG, bool FULL>
__global__ void __launch_bounds__(4096 >> DIG, 2) fft_pass_kernel(const PassParams p)
{
    using SH = Shape<STRIDED, DIG>;
    constexpr int RHO = SH::RHO, NST = SH::NST, LAST = NST - 1;
    extern __shared__ uint4 sm[];
    uint32_t *tws = reinterpret_cast<uint32_t *>(sm + SH::TILE_UNITS);
    const int tid = threadIdx.x;
    const int g = STRIDED ? (tid & 3) : 0, rg = STRIDED ? (tid >> 2) : tid;
    uint32_t tile_hi, mid;
    if (STRIDED) {
        tile_hi = blockIdx.x >> (p.lo - 4);
        mid = blockIdx.x & ((1u << (p.lo - 4)) - 1);
    } else {
        tile_hi = blockIdx.x;
        mid = 0;
    }
    tile_twiddles_async<RHO, SH::THREADS>(tws, p, tile_hi, tid);
    uint32_t v[SH::NE][4];
    if (!INV) {
        gmem_load<SH::S(0), STRIDED, DIG, SH::NE>(p, tile_hi, mid, rg, g, v);
        asm volatile("cp.async.wait_all;" ::: "memory");
        __syncthreads();
        run_stage<INV, STRIDED, DIG, FULL, 0>(v, p, sm, tws, rg, g, false, true);
        run_stage<INV, STRIDED, DIG, FULL, 1>(v, p, sm, tws, rg, g, true, true);
        if (NST == 4) run_stage<INV, STRIDED, DIG, FULL, (NST == 4 ? 2 : 1)>(v, p, sm, tws, rg, g, true, true);
        run_stage<INV, STRIDED, DIG, FULL, LAST>(v, p, sm, tws, rg, g, true, false);
        gmem_store<SH::S(LAST), STRIDED, DIG, SH::NE>(p, tile_hi, mid, rg, g, v);
    } else {
        gmem_load<SH::S(LAST), STRIDED, DIG, SH::NE>(p, tile_hi, mid, rg, g, v);
        asm volatile("cp.async.wait_all;" ::: "memory");
        __syncthreads();
        run_stage<INV, STRIDED, DIG, FULL, LAST>(v, p, sm, tws, rg, g, false, true);
        if (NST == 4) run_stage<INV, STRIDED, DIG, FULL, (NST == 4 ? 2 : 1)>(v, p, sm, tws, rg, g, true, true);
        run_stage<INV, STRIDED, DIG, FULL, 1>(v, p, sm, tws, rg, g, true, true);
        run_stage<INV, STRIDED, DIG, FULL, 0>(v, p, sm, tws, rg, g, true, false);
        if (p.scale2) {
#pragma unroll
            for (int d = 0; d < SH::NE; d++)
#pragma unroll
                for (int q = 0; q < 4; q++) v[d][q] = m31::mul2(v[d][q], p.scale2);
        }
        gmem_store<SH::S(0), STRIDED, DIG, SH::NE>(p, tile_hi, mid, rg, g, v);
    }
}

// ---------------------------------------------------------------- naive per-layer kernels (any size; used for
// log sizes below FAST_MIN_LOG and as the on-device cross-check of the tiled passes)
__global__ void fft_layer_kernel(ColSet cols, int n, int i, const uint32_t *tw2, uint32_t tree_len, int inverse)
{
    uint64_t b = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >> (n - 1)) return;
    uint32_t *v = cols.col(blockIdx.y);
    uint32_t h = (uint32_t)(b >> i), l = (uint32_t)b & ((1u << i) - 1);
    uint32_t i0 = (h << (i + 1)) + l, i1 = i0 + (1u << i);
    uint32_t t;
    if (i >= 1) t = tw2[tree_len - (1u << (n - i)) + h];
    else {
        t = tw2[tree_len - (1u << (n - 1)) + ((h >> 1) ^ 1)];
        if ((h ^ (h >> 1)) & 1) t = 0xfffffffeu - t;
    }
    uint32_t a = v[i0], c = v[i1];
    if (inverse) m31::ibutterfly(a, c, t); else m31::butterfly(a, c, t);
    v[i0] = a; v[i1] = c;
}
__global__ void scale_kernel(ColSet cols, int n, uint32_t scale2)
{
    uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >> n) return;
    uint32_t *v = cols.col(blockIdx.y);
    v[e] = m31::mul2(v[e], scale2);
}
// dst[e] = e < 2^poly_log ? src[e] : 0   (CpuCirclePoly.extend, circle.ts:71-82)
__global__ void extend_kernel(ColSet src, ColSet dst, int poly_log, int n)
{
    uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >> n) return;
    dst.col(blockIdx.y)[e] = (e >> poly_log) ? 0u : src.col(blockIdx.y)[e];
}
__global__ void copy_cols_kernel(ColSet src, ColSet dst, int n)
{
    uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >> n) return;
    dst.col(blockIdx.y)[e] = src.col(blockIdx.y)[e];
}
__global__ void swap57_kernel(ColSet cols)
{
    uint32_t *v = cols.col(blockIdx.x);
    uint32_t t = v[5]; v[5] = v[7]; v[7] = t;
}
// closed forms for log 1 and 2 (circle.ts:93-110, 153-185); x, y = half coset initial point, pre-doubled factors
__global__ void small_eval_kernel(ColSet cols, int n, uint32_t x2, uint32_t y2, uint32_t ny2)
{
    uint32_t *v = cols.col(blockIdx.x);
    if (n == 1) { uint32_t a = v[0], b = v[1]; m31::butterfly(a, b, y2); v[0] = a; v[1] = b; return; }
    uint32_t v0 = v[0], v1 = v[1], v2 = v[2], v3 = v[3];
    m31::butterfly(v0, v2, x2); m31::butterfly(v1, v3, x2);
    m31::butterfly(v0, v1, y2); m31::butterfly(v2, v3, ny2);
    v[0] = v0; v[1] = v1; v[2] = v2; v[3] = v3;
}
__global__ void small_interp_kernel(ColSet cols, int n, uint32_t xi2, uint32_t yi2, uint32_t nyi2, uint32_t ninv2)
{
    uint32_t *v = cols.col(blockIdx.x);
    if (n == 1) {
        uint32_t a = v[0], b = v[1];
        m31::ibutterfly(a, b, yi2);
        v[0] = m31::mul2(a, ninv2); v[1] = m31::mul2(b, ninv2);
        return;
    }
    uint32_t v0 = v[0], v1 = v[1], v2 = v[2], v3 = v[3];
    m31::ibutterfly(v0, v1, yi2); m31::ibutterfly(v2, v3, nyi2);
    m31::ibutterfly(v0, v2, xi2); m31::ibutterfly(v1, v3, xi2);
    v[0] = m31::mul2(v0, ninv2); v[1] = m31::mul2(v1, ninv2); v[2] = m31::mul2(v2, ninv2); v[3] = m31::mul2(v3, ninv2);
}

static bool g_force_naive = false;

static int g_dig = 4;   // rows per thread = 2^g_dig (tstwo_debug_set_fft_radix): 4 -> 256 threads (default), 3 -> 512 threads

template <bool INV, bool STRIDED, int DIG, bool FULL> static cudaError_t set_attr()
{
    return cudaFuncSetAttribute(fft_pass_kernel<INV, STRIDED, DIG, FULL>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                Shape<STRIDED, DIG>::SMEM);
}
template <int DIG> static int set_attrs()
{
    TS_CUDA((set_attr<false, false, DIG, true>())); TS_CUDA((set_attr<true, false, DIG, true>()));
    TS_CUDA((set_attr<false, true, DIG, true>()));  TS_CUDA((set_attr<true, true, DIG, true>()));
    TS_CUDA((set_attr<false, true, DIG, false>())); TS_CUDA((set_attr<true, true, DIG, false>()));
    return 0;
}
static int ensure_smem_attr()
{
    if (ctx().attr_fft) return 0;   // per device: cudaFuncSetAttribute applies to the current device
    int rc = set_attrs<3>();
    if (rc) return rc;
    if ((rc = set_attrs<4>())) return rc;
    ctx().attr_fft = true;
    return 0;
}

struct Pass { int lo, j_min, win; };
// passes in ascending layer order for L real layers (tests/proto_fft.py): one contiguous 12-layer pass, then
// strided 10-layer windows; the top window is aligned to L and skips the layers already done.
// The TMA passes (fft_tma.cu) use a 14-layer contiguous pass and 8-layer strided windows instead.
constexpr int W0 = Shape<false, 3>::RHO, W1 = Shape<true, 3>::RHO;
constexpr int MAX_PASSES = 4;
static int g_rows9 = 1;        // tstwo_debug_set_fft_rows9: 0 = never plan the 512-row pass
static int plan_passes(int L, Pass out[MAX_PASSES], int w0, int w1)
{
    int np = 0, done = w0;
    out[np++] = { 0, 0, 0 };
    if (L <= done) return np;
    // exactly nine layers left above the bulk-copy contiguous pass (L = 23: config 5's trace): ONE 512-row pass instead of an
    // 8-layer window plus a pass that moves every byte again for a single layer (100 x 2^23: 13.7 -> ~10.5 ms per LDE)
    if (g_rows9 && w0 == TMA_W0 && w1 == TMA_W1 && L - done == TMA_W9) { out[np++] = { done, 0, TMA_W9 }; return np; }
    while (L - done > w1) { out[np++] = { done, 0, 0 }; done += w1; }
    out[np++] = { L - w1, done - (L - w1), 0 };
    return np;
}

template <int DIG> static int launch_pass_t(bool inverse, const PassParams &p)
{
    constexpr int T = 4096 >> DIG;
    if (p.lo == 0) {
        dim3 grid(1u << (p.n - W0), (unsigned)((p.n_cols + 3) / 4));
        if (inverse) TS_LAUNCH((fft_pass_kernel<true, false, DIG, true>), grid, T, (Shape<false, DIG>::SMEM), p);
        else TS_LAUNCH((fft_pass_kernel<false, false, DIG, true>), grid, T, (Shape<false, DIG>::SMEM), p);
    } else {
        if (p.src.shard_log != 31 || p.dst.shard_log != 31) return fail(TSTWO_ERR_BAD_ARG, "strided FFT pass on a row-sharded column set");
        dim3 grid(1u << (p.n - W1 - 4), (unsigned)p.n_cols);
        const bool full = p.j_min == 0;
        if (inverse) {
            if (full) TS_LAUNCH((fft_pass_kernel<true, true, DIG, true>), grid, T, (Shape<true, DIG>::SMEM), p);
            else TS_LAUNCH((fft_pass_kernel<true, true, DIG, false>), grid, T, (Shape<true, DIG>::SMEM), p);
        } else {
            if (full) TS_LAUNCH((fft_pass_kernel<false, true, DIG, true>), grid, T, (Shape<true, DIG>::SMEM), p);
            else TS_LAUNCH((fft_pass_kernel<false, true, DIG, false>), grid, T, (Shape<true, DIG>::SMEM), p);
        }
    }
    return 0;
}
// 2: persistent passes fed by bulk copies (fft_tma.cu, default where the column sets allow it); 1: persistent passes
// with per-thread cp.async (fft_stream.cu); 0: one CTA per tile (this file)
static int g_stream = 2;
static int g_last_mode = -1;   // what the last tiled transform actually ran (tstwo_debug_last_fft_mode)
static int launch_pass(int mode, bool inverse, const PassParams &p)
{
    if (mode == 2) return tma_pass(inverse, p);
    if (mode == 1) return stream_pass(inverse, p);
    return g_dig == 4 ? launch_pass_t<4>(inverse, p) : launch_pass_t<3>(inverse, p);
}
// the bulk-copy passes are used when every pass of the transform can address its column sets (fft_tma.cu)
static int pick_mode(bool inverse, const ColSet &src, const ColSet &dst, const ColSet &work, int n_cols, int L)
{
    if (g_stream != 2) return g_stream;
    Pass ps[MAX_PASSES];
    const int np = plan_passes(L, ps, TMA_W0, TMA_W1);
    for (int k = 0; k < np; k++) {
        PassParams p;
        p.n_cols = n_cols; p.lo = ps[k].lo; p.win = ps[k].win;
        if (inverse) { p.src = (k == 0) ? src : dst; p.dst = dst; }
        else { p.src = (k == np - 1) ? src : work; p.dst = (k == 0) ? dst : work; }
        if (!tma_pass_supported(p)) return 1;
    }
    return 2;
}

static uint32_t dbl(uint32_t v) { return v + v; }

int fft_interpolate(const ColSet &src, const ColSet &dst, int n_cols, int n, const TwiddleTree *t, uint32_t flags)
{
    if (n_cols <= 0) return 0;
    const uint32_t tree_len = 1u << t->root_log;
    const uint64_t N = 1ull << n;
    const bool in_place = (src.base == dst.base && src.table == dst.table && src.stride == dst.stride);
    const uint32_t inv_n2 = dbl(host::inv((uint32_t)(N % host::P)));
    if (n >= FAST_MIN_LOG && !g_force_naive) {
        int rc = ensure_smem_attr();
        if (rc) return rc;
        const int mode = g_last_mode = pick_mode(true, src, dst, dst, n_cols, n);
        Pass ps[MAX_PASSES];
        int np = mode == 2 ? plan_passes(n, ps, TMA_W0, TMA_W1) : plan_passes(n, ps, W0, W1);
        for (int k = 0; k < np; k++) {
            PassParams p;
            p.src = (k == 0) ? src : dst; p.dst = dst; p.n_cols = n_cols; p.n = n; p.src_mask = (uint32_t)(N - 1);
            p.lo = ps[k].lo; p.j_min = ps[k].j_min; p.win = ps[k].win; p.tw2 = t->itw2; p.tree_len = tree_len;
            p.scale2 = (k == np - 1) ? inv_n2 : 0; p.zero = 0;
            rc = launch_pass(mode, true, p);
            if (rc) return rc;
        }
        return 0;
    }
    dim3 egrid((unsigned)((N + 255) / 256), (unsigned)n_cols);
    if (!in_place) TS_LAUNCH(copy_cols_kernel, egrid, 256, 0, src, dst, n);
    if ((flags & TSTWO_FLAG_QUIRKS) && n == 3) TS_LAUNCH(swap57_kernel, n_cols, 1, 0, dst);
    if (n <= 2) {
        // circle.ts:153-185
        uint32_t half_init = (t->root_init << (t->root_log - (n - 1))) & 0x7fffffffu;
        host::Pt ip = host::index_to_point(half_init);
        uint32_t xi = n == 2 ? host::inv(ip.x) : 0, yi = host::inv(ip.y);
        TS_LAUNCH(small_interp_kernel, n_cols, 1, 0, dst, n, dbl(xi), dbl(yi), dbl(host::neg(yi)), inv_n2);
        return 0;
    }
    dim3 bgrid((unsigned)((N / 2 + 255) / 256), (unsigned)n_cols);
    for (int i = 0; i < n; i++) TS_LAUNCH(fft_layer_kernel, bgrid, 256, 0, dst, n, i, t->itw2, tree_len, 1);
    TS_LAUNCH(scale_kernel, egrid, 256, 0, dst, n, inv_n2);
    return 0;
}

// `work`: plain (non-sharded) column set the strided passes run in; equals dst unless dst is row-sharded, in which
// case only the last (contiguous) pass writes dst.  pass_select (tiled sizes only): 0 = the whole transform, 1 = every
// pass but the last, 2 = the last pass alone (the one that writes dst) -- the multi-GPU path runs the two parts on different
// streams so that the pass whose stores cross NVLink overlaps the next column group's transforms.
int fft_evaluate(const ColSet &src, int L, const ColSet &dst, int n_cols, int n, const TwiddleTree *t, uint32_t flags,
                 const ColSet *work_in, int pass_select)
{
    if (n_cols <= 0) return 0;
    const ColSet &work = work_in ? *work_in : dst;
    const uint32_t tree_len = 1u << t->root_log;
    const uint64_t N = 1ull << n;
    if (L >= FAST_MIN_LOG && !g_force_naive) {
        int rc = ensure_smem_attr();
        if (rc) return rc;
        const int mode = g_last_mode = pick_mode(false, src, dst, work, n_cols, L);
        if (dst.shard_bases && mode != 2) return fail(TSTWO_ERR_BAD_ARG, "per-shard output bases need the bulk-copy FFT passes");
        Pass ps[MAX_PASSES];
        int np = mode == 2 ? plan_passes(L, ps, TMA_W0, TMA_W1) : plan_passes(L, ps, W0, W1);
        if (pass_select && np < 2) return fail(TSTWO_ERR_BAD_ARG, "split evaluation needs at least two passes");
        for (int k = np - 1; k >= 0; k--) {
            if ((pass_select == 1 && k == 0) || (pass_select == 2 && k != 0)) continue;
            PassParams p;
            const bool first = (k == np - 1);
            p.src = first ? src : work; p.dst = (k == 0) ? dst : work; p.n_cols = n_cols; p.n = n;
            p.src_mask = first ? (uint32_t)((1ull << L) - 1) : (uint32_t)(N - 1);
            p.lo = ps[k].lo; p.j_min = ps[k].j_min; p.win = ps[k].win; p.tw2 = t->tw2; p.tree_len = tree_len; p.scale2 = 0; p.zero = 0;
            rc = launch_pass(mode, false, p);
            if (rc) return rc;
        }
        return 0;
    }
    if (dst.shard_log != 31 || dst.shard_bases || pass_select) return fail(TSTWO_ERR_BAD_ARG, "row-sharded output needs poly_log >= %d", FAST_MIN_LOG);
    dim3 egrid((unsigned)((N + 255) / 256), (unsigned)n_cols);
    const bool in_place = (src.base == dst.base && src.table == dst.table && src.stride == dst.stride);
    if (!(in_place && L == n)) TS_LAUNCH(extend_kernel, egrid, 256, 0, src, dst, L, n);
    if (n <= 2) {
        uint32_t half_init = (t->root_init << (t->root_log - (n - 1))) & 0x7fffffffu;
        host::Pt ip = host::index_to_point(half_init);
        TS_LAUNCH(small_eval_kernel, n_cols, 1, 0, dst, n, dbl(ip.x), dbl(ip.y), dbl(host::neg(ip.y)));
        return 0;
    }
    dim3 bgrid((unsigned)((N / 2 + 255) / 256), (unsigned)n_cols);
    for (int i = n - 1; i >= 0; i--) TS_LAUNCH(fft_layer_kernel, bgrid, 256, 0, dst, n, i, t->tw2, tree_len, 0);
    if ((flags & TSTWO_FLAG_QUIRKS) && n == 3) TS_LAUNCH(swap57_kernel, n_cols, 1, 0, dst);
    return 0;
}

}  // namespace tstwo

using namespace tstwo;

extern "C" {

// test hook: force the per-layer kernels for every size (used by tests to cross-check the tiled passes on device)
int tstwo_debug_force_naive_fft(int on) { g_force_naive = on != 0; return 0; }
int tstwo_debug_set_fft_rows9(int on) { g_rows9 = on != 0; return 0; }
// tuning hook: rows per thread of the tiled passes (8 -> 512-thread CTAs, 16 -> 256-thread CTAs)
int tstwo_debug_set_fft_radix(int rows_per_thread) { g_dig = rows_per_thread == 8 ? 3 : 4; return 0; }
// tuning / test hook: 2 = bulk-copy persistent passes (default), 1 = cp.async persistent passes, 0 = one CTA per tile
// test hook: pass family used by the last tiled transform (2 bulk-copy, 1 cp.async persistent, 0 one CTA per tile)
int tstwo_debug_last_fft_mode(void) { return g_last_mode; }
int tstwo_debug_set_fft_stream(int mode) { g_stream = mode < 0 ? 0 : mode > 2 ? 2 : mode; return 0; }

static int check_tree(tstwo_twiddles h, uint32_t half_init, int half_log, TwiddleTree **out)
{
    TwiddleTree *t = get_tree(h);
    if (!t) return fail(TSTWO_ERR_BAD_ARG, "unknown twiddle tree handle");
    if (!is_doubling_of(half_init, half_log, t)) return fail(TSTWO_ERR_TWIDDLE_MISMATCH, "twiddle tree mismatch");
    *out = t;
    return 0;
}

int tstwo_interpolate_columns(const tstwo_dptr *evals, const tstwo_dptr *coeffs, int n_cols, int log_size,
                              uint32_t half_init, tstwo_twiddles tw, uint32_t flags)
{
    TS_REQUIRE_INIT();
    if (log_size < 1 || log_size > 30 || n_cols < 0) return fail(TSTWO_ERR_BAD_ARG, "bad interpolate arguments");
    TwiddleTree *t;
    int rc = check_tree(tw, half_init, log_size - 1, &t);
    if (rc) return rc;
    ColSet s, d;
    if ((rc = make_colset(evals, n_cols, 0, &s))) return rc;
    if ((rc = make_colset(coeffs, n_cols, 1, &d))) return rc;
    return fft_interpolate(s, d, n_cols, log_size, t, flags);
}

int tstwo_evaluate_columns(const tstwo_dptr *coeffs, int poly_log, const tstwo_dptr *evals, int n_cols, int domain_log,
                           uint32_t half_init, tstwo_twiddles tw, uint32_t flags)
{
    TS_REQUIRE_INIT();
    if (domain_log < 1 || domain_log > 30 || poly_log < 0 || n_cols < 0) return fail(TSTWO_ERR_BAD_ARG, "bad evaluate arguments");
    if (poly_log > domain_log) return fail(TSTWO_ERR_SIZE_MISMATCH, "log size too small");   // circle.ts:72
    TwiddleTree *t;
    int rc = check_tree(tw, half_init, domain_log - 1, &t);
    if (rc) return rc;
    ColSet s, d;
    if ((rc = make_colset(coeffs, n_cols, 0, &s))) return rc;
    if ((rc = make_colset(evals, n_cols, 1, &d))) return rc;
    return fft_evaluate(s, poly_log, d, n_cols, domain_log, t, flags);
}

}  // extern "C"
