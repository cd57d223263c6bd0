// libtstwo_cuda.so: lifecycle, column memory, bit-reverse and twiddle-tree entry points of include/tstwo_cuda.h.
#include <cstdarg>
#include <cstdlib>
#include <ctime>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <unordered_map>
#include <vector>

#include "common.cuh"
#include "circle_dev.cuh"

namespace tstwo { namespace { void release_cached_blocks(); std::mutex g_alloc_mu; } }   // column allocator, below (the mutex guards its maps)

namespace tstwo {

static thread_local char g_err[512] = "";
static Context g_ctxs[MAX_SLOTS];
static thread_local int g_cur = 0;
static int g_slots = 0;
Context &ctx() { return g_ctxs[g_cur]; }
Context &ctx_at(int slot) { return g_ctxs[slot]; }
int ctx_slots() { return g_slots; }
int ctx_select(int slot)
{
    const int prev = g_cur;
    if (slot != prev || true) {
        g_cur = slot;
        if (g_ctxs[slot].ready) cudaSetDevice(g_ctxs[slot].device);
    }
    return prev;
}

int fail(int code, const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
    return code;
}

static const int PTR_SLOTS = 4;
static const size_t PTR_SLOT_CAP = 4096;   // addresses per slot
void scatter_release();                     // side streams of tstwo_scatter_shards (defined with them)

int trace_level()
{
    static const int level = [] { const char *e = getenv("TSTWO_TRACE"); return e ? atoi(e) : 0; }();
    return level;
}
double trace_now()
{
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
}
void trace_span_end(const char *name, double t0)
{
    if (trace_level() >= 2 && ctx().ready) cudaStreamSynchronize(ctx().stream);
    fprintf(stderr, "tstwo: %-32s %9.3f ms%s\n", name, trace_now() - t0, trace_level() >= 2 ? " (stream drained)" : "");
}

int make_colset(const tstwo_dptr *cols, int n_cols, int slot, ColSet *out)
{
    out->base = nullptr; out->stride = 0; out->table = nullptr; out->aligned16 = true;
    if (n_cols <= 0) return 0;
    if (!cols) return fail(TSTWO_ERR_BAD_ARG, "null column list");
    for (int c = 0; c < n_cols; c++) out->aligned16 = out->aligned16 && (cols[c] & 15) == 0;
    bool uniform = true;
    int64_t stride = n_cols > 1 ? (int64_t)(cols[1] - cols[0]) : 0;
    for (int c = 1; c < n_cols && uniform; c++) uniform = (int64_t)(cols[c] - cols[c - 1]) == stride;
    if (uniform && stride >= 0 && (stride % 4) == 0) {
        out->base = reinterpret_cast<const uint32_t *>(cols[0]);
        out->stride = (uint64_t)stride / 4;
        return 0;
    }
    if ((size_t)n_cols > PTR_SLOT_CAP || slot >= PTR_SLOTS) return fail(TSTWO_ERR_BAD_ARG, "too many scattered columns (%d)", n_cols);
    Context &c = ctx();
    uint64_t *dst = c.d_ptr_table + (size_t)slot * PTR_SLOT_CAP;
    TS_CUDA(cudaMemcpyAsync(dst, cols, sizeof(uint64_t) * (size_t)n_cols, cudaMemcpyHostToDevice, c.stream));
    out->table = dst;
    return 0;
}

// ---------------------------------------------------------------- bit reverse (backend/cpu/index.ts:62-79)
__global__ void bit_reverse_kernel(uint32_t *a, int log_n)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >> log_n) return;
    uint32_t j = m31::bitrev((uint32_t)i, log_n);
    if (j > i) { uint32_t t = a[i]; a[i] = a[j]; a[j] = t; }
}

// Tiled form for 2^k >= 2^12 elements.  Index = (a : 6 high bits | m : k - 12 middle bits | b : 6 low bits); its bit reversal is
// (rev b | rev m | rev a), so the 64 x 64 tile of all (a, b) with middle bits m goes, transposed and with a and b reversed, into the
// tile with middle bits rev(m).  One CTA swaps the pair of tiles (m, rev m) through shared memory: every global access is a run
// of 64 consecutive words (256 B) instead of one scattered word per thread.
constexpr int BR_T = 6, BR_DIM = 1 << BR_T, BR_PAD = BR_DIM + 1;
__global__ void __launch_bounds__(256) bit_reverse_tiled_kernel(uint32_t *__restrict__ a, int log_n)
{
    __shared__ uint32_t sa[BR_DIM * BR_PAD], sb[BR_DIM * BR_PAD];
    const int mid_bits = log_n - 2 * BR_T;
    const uint32_t m = blockIdx.x, mr = mid_bits ? m31::bitrev(m, mid_bits) : 0u;
    if (m > mr) return;
    const int col = threadIdx.x & (BR_DIM - 1), row0 = threadIdx.x >> BR_T;      // 4 rows per sweep of 256 threads
    uint32_t *ta = a + ((uint64_t)m << BR_T), *tb = a + ((uint64_t)mr << BR_T);
    const uint64_t row_stride = 1ull << (log_n - BR_T);
#pragma unroll 4
    for (int r = row0; r < BR_DIM; r += 4) {
        sa[r * BR_PAD + col] = ta[r * row_stride + col];
        if (m != mr) sb[r * BR_PAD + col] = tb[r * row_stride + col];
    }
    __syncthreads();
    const int cr = (int)(__brev((unsigned)col) >> (32 - BR_T));
#pragma unroll 4
    for (int r = row0; r < BR_DIM; r += 4) {
        const int rr = (int)(__brev((unsigned)r) >> (32 - BR_T));
        // output (row r, column col) of tile rev(m) <- input (row rev(col), column rev(r)) of tile m
        tb[r * row_stride + col] = sa[cr * BR_PAD + rr];
        if (m != mr) ta[r * row_stride + col] = sb[cr * BR_PAD + rr];
    }
}

// ---------------------------------------------------------------- twiddles (backend/cpu/circle.ts:210-239)

// One thread per entry e of the 2^log tree buffer.  Layer i holds bitrev(x of first half of the i-times doubled
// coset); the last entry is the pad value 1.
__global__ void twiddle_tree_kernel(uint32_t *tw, uint32_t *itw, uint32_t *tw2, uint32_t *itw2, uint32_t init, int log)
{
    uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t n = 1ull << log;
    if (e >= n) return;
    uint32_t x;
    if (e == n - 1) x = 1;
    else {
        uint32_t r = (uint32_t)(n - e);              // 2 .. 2^log
        int m = 31 - __clz(r - 1);                   // layer half-size = 2^m
        uint32_t j = (1u << (m + 1)) - r;
        int i = log - 1 - m;                         // layer number
        uint32_t ci = (init << i) & 0x7fffffffu;     // init * 2^i mod 2^31
        uint32_t step = 1u << (31 - (m + 1));
        uint32_t idx = (ci + m31::bitrev(j, m) * step) & 0x7fffffffu;
        x = index_to_point(idx).x;
    }
    uint32_t ix = m31::inv(x);
    tw[e] = x; itw[e] = ix; tw2[e] = x + x; itw2[e] = ix + ix;
}

// out[i] = 2 / coord(init + bitrev(i, m) * step)  (pre-doubled inverse), i < 2^m; which: 0 = x, 1 = y
__global__ void coset_inv_coord_kernel(uint32_t *out2, uint32_t init, uint32_t step, int m, int which)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >> m) return;
    uint32_t idx = (init + m31::bitrev((uint32_t)i, m) * step) & 0x7fffffffu;
    m31::Pt p = index_to_point(idx);
    uint32_t v = m31::inv(which ? p.y : p.x);
    out2[i] = v + v;
}

int launch_coset_points(uint32_t *out2, int which, uint32_t init, int step_log, int m)
{
    uint64_t n = 1ull << m;
    uint32_t step = host::subgroup_gen(step_log);
    TS_LAUNCH(coset_inv_coord_kernel, (unsigned)((n + 255) / 256), 256, 0, out2, init, step, m, which);
    return 0;
}

static std::mutex g_tree_mu;
static std::unordered_map<uint64_t, TwiddleTree *> g_trees;
// Freed trees are kept (at most TW_SPARE_MAX, oldest dropped first) and handed out again by tstwo_twiddles_precompute for the same
// (device, coset): a prover that builds the tree of a domain inside every prove() (DomainEvaluationAccumulator.finalize,
// air/accumulator.ts:189-238) then pays neither the table kernel nor the device-wide synchronisation of cudaFree each time
// (15 ms of config 5's 35 ms prove()).  The tables are read-only, so reuse while earlier kernels still read them is safe.
static std::vector<TwiddleTree *> g_spare_trees;
constexpr size_t TW_SPARE_MAX = 4;
static void drop_spare_trees_locked()
{
    for (TwiddleTree *t : g_spare_trees) {
        if (g_ctxs[t->slot].ready) cudaSetDevice(g_ctxs[t->slot].device);
        cudaFree(t->tw);
        delete t;
    }
    g_spare_trees.clear();
    if (g_ctxs[g_cur].ready) cudaSetDevice(g_ctxs[g_cur].device);
}
static uint64_t g_next_tree = 1;

TwiddleTree *get_tree(tstwo_twiddles h)
{
    std::lock_guard<std::mutex> lk(g_tree_mu);
    auto it = g_trees.find(h);
    // a tree is device memory: only the slot it was built on may use it
    return it == g_trees.end() || it->second->slot != ctx().slot ? nullptr : it->second;
}
bool is_doubling_of(uint32_t init, int log, const TwiddleTree *t)
{
    if (log > t->root_log) return false;
    uint32_t v = t->root_init & 0x7fffffffu;
    for (int i = 0; i < t->root_log - log; i++) v = (v * 2u) & 0x7fffffffu;
    return v == (init & 0x7fffffffu);
}

}  // namespace tstwo

using namespace tstwo;

extern "C" {

}  // extern "C"
namespace tstwo {
namespace {
// brings up one slot on `device`: streams, pointer table, pool threshold, the per-device constant tables
int init_slot(int slot, int device)
{
    Context &c = g_ctxs[slot];
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        return fail(TSTWO_ERR_CUDA, "no CUDA device available (%s); libtstwo_cuda has no CPU fallback",
                    e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
    if (device < 0 || device >= n) return fail(TSTWO_ERR_BAD_ARG, "device %d out of range [0,%d)", device, n);
    TS_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    TS_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10)
        return fail(TSTWO_ERR_CUDA, "device is sm_%d%d; this library is built for sm_100a only", prop.major, prop.minor);
    c = Context();
    c.slot = slot;
    c.device = device;
    c.sm_count = prop.multiProcessorCount;
    TS_CUDA(cudaStreamCreateWithFlags(&c.own_stream, cudaStreamNonBlocking));
    TS_CUDA(cudaStreamCreateWithFlags(&c.copy_stream, cudaStreamNonBlocking));
    c.stream = c.own_stream;
    TS_CUDA(cudaMalloc(&c.d_ptr_table, sizeof(uint64_t) * PTR_SLOTS * PTR_SLOT_CAP));
    {   // temporaries of the gather / decommit calls come from the stream-ordered pool: keep what it holds across synchronisations
        // (the default threshold 0 hands the memory back at every sync, ~0.3 ms per call afterwards)
        cudaMemPool_t pool;
        uint64_t keep = 256ull << 20;
        TS_CUDA(cudaDeviceGetDefaultMemPool(&pool, device));
        TS_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));
    }
    TS_CUDA(upload_gen_table());
    TS_CUDA(elementwise_init());
    c.launches = 0;
    c.ready = true;
    return 0;
}
void shutdown_slot(int slot)
{
    Context &c = g_ctxs[slot];
    if (!c.ready) return;
    const int prev = ctx_select(slot);
    cudaDeviceSynchronize();
    {
        std::lock_guard<std::mutex> lock(g_alloc_mu);
        release_cached_blocks();
    }
    pipeline_release();
    if (c.ew_scratch) cudaFree(c.ew_scratch);
    cudaFree(c.d_ptr_table);
    cudaStreamDestroy(c.own_stream);
    cudaStreamDestroy(c.copy_stream);
    c = Context();
    g_cur = prev;
}
}  // namespace
void multi_release();   // multi.cu: per-device buffers of tstwo_lde_commit_multi
}  // namespace tstwo
extern "C" {

int tstwo_init(int device)
{
    Context &c = g_ctxs[0];
    if (c.ready && c.device == device && g_slots == 1) { ctx_select(0); return 0; }
    if (g_slots) tstwo_shutdown();
    g_cur = 0;
    int rc = init_slot(0, device);
    if (rc) return rc;
    g_slots = 1;
    return 0;
}

// One process, several devices (SURVEY 8b "Conventions": the reference host is ONE single-threaded process, so the multi-GPU
// fan-out lives inside the library).  Slot k drives device ids[k]; every pair gets peer access (NVLink).  Slot 0 is current.
int tstwo_init_devices(int n, const int *ids)
{
    // ids == NULL: the devices named by the environment, TSTWO_CUDA_DEVICES="0,1,2,3" (SURVEY 5); n = 0 takes all of them, else the first n
    int env_ids[MAX_SLOTS];
    if (!ids) {
        const char *e = getenv("TSTWO_CUDA_DEVICES");
        if (!e || !*e) return fail(TSTWO_ERR_BAD_ARG, "tstwo_init_devices: no device list given and TSTWO_CUDA_DEVICES is not set");
        int m = 0;
        for (const char *p = e; *p;) {
            char *end = nullptr;
            const long v = strtol(p, &end, 10);
            if (end == p || v < 0 || m >= MAX_SLOTS) return fail(TSTWO_ERR_BAD_ARG, "TSTWO_CUDA_DEVICES: cannot parse '%s'", e);
            env_ids[m++] = (int)v;
            p = (*end == ',') ? end + 1 : end;
            if (*end && *end != ',') return fail(TSTWO_ERR_BAD_ARG, "TSTWO_CUDA_DEVICES: cannot parse '%s'", e);
        }
        if (n == 0) n = m;
        if (n > m) return fail(TSTWO_ERR_BAD_ARG, "TSTWO_CUDA_DEVICES names %d devices, %d asked for", m, n);
        ids = env_ids;
    }
    if (n < 1 || n > MAX_SLOTS || !ids) return fail(TSTWO_ERR_BAD_ARG, "tstwo_init_devices: 1..%d devices", MAX_SLOTS);
    for (int i = 0; i < n; i++)
        for (int j = 0; j < i; j++)
            if (ids[i] == ids[j]) return fail(TSTWO_ERR_BAD_ARG, "device %d listed twice", ids[i]);
    if (g_slots) tstwo_shutdown();
    for (int i = 0; i < n; i++) {
        g_cur = i;
        int rc = init_slot(i, ids[i]);
        if (rc) { g_slots = i; tstwo_shutdown(); return rc; }
    }
    g_slots = n;
    for (int i = 0; i < n; i++) {
        TS_CUDA(cudaSetDevice(ids[i]));
        for (int j = 0; j < n; j++) {
            if (i == j) continue;
            int can = 0;
            TS_CUDA(cudaDeviceCanAccessPeer(&can, ids[i], ids[j]));
            if (!can) { tstwo_shutdown(); return fail(TSTWO_ERR_CUDA, "device %d cannot access device %d as a peer", ids[i], ids[j]); }
            cudaError_t e = cudaDeviceEnablePeerAccess(ids[j], 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
            else TS_CUDA(e);
        }
    }
    ctx_select(0);
    return 0;
}
int tstwo_device_count(void) { return g_slots; }
// the slot (not the CUDA ordinal) every following call of this thread works on
int tstwo_set_device(int slot)
{
    if (slot < 0 || slot >= g_slots || !g_ctxs[slot].ready) return fail(TSTWO_ERR_BAD_ARG, "device slot %d not initialised (%d slots)", slot, g_slots);
    ctx_select(slot);
    return 0;
}

int tstwo_shutdown(void)
{
    if (!g_slots && !g_ctxs[0].ready) return 0;
    multi_release();
    {
        std::lock_guard<std::mutex> lk(g_tree_mu);
        for (auto &kv : g_trees) {
            if (g_ctxs[kv.second->slot].ready) cudaSetDevice(g_ctxs[kv.second->slot].device);
            cudaFree(kv.second->tw);
            delete kv.second;
        }
        g_trees.clear();
        drop_spare_trees_locked();
    }
    scatter_release();
    for (int i = MAX_SLOTS - 1; i >= 0; i--) shutdown_slot(i);
    g_slots = 0;
    g_cur = 0;
    return 0;
}

const char *tstwo_last_error(void) { return g_err; }

int tstwo_set_stream(void *s)
{
    TS_REQUIRE_INIT();
    cudaStream_t next = s ? (cudaStream_t)s : ctx().own_stream;
    if (next != ctx().stream) {
        // the column allocator hands freed blocks out again without events: that is only safe while everything runs on ONE
        // stream, so work queued on the old stream (which may still touch cached blocks) is drained before the switch
        TS_CUDA(cudaStreamSynchronize(ctx().stream));
        ctx().stream = next;
    }
    return 0;
}
int tstwo_sync(void)
{
    TS_REQUIRE_INIT();
    TS_CUDA(cudaStreamSynchronize(ctx().stream));
    return 0;
}
int tstwo_device_info(int *sm_count, int *cc_major, int *cc_minor, uint64_t *total_mem)
{
    TS_REQUIRE_INIT();
    cudaDeviceProp prop;
    TS_CUDA(cudaGetDeviceProperties(&prop, ctx().device));
    if (sm_count) *sm_count = prop.multiProcessorCount;
    if (cc_major) *cc_major = prop.major;
    if (cc_minor) *cc_minor = prop.minor;
    if (total_mem) *total_mem = prop.totalGlobalMem;
    return 0;
}
uint64_t tstwo_launch_count(void) { return ctx().launches; }

// Column allocator.  The reference creates a fresh array for every result (Column.zeros / uninitialized in every fold, commit
// layer, interpolate ...): through cudaMalloc / cudaFree that costs a device synchronisation and ~0.1-1 ms per call, which at FRI's
// layer sizes is 50x the kernels (scripts/fri_bench.py).  Freed blocks are therefore kept in exact-size free lists and handed out
// again without touching the driver; a block is only ever reused by work enqueued later on the library stream, so no
// synchronisation is needed.  Blocks stay plain cudaMalloc memory (CUDA IPC export keeps working).
}  // extern "C"
namespace tstwo {
namespace {
struct BlockCache {                                                // one per device slot
    std::unordered_map<uint64_t, std::vector<void *>> free_blocks; // rounded size -> free blocks
    std::unordered_map<void *, uint64_t> block_size;               // every live or cached block -> rounded size
    uint64_t cached_bytes = 0;
};
BlockCache g_cache[MAX_SLOTS];
#define g_free_blocks (g_cache[ctx().slot].free_blocks)
#define g_block_size (g_cache[ctx().slot].block_size)
#define g_cached_bytes (g_cache[ctx().slot].cached_bytes)
constexpr uint64_t CACHE_CAP_BYTES = 24ull << 30;

void release_cached_blocks()   // caller holds g_alloc_mu
{
    cudaDeviceSynchronize();
    for (auto &kv : g_free_blocks)
        for (void *p : kv.second) { cudaFree(p); g_block_size.erase(p); }
    g_free_blocks.clear();
    g_cached_bytes = 0;
}
}  // namespace
}  // namespace tstwo
extern "C" {

int tstwo_malloc(uint64_t bytes, tstwo_dptr *out)
{
    TS_REQUIRE_INIT();
    if (!out) return fail(TSTWO_ERR_BAD_ARG, "null out");
    const uint64_t size = ((bytes ? bytes : 4) + 255) & ~255ull;
    std::lock_guard<std::mutex> lock(g_alloc_mu);
    auto it = g_free_blocks.find(size);
    if (it != g_free_blocks.end() && !it->second.empty()) {
        *out = (tstwo_dptr)it->second.back();
        it->second.pop_back();
        g_cached_bytes -= size;
        return 0;
    }
    void *p = nullptr;
    cudaError_t e = cudaMalloc(&p, size);
    if (e == cudaErrorMemoryAllocation) {          // give the cache back to the driver and try once more
        cudaGetLastError();
        release_cached_blocks();
        e = cudaMalloc(&p, size);
    }
    if (e == cudaErrorMemoryAllocation) { cudaGetLastError(); return fail(TSTWO_ERR_OOM, "cudaMalloc(%llu) out of memory", (unsigned long long)bytes); }
    TS_CUDA(e);
    g_block_size[p] = size;
    *out = (tstwo_dptr)p;
    return 0;
}
int tstwo_free(tstwo_dptr p)
{
    TS_REQUIRE_INIT();
    if (!p) return 0;
    std::lock_guard<std::mutex> lock(g_alloc_mu);
    auto it = g_block_size.find((void *)p);
    if (it != g_block_size.end() && g_cached_bytes + it->second <= CACHE_CAP_BYTES) {
        g_free_blocks[it->second].push_back((void *)p);
        g_cached_bytes += it->second;
        return 0;
    }
    if (it != g_block_size.end()) g_block_size.erase(it);
    TS_CUDA(cudaStreamSynchronize(ctx().stream));
    TS_CUDA(cudaFree((void *)p));
    return 0;
}
int tstwo_mem_trim(void)
{
    TS_REQUIRE_INIT();
    {
        std::lock_guard<std::mutex> lock(g_alloc_mu);
        release_cached_blocks();
    }
    TS_CUDA(cudaStreamSynchronize(ctx().stream));
    std::lock_guard<std::mutex> lk(g_tree_mu);
    drop_spare_trees_locked();
    return 0;
}
int tstwo_memset_zero(tstwo_dptr p, uint64_t bytes)
{
    TS_REQUIRE_INIT();
    TS_CUDA(cudaMemsetAsync((void *)p, 0, bytes, ctx().stream));
    return 0;
}
int tstwo_upload(tstwo_dptr dst, const void *host, uint64_t bytes)
{
    TS_REQUIRE_INIT();
    TS_CUDA(cudaMemcpyAsync((void *)dst, host, bytes, cudaMemcpyHostToDevice, ctx().stream));
    TS_CUDA(cudaStreamSynchronize(ctx().stream));   // host buffer is borrowed for the call only
    return 0;
}
int tstwo_download(void *host, tstwo_dptr src, uint64_t bytes)
{
    TS_REQUIRE_INIT();
    TS_CUDA(cudaMemcpyAsync(host, (const void *)src, bytes, cudaMemcpyDeviceToHost, ctx().stream));
    TS_CUDA(cudaStreamSynchronize(ctx().stream));
    return 0;
}
int tstwo_copy(tstwo_dptr dst, tstwo_dptr src, uint64_t bytes)
{
    TS_REQUIRE_INIT();
    TS_CUDA(cudaMemcpyAsync((void *)dst, (const void *)src, bytes, cudaMemcpyDeviceToDevice, ctx().stream));
    return 0;
}
// Multi-GPU exchange on the copy engines: the shards of a [n_shards][n_cols][rows] block (tstwo_lde_columns_sharded) go to their
// owners' leaf buffers by device-to-device copies on side streams, ordered after what the main stream holds now and running
// concurrently with what it gets next (the next column group's transforms) -- the SMs never wait on NVLink.
static cudaStream_t g_scatter_streams[4] = { nullptr, nullptr, nullptr, nullptr };
static cudaEvent_t g_scatter_fork = nullptr, g_scatter_done[4] = { nullptr, nullptr, nullptr, nullptr };
static bool g_scatter_pending = false;
int tstwo_scatter_shards(tstwo_dptr evals_sharded, int n_cols, uint64_t rows_per_shard, const tstwo_dptr *shard_bases, int n_shards,
                         int col_offset)
{
    TS_REQUIRE_INIT();
    if (!evals_sharded || n_cols <= 0 || !rows_per_shard || !shard_bases || n_shards <= 0 || col_offset < 0)
        return fail(TSTWO_ERR_BAD_ARG, "bad scatter arguments");
    Context &c = ctx();
    if (!g_scatter_fork) {
        TS_CUDA(cudaEventCreateWithFlags(&g_scatter_fork, cudaEventDisableTiming));
        for (int k = 0; k < 4; k++) {
            TS_CUDA(cudaStreamCreateWithFlags(&g_scatter_streams[k], cudaStreamNonBlocking));
            TS_CUDA(cudaEventCreateWithFlags(&g_scatter_done[k], cudaEventDisableTiming));
        }
    }
    TS_CUDA(cudaEventRecord(g_scatter_fork, c.stream));
    for (int k = 0; k < 4; k++) TS_CUDA(cudaStreamWaitEvent(g_scatter_streams[k], g_scatter_fork, 0));
    const uint64_t bytes = 4ull * n_cols * rows_per_shard;
    for (int d = 0; d < n_shards; d++) {
        if (!shard_bases[d]) return fail(TSTWO_ERR_BAD_ARG, "shard base %d is null", d);
        TS_CUDA(cudaMemcpyAsync((void *)(shard_bases[d] + 4ull * rows_per_shard * col_offset), (const void *)(evals_sharded + bytes * d), bytes,
                                cudaMemcpyDeviceToDevice, g_scatter_streams[d & 3]));
    }
    g_scatter_pending = true;
    return 0;
}
}  // extern "C"
void tstwo::scatter_release()
{
    for (int k = 0; k < 4; k++) {
        if (g_scatter_streams[k]) cudaStreamDestroy(g_scatter_streams[k]);
        if (g_scatter_done[k]) cudaEventDestroy(g_scatter_done[k]);
        g_scatter_streams[k] = nullptr; g_scatter_done[k] = nullptr;
    }
    if (g_scatter_fork) cudaEventDestroy(g_scatter_fork);
    g_scatter_fork = nullptr; g_scatter_pending = false;
}
extern "C" {
// the library's stream waits for every scatter copy issued so far
int tstwo_scatter_join(void)
{
    TS_REQUIRE_INIT();
    if (!g_scatter_pending) return 0;
    for (int k = 0; k < 4; k++) {
        TS_CUDA(cudaEventRecord(g_scatter_done[k], g_scatter_streams[k]));
        TS_CUDA(cudaStreamWaitEvent(ctx().stream, g_scatter_done[k], 0));
    }
    g_scatter_pending = false;
    return 0;
}
// gather for decommitments (vcs/prover.ts:32-109 reads childLayerHashes[j] and column[nodeIndex] one by one): the host
// walks the query tree (it depends on the queries only, not on data), then ONE launch collects every witness word.
__global__ void gather_words_kernel(const uint64_t *__restrict__ addrs, uint64_t total, int words, uint32_t *__restrict__ out)
{
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= total) return;
    out[t] = reinterpret_cast<const uint32_t *>(addrs[t / words])[t % words];
}
int tstwo_gather_words(const tstwo_dptr *addrs, int n, int words_each, uint32_t *host_out)
{
    TS_REQUIRE_INIT();
    if (n < 0 || words_each <= 0 || words_each > 64 || (n && (!addrs || !host_out))) return fail(TSTWO_ERR_BAD_ARG, "bad gather arguments");
    if (n == 0) return 0;
    for (int k = 0; k < n; k++)
        if (!addrs[k] || (addrs[k] & 3)) return fail(TSTWO_ERR_BAD_ARG, "gather address %d is null or not word aligned", k);
    Context &c = ctx();
    const uint64_t total = (uint64_t)n * words_each;
    uint64_t *d_addrs = nullptr;
    uint32_t *d_out = nullptr;
    TS_CUDA(cudaMallocAsync(&d_addrs, sizeof(uint64_t) * n, c.stream));
    TS_CUDA(cudaMallocAsync(&d_out, 4 * total, c.stream));
    TS_CUDA(cudaMemcpyAsync(d_addrs, addrs, sizeof(uint64_t) * n, cudaMemcpyHostToDevice, c.stream));
    TS_LAUNCH(gather_words_kernel, (unsigned)((total + 255) / 256), 256, 0, d_addrs, total, words_each, d_out);
    TS_CUDA(cudaMemcpyAsync(host_out, d_out, 4 * total, cudaMemcpyDeviceToHost, c.stream));
    TS_CUDA(cudaFreeAsync(d_addrs, c.stream));
    TS_CUDA(cudaFreeAsync(d_out, c.stream));
    TS_CUDA(cudaStreamSynchronize(c.stream));
    return 0;
}
// MerkleProver.decommit (vcs/prover.ts:32-109): the walk over the query tree on the host (nextDecommitmentNode,
// vcs/utils.ts:40-57).  Appends the device address of every witness hash (in order) and of every column value in node order
// (is_query: the value belongs to a queried node, else it is column witness).
static int decommit_walk(tstwo_dptr tree, int max_log, const tstwo_dptr *cols, const int *col_logs, int n_cols, const int *query_offsets,
                         const uint32_t *query_positions, std::vector<uint64_t> &hash_addrs, std::vector<uint64_t> &value_addrs,
                         std::vector<uint8_t> &value_is_query)
{
    for (int k = 0; k < n_cols; k++)
        if (col_logs[k] < 0 || col_logs[k] > max_log || !cols[k]) return fail(TSTWO_ERR_BAD_ARG, "column %d does not fit the tree", k);
    // columns by size, largest first, stable (Array.prototype.sort in the reference)
    std::vector<std::vector<int>> by_log(max_log + 1);
    for (int k = 0; k < n_cols; k++) by_log[col_logs[k]].push_back(k);
    std::vector<uint32_t> last, nxt;
    for (int log = max_log; log >= 0; log--) {
        const uint32_t *direct = query_positions + query_offsets[log];
        const size_t n_direct = (size_t)(query_offsets[log + 1] - query_offsets[log]);
        const bool has_child = log < max_log;
        const uint64_t child_base = tree + 32ull * ((2ull << log) - 1);   // layer log + 1
        size_t pi = 0, di = 0;
        nxt.clear();
        while (pi < last.size() || di < n_direct) {
            uint32_t node = 0xffffffffu;
            if (pi < last.size()) node = last[pi] >> 1;
            if (di < n_direct && direct[di] < node) node = direct[di];
            if (has_child)
                for (uint32_t ch = 2 * node; ch <= 2 * node + 1; ch++) {
                    if (pi < last.size() && last[pi] == ch) pi++;       // the verifier recomputes this child itself
                    else hash_addrs.push_back(child_base + 32ull * ch);
                }
            const bool queried = di < n_direct && direct[di] == node;
            if (queried) di++;
            if (node >> log) return fail(TSTWO_ERR_BAD_ARG, "MerkleProver.decommit: query %u is outside layer %d", node, log);
            for (int k : by_log[log]) { value_addrs.push_back(cols[k] + 4ull * node); value_is_query.push_back(queried); }
            nxt.push_back(node);
        }
        last.swap(nxt);
    }
    return 0;
}
// hashes (8 words each) then single words, gathered by two launches into `words` ([8 nh | nv]); one synchronisation
static int gather_hashes_and_words(const std::vector<uint64_t> &hash_addrs, const std::vector<uint64_t> &value_addrs, std::vector<uint32_t> &words)
{
    const size_t nh = hash_addrs.size(), nv = value_addrs.size();
    words.assign(8 * nh + nv, 0u);
    if (nh + nv == 0) return 0;
    Context &c = ctx();
    uint64_t *d_addrs = nullptr;
    uint32_t *d_out = nullptr;
    std::vector<uint64_t> addrs(hash_addrs);
    addrs.insert(addrs.end(), value_addrs.begin(), value_addrs.end());
    TS_CUDA(cudaMallocAsync(&d_addrs, sizeof(uint64_t) * addrs.size(), c.stream));
    TS_CUDA(cudaMallocAsync(&d_out, 4 * words.size(), c.stream));
    TS_CUDA(cudaMemcpyAsync(d_addrs, addrs.data(), sizeof(uint64_t) * addrs.size(), cudaMemcpyHostToDevice, c.stream));
    if (nh) TS_LAUNCH(gather_words_kernel, (unsigned)((8 * nh + 255) / 256), 256, 0, d_addrs, (uint64_t)8 * nh, 8, d_out);
    if (nv) TS_LAUNCH(gather_words_kernel, (unsigned)((nv + 255) / 256), 256, 0, d_addrs + nh, (uint64_t)nv, 1, d_out + 8 * nh);
    TS_CUDA(cudaMemcpyAsync(words.data(), d_out, 4 * words.size(), cudaMemcpyDeviceToHost, c.stream));
    TS_CUDA(cudaFreeAsync(d_addrs, c.stream));
    TS_CUDA(cudaFreeAsync(d_out, c.stream));
    TS_CUDA(cudaStreamSynchronize(c.stream));
    return 0;
}
int tstwo_merkle_decommit(tstwo_dptr tree, int max_log, const tstwo_dptr *cols, const int *col_logs, int n_cols, const int *query_offsets,
                          const uint32_t *query_positions, uint32_t *queried_values, uint8_t *hash_witness, uint32_t *column_witness,
                          uint64_t counts[3])
{
    TS_REQUIRE_INIT();
    if (!tree || max_log < 0 || max_log > 30 || n_cols < 0 || !query_offsets || !counts) return fail(TSTWO_ERR_BAD_ARG, "bad decommit arguments");
    std::vector<uint64_t> hash_addrs, value_addrs;
    std::vector<uint8_t> value_is_query;
    int rc = decommit_walk(tree, max_log, cols, col_logs, n_cols, query_offsets, query_positions, hash_addrs, value_addrs, value_is_query);
    if (rc) return rc;
    uint64_t n_q = 0;
    for (uint8_t q : value_is_query) n_q += q;
    counts[0] = n_q; counts[1] = hash_addrs.size(); counts[2] = value_addrs.size() - n_q;
    if (!queried_values && !hash_witness && !column_witness) return 0;      // sizing call
    if ((n_q && !queried_values) || (counts[1] && !hash_witness) || (counts[2] && !column_witness))
        return fail(TSTWO_ERR_BAD_ARG, "decommit output buffer missing");
    const size_t nh = hash_addrs.size(), nv = value_addrs.size();
    std::vector<uint32_t> words;
    if ((rc = gather_hashes_and_words(hash_addrs, value_addrs, words))) return rc;
    if (nh) memcpy(hash_witness, words.data(), 32 * nh);
    size_t qi = 0, wi = 0;
    for (size_t k = 0; k < nv; k++) {
        if (value_is_query[k]) queried_values[qi++] = words[8 * nh + k];
        else column_witness[wi++] = words[8 * nh + k];
    }
    return 0;
}

// FriProver.decommitOnQueries' loop over the inner layers (fri.ts:769-785 with FriInnerLayerProver.decommit :533-541) in ONE call:
// per layer the decommitment positions and FRI witness positions of the folded queries (computeDecommitmentPositionsAndWitnessEvals,
// fri.ts:346-384, fold step 1), the Merkle walk of that layer's tree, then ALL witness evaluations, witness hashes and column
// values of ALL layers are fetched by two gather launches and one synchronisation (23 layers of config 4 used to be 23 x 2 calls
// with a synchronisation each plus the host's per-layer bookkeeping).  Outputs are concatenated layer by layer; counts[l] =
// {FRI witness evaluations (4 words each), witness hashes, column-witness words} of layer l.
int tstwo_fri_decommit_layers(int n_layers, const tstwo_dptr *trees, const tstwo_dptr *evals, const int *logs, const uint32_t *queries,
                              int n_queries, uint32_t *fri_witness, uint64_t cap_fri_witness, uint8_t *hash_witness, uint64_t cap_hashes,
                              uint32_t *column_witness, uint64_t cap_column_witness, uint64_t *counts)
{
    TS_REQUIRE_INIT();
    if (n_layers < 0 || n_queries < 0 || !counts || (n_layers && (!trees || !evals || !logs)) || (n_queries && !queries))
        return fail(TSTWO_ERR_BAD_ARG, "bad fri_decommit_layers arguments");
    std::vector<uint32_t> q(queries, queries + n_queries), pos, folded;
    for (int k = 1; k < n_queries; k++)
        if (q[k] <= q[k - 1]) return fail(TSTWO_ERR_BAD_ARG, "fri_decommit_layers: queries must be sorted and distinct");
    std::vector<uint64_t> hash_addrs, value_addrs, wit_addrs;
    std::vector<uint8_t> value_is_query;
    std::vector<size_t> h_end((size_t)n_layers), v_end((size_t)n_layers), w_end((size_t)n_layers);
    std::vector<int> offsets;
    for (int l = 0; l < n_layers; l++) {
        const int log = logs[l];
        if (log < 1 || log > 30 || !trees[l]) return fail(TSTWO_ERR_BAD_ARG, "fri_decommit_layers: bad layer %d", l);
        if (!q.empty() && (q.back() >> log)) return fail(TSTWO_ERR_BAD_ARG, "fri_decommit_layers: query position outside layer %d", l);
        // every position of every folding coset (pairs) that holds a query; those that are not queries go into the FRI witness
        pos.clear();
        for (size_t i = 0; i < q.size();) {
            const uint32_t coset = q[i] >> 1;
            bool has[2] = { false, false };
            while (i < q.size() && (q[i] >> 1) == coset) { has[q[i] & 1] = true; i++; }
            for (uint32_t b = 0; b < 2; b++) {
                const uint32_t p = 2 * coset + b;
                pos.push_back(p);
                if (!has[b]) for (int k = 0; k < 4; k++) wit_addrs.push_back(evals[4 * l + k] + 4ull * p);
            }
        }
        offsets.assign((size_t)log + 2, 0);
        offsets[(size_t)log + 1] = (int)pos.size();
        int col_logs[4] = { log, log, log, log };
        int rc = decommit_walk(trees[l], log, evals + 4 * l, col_logs, 4, offsets.data(), pos.data(), hash_addrs, value_addrs, value_is_query);
        if (rc) return rc;
        h_end[l] = hash_addrs.size(); v_end[l] = value_addrs.size(); w_end[l] = wit_addrs.size();
        folded.clear();
        for (uint32_t p : q) if (folded.empty() || folded.back() != (p >> 1)) folded.push_back(p >> 1);
        q.swap(folded);
    }
    // counts + capacity checks before any device work
    uint64_t tot_w = 0, tot_h = 0, tot_c = 0;
    for (int l = 0; l < n_layers; l++) {
        const size_t v0 = l ? v_end[l - 1] : 0;
        uint64_t cw = 0;
        for (size_t k = v0; k < v_end[l]; k++) cw += !value_is_query[k];
        counts[3 * l] = (w_end[l] - (l ? w_end[l - 1] : 0)) / 4;
        counts[3 * l + 1] = h_end[l] - (l ? h_end[l - 1] : 0);
        counts[3 * l + 2] = cw;
        tot_w += counts[3 * l]; tot_h += counts[3 * l + 1]; tot_c += cw;
    }
    if (tot_w > cap_fri_witness || tot_h > cap_hashes || tot_c > cap_column_witness)
        return fail(TSTWO_ERR_BAD_ARG, "fri_decommit_layers: output capacity too small (need %llu evaluations, %llu hashes, %llu words)",
                    (unsigned long long)tot_w, (unsigned long long)tot_h, (unsigned long long)tot_c);
    if ((tot_w && !fri_witness) || (tot_h && !hash_witness) || (tot_c && !column_witness))
        return fail(TSTWO_ERR_BAD_ARG, "fri_decommit_layers: output buffer missing");
    // one gather: [hashes | witness evaluation words | column values]; only the column WITNESS words are kept of the last group
    std::vector<uint64_t> word_addrs(wit_addrs);
    std::vector<size_t> cw_index;
    for (size_t k = 0; k < value_addrs.size(); k++)
        if (!value_is_query[k]) { word_addrs.push_back(value_addrs[k]); cw_index.push_back(k); }
    std::vector<uint32_t> words;
    int rc = gather_hashes_and_words(hash_addrs, word_addrs, words);
    if (rc) return rc;
    const size_t nh = hash_addrs.size();
    if (nh) memcpy(hash_witness, words.data(), 32 * nh);
    if (!wit_addrs.empty()) memcpy(fri_witness, words.data() + 8 * nh, 4 * wit_addrs.size());
    if (tot_c) memcpy(column_witness, words.data() + 8 * nh + wit_addrs.size(), 4 * (size_t)tot_c);
    return 0;
}
int tstwo_bit_reverse(tstwo_dptr col, uint64_t len)
{
    TS_REQUIRE_INIT();
    if (len == 0 || (len & (len - 1))) return fail(TSTWO_ERR_NOT_POW2, "length is not power of two");
    int log_n = 0;
    while ((1ull << log_n) < len) log_n++;
    if (log_n > 31) return fail(TSTWO_ERR_BAD_ARG, "column too long");
    if (log_n >= 2 * BR_T)
        TS_LAUNCH(bit_reverse_tiled_kernel, 1u << (log_n - 2 * BR_T), 256, 0, (uint32_t *)col, log_n);
    else
        TS_LAUNCH(bit_reverse_kernel, (unsigned)((len + 255) / 256), 256, 0, (uint32_t *)col, log_n);
    return 0;
}

int tstwo_twiddles_precompute(uint32_t init, int log, tstwo_twiddles *out)
{
    TS_REQUIRE_INIT();
    if (!out || log < 0 || log > 30) return fail(TSTWO_ERR_BAD_ARG, "bad twiddle request (log %d)", log);
    {
        std::lock_guard<std::mutex> lk(g_tree_mu);
        for (size_t i = 0; i < g_spare_trees.size(); i++) {
            TwiddleTree *sp = g_spare_trees[i];
            if (sp->slot == ctx().slot && sp->root_init == (init & 0x7fffffffu) && sp->root_log == log) {
                g_spare_trees.erase(g_spare_trees.begin() + (long)i);
                uint64_t h = g_next_tree++;
                g_trees[h] = sp;
                *out = h;
                return 0;
            }
        }
    }
    uint64_t n = 1ull << log;
    uint32_t *buf = nullptr;
    cudaError_t e = cudaMalloc(&buf, n * 4 * 4);
    if (e == cudaErrorMemoryAllocation) { cudaGetLastError(); return fail(TSTWO_ERR_OOM, "twiddle tree out of memory"); }
    TS_CUDA(e);
    TwiddleTree *t = new TwiddleTree{ ctx().slot, init & 0x7fffffffu, log, buf, buf + n, buf + 2 * n, buf + 3 * n };
    TS_LAUNCH(twiddle_tree_kernel, (unsigned)((n + 127) / 128), 128, 0, t->tw, t->itw, t->tw2, t->itw2, t->root_init, log);
    std::lock_guard<std::mutex> lk(g_tree_mu);
    uint64_t h = g_next_tree++;
    g_trees[h] = t;
    *out = h;
    return 0;
}
int tstwo_twiddles_free(tstwo_twiddles h)
{
    TS_REQUIRE_INIT();
    TwiddleTree *t = get_tree(h);
    if (!t) return fail(TSTWO_ERR_BAD_ARG, "unknown twiddle tree handle");
    TwiddleTree *victim = nullptr;
    {
        std::lock_guard<std::mutex> lk(g_tree_mu);
        g_trees.erase(h);
        g_spare_trees.push_back(t);
        if (g_spare_trees.size() > TW_SPARE_MAX) { victim = g_spare_trees.front(); g_spare_trees.erase(g_spare_trees.begin()); }
    }
    if (victim) {
        const int back = ctx().slot;
        if (victim->slot != back) ctx_select(victim->slot);
        cudaStreamSynchronize(ctx().stream);           // kernels enqueued earlier may still read the tables
        cudaError_t e = cudaFree(victim->tw);
        delete victim;
        if (ctx().slot != back) ctx_select(back);
        TS_CUDA(e);
    }
    return 0;
}
int tstwo_twiddles_info(tstwo_twiddles h, uint32_t *root_init, int *root_log, tstwo_dptr *tw, tstwo_dptr *itw)
{
    TS_REQUIRE_INIT();
    TwiddleTree *t = get_tree(h);
    if (!t) return fail(TSTWO_ERR_BAD_ARG, "unknown twiddle tree handle");
    if (root_init) *root_init = t->root_init;
    if (root_log) *root_log = t->root_log;
    if (tw) *tw = (tstwo_dptr)t->tw;
    if (itw) *itw = (tstwo_dptr)t->itw;
    return 0;
}

}  // extern "C"
