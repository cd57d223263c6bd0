// The metric path in one call: Stwo's CommitmentTreeProver::new as quoted in the reference at
// packages/core/src/pcs/prover.ts:215-235 (interpolate -> evaluate on the blown-up canonic domain -> Merkle commit),
// i.e. PolyOps.interpolateColumns + evaluatePolynomials (poly/circle/ops.ts:73-101) + MerkleProver.commit
// (vcs/prover.ts:13-30) on device-resident columns.
#include <cstdlib>
#include <cstring>
#include <vector>

#include "common.cuh"
#include "host_m31.h"

using namespace tstwo;

namespace {
// scratch for coefficients the caller does not want to keep (they may not alias the evals they are extended into:
// the coset passes of different CTAs read the same coefficients)
// (per device: Context::coeff_scratch)
#define g_coeff_scratch (ctx().coeff_scratch)
#define g_coeff_scratch_words (ctx().coeff_scratch_words)
int coeff_scratch(size_t words, uint32_t **out)
{
    if (words > g_coeff_scratch_words) {
        if (g_coeff_scratch) { TS_CUDA(cudaStreamSynchronize(ctx().stream)); TS_CUDA(cudaFree(g_coeff_scratch)); g_coeff_scratch = nullptr; g_coeff_scratch_words = 0; }
        cudaError_t e = cudaMalloc(&g_coeff_scratch, words * 4);
        if (e == cudaErrorMemoryAllocation) { cudaGetLastError(); return fail(TSTWO_ERR_OOM, "coefficient scratch out of memory"); }
        TS_CUDA(e);
        g_coeff_scratch_words = words;
    }
    *out = g_coeff_scratch;
    return 0;
}

int lde_commit_device(const uint32_t *trace, int n_cols, int log_size, int log_blowup, TwiddleTree *t, uint32_t *coeffs,
                      uint32_t *evals, uint8_t *layers, uint8_t *root, uint32_t flags, bool sync_root)
{
    const int ext = log_size + log_blowup;
    const uint64_t n = 1ull << log_size, ne = 1ull << ext;
    int rc;
    if (!coeffs && log_blowup > 0 && (rc = coeff_scratch((size_t)n_cols * n, &coeffs))) return rc;
    ColSet src = { trace, n, nullptr };
    ColSet cf = coeffs ? ColSet{ coeffs, n, nullptr } : ColSet{ evals, ne, nullptr };   // blowup 0: in place in evals
    ColSet ev = { evals, ne, nullptr };
    Span commitment("Commitment");
    {
        Span s("Interpolation for commitment");
        rc = fft_interpolate(src, cf, n_cols, log_size, t, flags);
    }
    if (rc) return rc;
    {
        Span s("Extension");
        rc = fft_evaluate(cf, log_size, ev, n_cols, ext, t, flags);
    }
    if (rc) return rc;
    if (layers) {
        Span s("Merkle");
        const uint8_t *prev = nullptr;
        ColSet none = { nullptr, 0, nullptr };
        for (int log = ext; log >= 0; log--) {
            if (log < ext && log <= 9 && merkle_tail(layers, log) == log) break;   // the bottom layers in one launch
            uint8_t *dst = layers + 32ull * ((1ull << log) - 1);
            rc = merkle_commit_layer(log, prev, log == ext ? ev : none, log == ext ? n_cols : 0, dst);
            if (rc) return rc;
            prev = dst;
        }
        if (root) {
            TS_CUDA(cudaMemcpyAsync(root, layers, 32, cudaMemcpyDeviceToHost, ctx().stream));
            if (sync_root) TS_CUDA(cudaStreamSynchronize(ctx().stream));
        }
    }
    return 0;
}

int check_args(int n_cols, int log_size, int log_blowup, tstwo_twiddles tw, TwiddleTree **t)
{
    if (n_cols < 1 || log_size < 1 || log_blowup < 0 || log_size + log_blowup > 30)
        return fail(TSTWO_ERR_BAD_ARG, "bad lde_commit arguments");
    *t = get_tree(tw);
    if (!*t) return fail(TSTWO_ERR_BAD_ARG, "unknown twiddle tree handle");
    const int ext = log_size + log_blowup;
    // CanonicCoset(k).half_coset() = half_odds(k-1): initial index 2^(31-(k-1)-2) (poly/circle/canonic.ts, circle.ts:232-234)
    if (!is_doubling_of(host::subgroup_gen(ext - 1 + 2), ext - 1, *t) ||
        !is_doubling_of(host::subgroup_gen(log_size - 1 + 2), log_size - 1, *t))
        return fail(TSTWO_ERR_TWIDDLE_MISMATCH, "twiddle tree mismatch");
    return 0;
}
}  // namespace

extern "C" {

int tstwo_lde_commit(tstwo_dptr trace, int n_cols, int log_size, int log_blowup, tstwo_twiddles tw, tstwo_dptr coeffs,
                     tstwo_dptr evals, tstwo_dptr layers, uint8_t root[32], uint32_t flags)
{
    TS_REQUIRE_INIT();
    TwiddleTree *t;
    int rc = check_args(n_cols, log_size, log_blowup, tw, &t);
    if (rc) return rc;
    if (!trace || !evals) return fail(TSTWO_ERR_BAD_ARG, "null trace/evals");
    return lde_commit_device((const uint32_t *)trace, n_cols, log_size, log_blowup, t, (uint32_t *)coeffs, (uint32_t *)evals,
                             (uint8_t *)layers, root, flags, true);
}

// Column-sharded LDE for the multi-GPU path: this rank's n_cols columns are interpolated and extended, and the
// evaluations are written grouped by destination rank, [n_shards][n_cols][2^(log_size+log_blowup)/n_shards], so that
// one all-to-all (NCCL or peer stores) hands every rank all columns of its contiguous row range -- exactly the
// column-major block the leaf-hash kernel reads.
int tstwo_lde_columns_sharded(tstwo_dptr trace, int n_cols, int log_size, int log_blowup, tstwo_twiddles tw, tstwo_dptr coeffs,
                              tstwo_dptr evals, int n_shards, uint32_t flags)
{
    TS_REQUIRE_INIT();
    TwiddleTree *t;
    int rc = check_args(n_cols, log_size, log_blowup, tw, &t);
    if (rc) return rc;
    const int ext = log_size + log_blowup;
    int sl = 0;
    while ((1 << sl) < n_shards) sl++;
    if (n_shards < 1 || (1 << sl) != n_shards || ext - sl < 12) return fail(TSTWO_ERR_BAD_ARG, "bad shard count %d", n_shards);
    if (!trace || !evals) return fail(TSTWO_ERR_BAD_ARG, "null trace/evals");
    const uint64_t n = 1ull << log_size, rows = 1ull << (ext - sl);
    // scratch: [coefficients (unless kept by the caller)] [plain evaluations the strided passes work in]
    const uint64_t ne = 1ull << ext;
    uint32_t *scr;
    if ((rc = coeff_scratch((size_t)n_cols * (n + ne), &scr))) return rc;
    uint32_t *cf = coeffs ? (uint32_t *)coeffs : scr;
    ColSet work = { scr + (size_t)n_cols * n, ne, nullptr };
    ColSet src = { (const uint32_t *)trace, n, nullptr };
    ColSet cfs = { cf, n, nullptr };
    ColSet ev = { (const uint32_t *)evals, rows, nullptr };
    ev.shard_log = ext - sl;
    ev.shard_adj = (uint64_t)n_cols * rows - rows;
    if ((rc = fft_interpolate(src, cfs, n_cols, log_size, t, flags))) return rc;
    return fft_evaluate(cfs, log_size, ev, n_cols, ext, t, flags, &work);
}

}  // extern "C" (reopened below)

// Multi-GPU, fused exchange: like tstwo_lde_columns_sharded, but the last FFT pass stores the rows of shard d straight
// into shard_bases[d] -- the leaf buffer [total columns][rows] of rank d, mapped into this process with
// tstwo_ipc_import -- at column col_offset + c.  No send buffer, no all-to-all: the stores travel over NVLink while the
// butterflies of the next tiles run.  The caller orders the step with one (tiny) collective before the receivers hash.
//
// col_groups > 1 (tstwo_lde_columns_to_peers_overlapped): the NVLink-bound pass of column group k runs on a side stream
// with at most peer_sms CTAs while the other SMs interpolate and extend group k + 1 on the library's stream -- the
// exchange (7/8 of the evaluations at 8 GPUs, >= 1 ms at 900 GB/s) then overlaps ALL of the LDE's arithmetic, not just the
// 14 layers of the pass that issues the stores.  Same bytes, same destinations: results are identical for any grouping.
namespace {
int g_overlap_taper = 100;     // percent of an equal share given to the first and the last column group (tstwo_debug_set_overlap_taper)
// per device (Context): peer_stream; peer_ev[0] group ready for its last pass (main -> side), [1] side drained
#define g_peer_stream (ctx().peer_stream)
#define g_peer_ev (ctx().peer_ev)

// Column-interleaved ownership (tstwo_lde_columns_to_peers_halves): the first half of the column groups goes to columns col_offset..,
// the second half to col_offset_hi.., the side stream publishes `value` to half_flags once the first half's stores are complete and
// to arrive_flags once all are, the last group's storing pass keeps to peer_sms SMs and the side stream is NOT joined (the caller
// hashes the first half of the columns beside it and then calls tstwo_peer_join).
struct Halves {
    int col_offset_hi = -1;
    const tstwo_dptr *half_flags = nullptr, *arrive_flags = nullptr;
    int n_flags = 0;
    uint32_t value = 0;
};
int signal_flags(const tstwo_dptr *flag_addrs, int n, uint32_t value);   // below: the kernel of tstwo_peer_signal on ctx().stream

int lde_to_peers(tstwo_dptr trace, int n_cols, int log_size, int log_blowup, tstwo_twiddles tw, tstwo_dptr coeffs,
                 tstwo_dptr work, const tstwo_dptr *shard_bases, int n_shards, int col_offset, uint32_t flags,
                 int col_groups, int peer_sms, const Halves *hv = nullptr)
{
    TwiddleTree *t;
    int rc = check_args(n_cols, log_size, log_blowup, tw, &t);
    if (rc) return rc;
    const int ext = log_size + log_blowup;
    int sl = 0;
    while ((1 << sl) < n_shards) sl++;
    if (n_shards < 1 || n_shards > 64 || (1 << sl) != n_shards || ext - sl < 14) return fail(TSTWO_ERR_BAD_ARG, "bad shard count %d", n_shards);
    if (!trace || !work || !shard_bases || col_offset < 0) return fail(TSTWO_ERR_BAD_ARG, "null trace/work/shard bases");
    Context &c = ctx();
    if (col_groups < 1 || peer_sms < 0 || peer_sms >= c.sm_count) return fail(TSTWO_ERR_BAD_ARG, "bad overlap parameters (%d groups, %d SMs)", col_groups, peer_sms);
    if (log_size < 15 || peer_sms == 0) col_groups = 1;          // a one-pass evaluation has nothing to split off
    if (col_groups > n_cols) col_groups = n_cols;
    if (col_groups > 32) col_groups = 32;
    if (hv && (col_groups < 2 || (col_groups & 1) || n_cols % col_groups || hv->col_offset_hi < 0 || hv->n_flags < 1 || hv->n_flags > 64 ||
               !hv->half_flags || !hv->arrive_flags))
        return fail(TSTWO_ERR_BAD_ARG, "column halves need an even number of equal column groups on a side stream (%d columns, %d groups, %d SMs) and flags",
                    n_cols, col_groups, peer_sms);
    const uint64_t n = 1ull << log_size, ne = 1ull << ext, rows = 1ull << (ext - sl);
    uint32_t *scr = nullptr;
    if (!coeffs && (rc = coeff_scratch((size_t)n_cols * n, &scr))) return rc;
    uint32_t *cf = coeffs ? (uint32_t *)coeffs : scr;
    for (int d = 0; d < n_shards; d++)
        if (!shard_bases[d] || (shard_bases[d] & 15)) return fail(TSTWO_ERR_BAD_ARG, "shard base %d null or not 16-byte aligned", d);
    if (col_groups > 1 && !g_peer_stream) {
        if (const char *e = getenv("TSTWO_OVERLAP_TAPER")) { const int v = atoi(e); if (v >= 1 && v <= 100) g_overlap_taper = v; }
        int lo_pri = 0, hi_pri = 0;
        TS_CUDA(cudaDeviceGetStreamPriorityRange(&lo_pri, &hi_pri));
        TS_CUDA(cudaStreamCreateWithPriority(&g_peer_stream, cudaStreamNonBlocking, hi_pri));
        for (auto &e : g_peer_ev) TS_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    }
    cudaStream_t main_stream = c.stream;
    // per-shard bases, advanced to each group's first column, in the library's device pointer table (slot 2, 64 per group)
    uint64_t host_bases[32 * 64];
    // Column groups.  Equal groups leave the first group's main passes and the last group's storing pass without a partner; with
    // g_overlap_taper < 100 the first and the last group get that percentage of an equal share (the middle ones the rest), and
    // every phase splits the SMs in proportion to the work of its two sides, so head and tail shrink while the phases stay balanced.
    int group_cols[32];
    {
        int edge = col_groups >= 3 ? (int)(((long long)n_cols * g_overlap_taper + 50LL * col_groups) / (100LL * col_groups)) : 0;
        if (col_groups >= 3 && edge < 1) edge = 1;
        if (col_groups < 3 || g_overlap_taper >= 100 || n_cols - 2 * edge < col_groups - 2 || hv) edge = 0;
        int left = n_cols - 2 * edge, c = 0;
        for (int k = 0; k < col_groups; k++) {
            if (edge && (k == 0 || k == col_groups - 1)) group_cols[k] = edge;
            else {
                const int mid_left = edge ? col_groups - 2 - (k - 1) : col_groups - k;
                group_cols[k] = (left + mid_left - 1) / mid_left;
                left -= group_cols[k];
            }
            c += group_cols[k];
        }
        if (c != n_cols) return fail(TSTWO_ERR_BAD_ARG, "internal: column groups do not add up (%d of %d)", c, n_cols);
    }
    int c0 = 0;
    for (int k = 0; k < col_groups; k++) {
        const int nc = group_cols[k];
        const int dst_col = (hv && k >= col_groups / 2) ? hv->col_offset_hi + (c0 - n_cols / 2) : col_offset + c0;
        for (int d = 0; d < n_shards; d++) host_bases[64 * k + d] = shard_bases[d] + (uint64_t)dst_col * rows * 4;
        c0 += nc;
    }
    // SMs of group k's storing pass while group k + 1's other passes run beside it: peer_sms is the storing pass's share of the
    // LDE's integer work for EQUAL groups; for unequal neighbours the split follows their column counts
    auto side_sms = [&](int k) -> int {
        if (k >= col_groups - 1) return hv ? peer_sms : 0;          // halves: the caller hashes beside the last storing pass
        const double ws = (double)peer_sms * group_cols[k], wm = (double)(c.sm_count - peer_sms) * group_cols[k + 1];
        int p = (int)(c.sm_count * ws / (ws + wm) + 0.5);
        if (p < 8) p = 8;
        if (p > c.sm_count - 8) p = c.sm_count - 8;
        return p;
    };
    uint64_t *dev_bases = c.d_ptr_table + 2 * 4096;
    TS_CUDA(cudaMemcpyAsync(dev_bases, host_bases, sizeof(uint64_t) * 64 * (size_t)col_groups, cudaMemcpyHostToDevice, main_stream));
    c0 = 0;
    for (int k = 0; k < col_groups && !rc; k++) {
        const int nc = group_cols[k];
        ColSet src = { (const uint32_t *)trace + (uint64_t)c0 * n, n, nullptr };
        ColSet cfs = { cf + (uint64_t)c0 * n, n, nullptr };
        ColSet wk = { (const uint32_t *)work + (uint64_t)c0 * ne, ne, nullptr };
        ColSet ev = { nullptr, rows, nullptr };
        ev.shard_log = ext - sl;
        ev.shard_bases = dev_bases + 64 * k;
        if (col_groups == 1) {
            { Span s("Interpolation for commitment"); rc = fft_interpolate(src, cfs, nc, log_size, t, flags); }
            if (rc) break;
            Span s("Extension");
            rc = fft_evaluate(cfs, log_size, ev, nc, ext, t, flags, &wk);
            break;
        }
        // every pass but the last on the library's stream; while an earlier group's last pass may be running they leave it peer_sms SMs
        c.grid_limit = k > 0 ? c.sm_count - side_sms(k - 1) : 0;
        if (!(rc = fft_interpolate(src, cfs, nc, log_size, t, flags))) rc = fft_evaluate(cfs, log_size, ev, nc, ext, t, flags, &wk, 1);
        if (rc) break;
        cudaError_t e = cudaEventRecord(g_peer_ev[0], main_stream);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(g_peer_stream, g_peer_ev[0], 0);
        if (e != cudaSuccess) { rc = fail(TSTWO_ERR_CUDA, "event record/wait failed: %s", cudaGetErrorString(e)); break; }
        // the pass whose stores cross NVLink, on the side stream; the last group's has the GPU to itself
        c.stream = g_peer_stream;
        c.grid_limit = side_sms(k);
        rc = fft_evaluate(cfs, log_size, ev, nc, ext, t, flags, &wk, 2);
        if (!rc && hv && k == col_groups / 2 - 1) rc = signal_flags(hv->half_flags, hv->n_flags, hv->value);
        if (!rc && hv && k == col_groups - 1) rc = signal_flags(hv->arrive_flags, hv->n_flags, hv->value);
        c.stream = main_stream;
        c0 += nc;
    }
    c.stream = main_stream;
    c.grid_limit = 0;
    if (col_groups > 1 && (!hv || rc)) {                          // join, also on the error path: the side stream may hold work
        cudaError_t e = cudaEventRecord(g_peer_ev[1], g_peer_stream);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(main_stream, g_peer_ev[1], 0);
        if (e != cudaSuccess && !rc) rc = fail(TSTWO_ERR_CUDA, "event record/wait failed: %s", cudaGetErrorString(e));
    }
    return rc;
}
}  // namespace

namespace tstwo {
void pipeline_release()
{
    if (g_coeff_scratch) cudaFree(g_coeff_scratch);
    g_coeff_scratch = nullptr; g_coeff_scratch_words = 0;
    if (g_peer_stream) cudaStreamDestroy(g_peer_stream);
    g_peer_stream = nullptr;
    for (auto &e : g_peer_ev) { if (e) cudaEventDestroy(e); e = nullptr; }
}
}  // namespace tstwo

extern "C" {
int tstwo_debug_set_overlap_taper(int percent)
{
    if (percent < 1 || percent > 100) return fail(TSTWO_ERR_BAD_ARG, "taper must be 1..100 percent");
    g_overlap_taper = percent;
    return 0;
}
int tstwo_lde_columns_to_peers(tstwo_dptr trace, int n_cols, int log_size, int log_blowup, tstwo_twiddles tw, tstwo_dptr coeffs,
                               tstwo_dptr work, const tstwo_dptr *shard_bases, int n_shards, int col_offset, uint32_t flags)
{
    TS_REQUIRE_INIT();
    return lde_to_peers(trace, n_cols, log_size, log_blowup, tw, coeffs, work, shard_bases, n_shards, col_offset, flags, 1, 0);
}
int tstwo_lde_columns_to_peers_overlapped(tstwo_dptr trace, int n_cols, int log_size, int log_blowup, tstwo_twiddles tw,
                                          tstwo_dptr coeffs, tstwo_dptr work, const tstwo_dptr *shard_bases, int n_shards,
                                          int col_offset, uint32_t flags, int col_groups, int peer_sms)
{
    TS_REQUIRE_INIT();
    return lde_to_peers(trace, n_cols, log_size, log_blowup, tw, coeffs, work, shard_bases, n_shards, col_offset, flags,
                        col_groups, peer_sms);
}
int tstwo_lde_columns_to_peers_halves(tstwo_dptr trace, int n_cols, int log_size, int log_blowup, tstwo_twiddles tw,
                                      tstwo_dptr coeffs, tstwo_dptr work, const tstwo_dptr *shard_bases, int n_shards,
                                      int col_offset_lo, int col_offset_hi, uint32_t flags, int col_groups, int peer_sms,
                                      const tstwo_dptr *half_flag_addrs, const tstwo_dptr *arrive_flag_addrs, int n_flags, uint32_t value)
{
    TS_REQUIRE_INIT();
    Halves hv;
    hv.col_offset_hi = col_offset_hi;
    hv.half_flags = half_flag_addrs; hv.arrive_flags = arrive_flag_addrs; hv.n_flags = n_flags; hv.value = value;
    return lde_to_peers(trace, n_cols, log_size, log_blowup, tw, coeffs, work, shard_bases, n_shards, col_offset_lo, flags,
                        col_groups, peer_sms, &hv);
}
int tstwo_peer_join(void)
{
    TS_REQUIRE_INIT();
    Context &c = ctx();
    if (!g_peer_stream) return 0;                                 // nothing was ever put on a side stream
    TS_CUDA(cudaEventRecord(g_peer_ev[1], g_peer_stream));
    TS_CUDA(cudaStreamWaitEvent(c.stream, g_peer_ev[1], 0));
    return 0;
}
}

extern "C" {
// CUDA IPC plumbing for the above (one process per GPU): export a tstwo_malloc'ed buffer, import a peer's.
int tstwo_ipc_export(tstwo_dptr ptr, uint8_t handle[64])
{
    TS_REQUIRE_INIT();
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
    cudaIpcMemHandle_t h;
    TS_CUDA(cudaIpcGetMemHandle(&h, (void *)ptr));
    memcpy(handle, &h, 64);
    return 0;
}
int tstwo_ipc_import(const uint8_t handle[64], tstwo_dptr *out)
{
    TS_REQUIRE_INIT();
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, 64);
    void *p = nullptr;
    TS_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    *out = (tstwo_dptr)p;
    return 0;
}
int tstwo_ipc_close(tstwo_dptr ptr)
{
    TS_REQUIRE_INIT();
    TS_CUDA(cudaIpcCloseMemHandle((void *)ptr));
    return 0;
}

// ---- cross-device flags (see the header).  A kernel boundary orders the flag store after every store of the kernels before it
// on the stream (the peer stores of the last FFT pass included); fence.sys + st.release.sys publish it to the peer, whose
// polling kernel acquires it before the hashing kernels that follow it on ITS stream start.
namespace {
struct FlagAddrs { uint32_t *a[64]; };
__global__ void peer_signal_kernel(FlagAddrs fa, int n, uint32_t value)
{
    if ((int)threadIdx.x < n) {
        __threadfence_system();
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(fa.a[threadIdx.x]), "r"(value) : "memory");
    }
}
__global__ void peer_wait_kernel(const uint32_t *flags, int n, uint32_t value, unsigned long long timeout_ns)
{
    if ((int)threadIdx.x < n) {
        unsigned long long t0;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        for (;;) {
            uint32_t v;
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flags + threadIdx.x) : "memory");
            if ((int32_t)(v - value) >= 0) break;
            __nanosleep(200);
            unsigned long long t1;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
            if (t1 - t0 > timeout_ns) __trap();          // a peer that never arrives is an error, not a hung device
        }
    }
    __syncwarp();
    __threadfence_system();
}
__global__ void peer_put_kernel(uint32_t *dst, const uint32_t *src, int n_words, uint32_t *flag, uint32_t value)
{
    for (int i = threadIdx.x; i < n_words; i += blockDim.x) dst[i] = src[i];
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(flag), "r"(value) : "memory");
}
}  // namespace

extern "C++" {
namespace {
int signal_flags(const tstwo_dptr *flag_addrs, int n, uint32_t value)
{
    if (!flag_addrs || n < 1 || n > 64) return fail(TSTWO_ERR_BAD_ARG, "tstwo_peer_signal: 1..64 flag addresses");
    FlagAddrs fa;
    memset(&fa, 0, sizeof fa);
    for (int i = 0; i < n; i++) {
        if (!flag_addrs[i] || (flag_addrs[i] & 3)) return fail(TSTWO_ERR_BAD_ARG, "tstwo_peer_signal: flag %d null or unaligned", i);
        fa.a[i] = (uint32_t *)flag_addrs[i];
    }
    TS_LAUNCH(peer_signal_kernel, 1, 64, 0, fa, n, value);
    return 0;
}
}  // namespace
}  // extern "C++"
int tstwo_peer_signal(const tstwo_dptr *flag_addrs, int n, uint32_t value)
{
    TS_REQUIRE_INIT();
    return signal_flags(flag_addrs, n, value);
}
int tstwo_peer_wait(tstwo_dptr flags, int n, uint32_t value, uint32_t timeout_ms)
{
    TS_REQUIRE_INIT();
    if (!flags || (flags & 3) || n < 1 || n > 64) return fail(TSTWO_ERR_BAD_ARG, "tstwo_peer_wait: 1..64 aligned flags");
    TS_LAUNCH(peer_wait_kernel, 1, 64, 0, (const uint32_t *)flags, n, value, (unsigned long long)(timeout_ms ? timeout_ms : 4000) * 1000000ull);
    return 0;
}
int tstwo_peer_put(tstwo_dptr dst, tstwo_dptr src, int n_words, tstwo_dptr flag_addr, uint32_t value)
{
    TS_REQUIRE_INIT();
    if (!dst || !src || !flag_addr || ((dst | src | flag_addr) & 3) || n_words < 1 || n_words > 1024)
        return fail(TSTWO_ERR_BAD_ARG, "tstwo_peer_put: bad arguments");
    TS_LAUNCH(peer_put_kernel, 1, 64, 0, (uint32_t *)dst, (const uint32_t *)src, n_words, (uint32_t *)flag_addr, value);
    return 0;
}

// Host-resident trace: the columns are copied in chunks on a copy stream while the previous chunk's interpolate /
// evaluate kernels run; the Merkle commit runs once all columns are extended.
int tstwo_lde_commit_host(const uint32_t *host_trace, int n_cols, int log_size, int log_blowup, tstwo_twiddles tw,
                          tstwo_dptr trace_staging, tstwo_dptr coeffs, tstwo_dptr evals, tstwo_dptr layers,
                          uint8_t root[32], uint32_t flags)
{
    TS_REQUIRE_INIT();
    TwiddleTree *t;
    int rc = check_args(n_cols, log_size, log_blowup, tw, &t);
    if (rc) return rc;
    if (!host_trace || !trace_staging || !evals) return fail(TSTWO_ERR_BAD_ARG, "null trace/staging/evals");
    Context &c = ctx();
    const int ext = log_size + log_blowup;
    const uint64_t n = 1ull << log_size, ne = 1ull << ext;
    const int chunk = 16;   // columns per H2D chunk = one 64-byte BLAKE2s block of every leaf
    const int n_chunks = (n_cols + chunk - 1) / chunk;
    std::vector<cudaEvent_t> ev((size_t)n_chunks);
    uint32_t *stage = (uint32_t *)trace_staging;
    for (int k = 0; k < n_chunks; k++) {
        const int c0 = k * chunk, nc = (n_cols - c0 < chunk) ? n_cols - c0 : chunk;
        TS_CUDA(cudaEventCreateWithFlags(&ev[k], cudaEventDisableTiming));
        TS_CUDA(cudaMemcpyAsync(stage + (uint64_t)c0 * n, host_trace + (uint64_t)c0 * n, (uint64_t)nc * n * 4,
                                cudaMemcpyHostToDevice, c.copy_stream));
        TS_CUDA(cudaEventRecord(ev[k], c.copy_stream));
    }
    for (int k = 0; k < n_chunks; k++) {
        const int c0 = k * chunk, nc = (n_cols - c0 < chunk) ? n_cols - c0 : chunk;
        TS_CUDA(cudaStreamWaitEvent(c.stream, ev[k], 0));
        uint32_t *cf = coeffs ? (uint32_t *)coeffs + (uint64_t)c0 * n : nullptr;
        rc = lde_commit_device(stage + (uint64_t)c0 * n, nc, log_size, log_blowup, t, cf, (uint32_t *)evals + (uint64_t)c0 * ne,
                               nullptr, nullptr, flags, false);
        if (rc) return rc;
        if (layers) {
            // the leaf hashes absorb this chunk's 16 columns now (one BLAKE2s block per row), so that only the last
            // chunk's block and the node layers are left once the last column has crossed PCIe
            ColSet evs = { (const uint32_t *)evals + (uint64_t)c0 * ne, ne, nullptr };
            rc = merkle_leaf_absorb(ext, evs, c0, nc, n_cols, (uint8_t *)layers + 32ull * ((1ull << ext) - 1));
            if (rc) return rc;
        }
    }
    if (layers) {
        const uint8_t *prev = (const uint8_t *)layers + 32ull * ((1ull << ext) - 1);
        ColSet none = { nullptr, 0, nullptr };
        for (int log = ext - 1; log >= 0; log--) {
            if (log <= 9 && merkle_tail((uint8_t *)layers, log) == log) break;
            uint8_t *dst = (uint8_t *)layers + 32ull * ((1ull << log) - 1);
            rc = merkle_commit_layer(log, prev, none, 0, dst);
            if (rc) return rc;
            prev = dst;
        }
        if (root) TS_CUDA(cudaMemcpyAsync(root, (const void *)layers, 32, cudaMemcpyDeviceToHost, c.stream));
    }
    TS_CUDA(cudaStreamSynchronize(c.stream));
    for (auto e : ev) cudaEventDestroy(e);
    return 0;
}

}  // extern "C"
