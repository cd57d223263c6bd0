// The metric path in one call: Stwo's CommitmentTreeProver::new as quoted in the reference at
// packages/core/src/pcs/prover.ts:215-235 (interpolate -> evaluate on the blown-up canonic domain -> Merkle commit),
// i.e. PolyOps.interpolateColumns + evaluatePolynomials (poly/circle/ops.ts:73-101) + MerkleProver.commit
// (vcs/prover.ts:13-30) on device-resident columns.
#include <cstring>
#include <vector>

#include "common.cuh"
#include "host_m31.h"

using namespace tstwo;

namespace {
// scratch for coefficients the caller does not want to keep (they may not alias the evals they are extended into:
// the coset passes of different CTAs read the same coefficients)
uint32_t *g_coeff_scratch = nullptr;
size_t g_coeff_scratch_words = 0;
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
    rc = fft_interpolate(src, cf, n_cols, log_size, t, flags);
    if (rc) return rc;
    rc = fft_evaluate(cf, log_size, ev, n_cols, ext, t, flags);
    if (rc) return rc;
    if (layers) {
        const uint8_t *prev = nullptr;
        ColSet none = { nullptr, 0, nullptr };
        for (int log = ext; log >= 0; log--) {
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

// Multi-GPU, fused exchange: like tstwo_lde_columns_sharded, but the last FFT pass stores the rows of shard d straight
// into shard_bases[d] -- the leaf buffer [total columns][rows] of rank d, mapped into this process with
// tstwo_ipc_import -- at column col_offset + c.  No send buffer, no all-to-all: the stores travel over NVLink while the
// butterflies of the next tiles run.  The caller orders the step with one (tiny) collective before the receivers hash.
int tstwo_lde_columns_to_peers(tstwo_dptr trace, int n_cols, int log_size, int log_blowup, tstwo_twiddles tw, tstwo_dptr coeffs,
                               tstwo_dptr work, const tstwo_dptr *shard_bases, int n_shards, int col_offset, uint32_t flags)
{
    TS_REQUIRE_INIT();
    TwiddleTree *t;
    int rc = check_args(n_cols, log_size, log_blowup, tw, &t);
    if (rc) return rc;
    const int ext = log_size + log_blowup;
    int sl = 0;
    while ((1 << sl) < n_shards) sl++;
    if (n_shards < 1 || n_shards > 64 || (1 << sl) != n_shards || ext - sl < 14) return fail(TSTWO_ERR_BAD_ARG, "bad shard count %d", n_shards);
    if (!trace || !work || !shard_bases || col_offset < 0) return fail(TSTWO_ERR_BAD_ARG, "null trace/work/shard bases");
    const uint64_t n = 1ull << log_size, ne = 1ull << ext, rows = 1ull << (ext - sl);
    uint32_t *scr = nullptr;
    if (!coeffs && (rc = coeff_scratch((size_t)n_cols * n, &scr))) return rc;
    uint32_t *cf = coeffs ? (uint32_t *)coeffs : scr;
    // per-shard bases, already advanced to this rank's first column, in the library's device pointer table (slot 2)
    Context &c = ctx();
    uint64_t host_bases[64];
    for (int d = 0; d < n_shards; d++) {
        if (!shard_bases[d] || (shard_bases[d] & 15)) return fail(TSTWO_ERR_BAD_ARG, "shard base %d null or not 16-byte aligned", d);
        host_bases[d] = shard_bases[d] + (uint64_t)col_offset * rows * 4;
    }
    uint64_t *dev_bases = c.d_ptr_table + 2 * 4096;
    TS_CUDA(cudaMemcpyAsync(dev_bases, host_bases, sizeof(uint64_t) * (size_t)n_shards, cudaMemcpyHostToDevice, c.stream));
    ColSet src = { (const uint32_t *)trace, n, nullptr };
    ColSet cfs = { cf, n, nullptr };
    ColSet wk = { (const uint32_t *)work, ne, nullptr };
    ColSet ev = { nullptr, rows, nullptr };
    ev.shard_log = ext - sl;
    ev.shard_bases = dev_bases;
    if ((rc = fft_interpolate(src, cfs, n_cols, log_size, t, flags))) return rc;
    return fft_evaluate(cfs, log_size, ev, n_cols, ext, t, flags, &wk);
}

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
