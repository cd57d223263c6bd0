// Multi-GPU LDE + Blake2s commit driven by ONE host thread of ONE process (SURVEY 8b "Conventions", 8e): the reference host is
// a single-threaded Bun process that calls MerkleProver.commit inline (packages/core/src/vcs/prover.ts:13-30 from
// prover/index.ts:582-662), so the fan-out over the devices of tstwo_init_devices lives inside the library:
//
//   device d  (its own stream)   LDE of columns [d * n_cols/G, (d+1) * n_cols/G); the last FFT pass stores every row range
//                                straight into its owner's leaf buffer over NVLink peer memory (cudaDeviceEnablePeerAccess,
//                                no CUDA IPC, no NCCL) -- the fused exchange of tstwo_lde_columns_to_peers_overlapped
//   event fan-in                 device d waits for the "LDE done" events of ALL devices (cross-device cudaStreamWaitEvent),
//   device d                     hashes its 2^ext/G leaves, builds its subtree, copies its root to device 0,
//   device 0                     waits for the G roots, finishes the top log2 G layers, root to the host.
//
// Every call is enqueued asynchronously device by device; the only host wait is the final one for the root.
#include <condition_variable>
#include <cstring>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

#include "common.cuh"
#include "host_m31.h"

using namespace tstwo;

namespace {

// The CALLER is one host thread (the reference host is single-threaded), but enqueuing ~70 launches / event operations per
// device costs ~0.35 ms of host time each: at 8 devices that alone is the whole 2.9 ms step.  The library therefore keeps one
// helper thread per extra device; for_each_device(fn) runs fn(slot) for all slots concurrently (slot 0 on the caller's thread)
// and returns when all are done.  Helpers only enqueue; they never outlive tstwo_shutdown.
class Fanout {
public:
    int run(int n, const std::function<int(int)> &fn)
    {
        ensure(n);
        {
            std::lock_guard<std::mutex> lk(mu_);
            fn_ = &fn; pending_ = n - 1; gen_++;
            for (int i = 0; i < n; i++) rc_[i] = 0;
        }
        cv_.notify_all();
        rc_[0] = fn(0);
        std::unique_lock<std::mutex> lk(mu_);
        done_.wait(lk, [&] { return pending_ == 0; });
        for (int i = 0; i < n; i++)
            if (rc_[i]) { if (i) fail(rc_[i], "%s", err_[i]); return rc_[i]; }
        return 0;
    }
    void stop()
    {
        {
            std::lock_guard<std::mutex> lk(mu_);
            quit_ = true; gen_++;
        }
        cv_.notify_all();
        for (auto &t : th_) t.join();
        th_.clear(); quit_ = false; n_ = 1;
    }
private:
    void ensure(int n)
    {
        if (n == n_) return;
        stop();
        n_ = n;
        const uint64_t g0 = gen_;
        for (int i = 1; i < n; i++) th_.emplace_back([this, i, g0] { worker(i, g0); });
    }
    void worker(int slot, uint64_t seen)
    {
        for (;;) {
            const std::function<int(int)> *fn;
            {
                std::unique_lock<std::mutex> lk(mu_);
                cv_.wait(lk, [&] { return gen_ != seen; });
                seen = gen_;
                if (quit_) return;
                fn = fn_;
            }
            int rc = (*fn)(slot);
            if (rc) { strncpy(err_[slot], tstwo_last_error(), sizeof err_[slot] - 1); err_[slot][sizeof err_[slot] - 1] = 0; }
            std::lock_guard<std::mutex> lk(mu_);
            rc_[slot] = rc;
            if (--pending_ == 0) done_.notify_one();
        }
    }
    std::mutex mu_;
    std::condition_variable cv_, done_;
    std::vector<std::thread> th_;
    const std::function<int(int)> *fn_ = nullptr;
    uint64_t gen_ = 0;
    int n_ = 1, pending_ = 0, rc_[MAX_SLOTS] = { 0 };
    char err_[MAX_SLOTS][256];
    bool quit_ = false;
} g_fan;

struct Dev {
    tstwo_twiddles tw = 0;
    uint32_t *stage = nullptr, *work = nullptr, *leaf = nullptr;   // host-fed staging [lc][n], evaluations scratch [lc][ne], leaf buffer [n_cols][rows]
    uint8_t *layers = nullptr;                                     // this device's subtree, root first
    cudaEvent_t lde_done = nullptr, sub_done = nullptr, stage_free = nullptr, half_done = nullptr;
    std::vector<cudaEvent_t> chunk_ev;
};
struct Multi {
    int n_dev = 0, n_cols = 0, log_size = 0, log_blowup = 0;
    bool has_stage = false, ran = false;
    Dev d[MAX_SLOTS];
    uint8_t *roots = nullptr, *top = nullptr;    // device 0: [G][32] subtree roots, top layers (G - 1 hashes, root first)
    uint8_t *host_root = nullptr;                // pinned
} g_m;

void release_all()
{
    for (int i = 0; i < g_m.n_dev; i++) {
        if (!ctx_at(i).ready) continue;
        DeviceScope sc(i);
        Dev &v = g_m.d[i];
        cudaDeviceSynchronize();
        if (v.tw) tstwo_twiddles_free(v.tw);
        cudaFree(v.stage); cudaFree(v.work); cudaFree(v.leaf); cudaFree(v.layers);
        for (cudaEvent_t e : { v.lde_done, v.sub_done, v.stage_free, v.half_done }) if (e) cudaEventDestroy(e);
        for (cudaEvent_t e : v.chunk_ev) cudaEventDestroy(e);
        if (i == 0) { cudaFree(g_m.roots); cudaFree(g_m.top); }
        v = Dev();
    }
    if (g_m.host_root) cudaFreeHost(g_m.host_root);
    g_m = Multi();
}

int dev_alloc(void **p, size_t bytes, const char *what)
{
    cudaError_t e = cudaMalloc(p, bytes);
    if (e == cudaErrorMemoryAllocation) { cudaGetLastError(); return fail(TSTWO_ERR_OOM, "tstwo_lde_commit_multi: %s (%zu bytes) out of memory", what, bytes); }
    TS_CUDA(e);
    return 0;
}

int ensure(int n_cols, int log_size, int log_blowup, bool need_stage)
{
    const int G = ctx_slots();
    if (g_m.n_dev == G && g_m.n_cols == n_cols && g_m.log_size == log_size && g_m.log_blowup == log_blowup && (g_m.has_stage || !need_stage))
        return 0;
    release_all();
    const int ext = log_size + log_blowup;
    int sl = 0;
    while ((1 << sl) < G) sl++;
    const uint64_t n = 1ull << log_size, ne = 1ull << ext, rows = ne >> sl;
    const int lc = n_cols / G;
    g_m.n_dev = G; g_m.n_cols = n_cols; g_m.log_size = log_size; g_m.log_blowup = log_blowup; g_m.has_stage = need_stage;
    for (int i = 0; i < G; i++) {
        DeviceScope sc(i);
        Dev &v = g_m.d[i];
        int rc = tstwo_twiddles_precompute(host::subgroup_gen(ext - 1 + 2), ext - 1, &v.tw);
        if (rc) return rc;
        if ((rc = dev_alloc((void **)&v.work, (size_t)lc * ne * 4, "evaluation scratch"))) return rc;
        if ((rc = dev_alloc((void **)&v.leaf, (size_t)n_cols * rows * 4, "leaf buffer"))) return rc;
        if ((rc = dev_alloc((void **)&v.layers, 32ull * (2 * rows - 1), "subtree"))) return rc;
        if (need_stage && (rc = dev_alloc((void **)&v.stage, (size_t)lc * n * 4, "trace staging"))) return rc;
        for (cudaEvent_t *e : { &v.lde_done, &v.sub_done, &v.stage_free, &v.half_done }) TS_CUDA(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
        if (i == 0) {
            if ((rc = dev_alloc((void **)&g_m.roots, 32ull * G, "subtree roots"))) return rc;
            if ((rc = dev_alloc((void **)&g_m.top, 32ull * (G > 1 ? G : 2), "top layers"))) return rc;
        }
    }
    TS_CUDA(cudaMallocHost((void **)&g_m.host_root, 32));
    return 0;
}

// column groups / SMs of the NVLink-bound pass for the overlapped exchange (profiles/r01_overlap.txt)
void overlap_params(int G, int *groups, int *peer_sms) { *groups = G > 1 ? 4 : 1; *peer_sms = G >= 8 ? 66 : G > 1 ? 62 : 0; }   // profiles/r02_multi.txt

int check(int n_cols, int log_size, int log_blowup)
{
    const int G = ctx_slots();
    if (G < 1) return fail(TSTWO_ERR_NOT_INIT, "tstwo_init_devices() has not been called");
    if (G & (G - 1)) return fail(TSTWO_ERR_BAD_ARG, "the number of devices (%d) must be a power of two", G);
    if (n_cols < G || n_cols % G) return fail(TSTWO_ERR_BAD_ARG, "n_cols (%d) must be a multiple of the number of devices (%d)", n_cols, G);
    int sl = 0;
    while ((1 << sl) < G) sl++;
    if (log_size < 1 || log_blowup < 0 || log_size + log_blowup > 30 || log_size + log_blowup - sl < 14)
        return fail(TSTWO_ERR_BAD_ARG, "bad sizes: every device needs at least 2^14 rows of the evaluation domain");
    return 0;
}

// steps 2..4 of the header comment, after every device's LDE has been enqueued and lde_done recorded
int subtrees_and_top(uint8_t root[32], bool first_half_absorbed = false)
{
    const int G = g_m.n_dev, n_cols = g_m.n_cols, ext = g_m.log_size + g_m.log_blowup;
    int sl = 0;
    while ((1 << sl) < G) sl++;
    const uint64_t rows = (1ull << ext) >> sl;
    int rc = g_fan.run(G, [&](int i) -> int {
        std::vector<tstwo_dptr> ptrs((size_t)n_cols);
        std::vector<int> logs((size_t)n_cols, ext - sl);
        DeviceScope sc(i);
        Context &c = ctx();
        Dev &v = g_m.d[i];
        for (int e = 0; e < G; e++) TS_CUDA(cudaStreamWaitEvent(c.stream, g_m.d[e].lde_done, 0));
        int r;
        if (first_half_absorbed) {                             // columns [0, n_cols / 2) are in the leaf slots' chaining values already
            r = tstwo_merkle_leaf_absorb((tstwo_dptr)(v.leaf + (uint64_t)(n_cols / 2) * rows), rows, n_cols / 2, n_cols / 2, n_cols, ext - sl, (tstwo_dptr)v.layers);
            if (!r) r = tstwo_merkle_commit_nodes(ext - sl, (tstwo_dptr)v.layers, nullptr);
        } else {
            for (int k = 0; k < n_cols; k++) ptrs[k] = (tstwo_dptr)(v.leaf + (uint64_t)k * rows);
            r = tstwo_merkle_commit(ptrs.data(), logs.data(), n_cols, (tstwo_dptr)v.layers, nullptr);
        }
        if (r) return r;
        TS_CUDA(cudaMemcpyPeerAsync(g_m.roots + 32 * i, ctx_at(0).device, v.layers, c.device, 32, c.stream));
        TS_CUDA(cudaEventRecord(v.sub_done, c.stream));
        return 0;
    });
    if (rc) return rc;
    DeviceScope sc(0);
    Context &c = ctx();
    for (int e = 1; e < G; e++) TS_CUDA(cudaStreamWaitEvent(c.stream, g_m.d[e].sub_done, 0));
    const uint8_t *prev = g_m.roots;
    for (int k = sl - 1; k >= 0; k--) {                        // layer k has 2^k hashes, at byte 32 (2^k - 1) of `top`
        uint8_t *dst = g_m.top + 32ull * ((1ull << k) - 1);
        rc = tstwo_merkle_commit_layer(k, (tstwo_dptr)prev, nullptr, 0, (tstwo_dptr)dst);
        if (rc) return rc;
        prev = dst;
    }
    TS_CUDA(cudaMemcpyAsync(g_m.host_root, prev, 32, cudaMemcpyDeviceToHost, c.stream));
    TS_CUDA(cudaStreamSynchronize(c.stream));
    if (root) memcpy(root, g_m.host_root, 32);
    g_m.ran = true;
    return 0;
}

}  // namespace

namespace tstwo {
void multi_release() { release_all(); g_fan.stop(); }
}

extern "C" {

// Device-resident shards: trace_shards[d] = columns [d * n_cols/G, (d+1) * n_cols/G) as [n_cols/G][2^log_size] words ON device
// slot d.  The library owns every other buffer (allocated at the first call for a shape, kept until the shape changes).
int tstwo_lde_commit_multi_dev(const tstwo_dptr *trace_shards, int n_cols, int log_size, int log_blowup, uint8_t root[32], uint32_t flags)
{
    int rc = check(n_cols, log_size, log_blowup);
    if (rc) return rc;
    if (!trace_shards) return fail(TSTWO_ERR_BAD_ARG, "null trace shards");
    if ((rc = ensure(n_cols, log_size, log_blowup, false))) return rc;
    const int G = g_m.n_dev, lc = n_cols / G;
    int groups, peer_sms;
    overlap_params(G, &groups, &peer_sms);
    tstwo_dptr bases[MAX_SLOTS];
    for (int i = 0; i < G; i++) bases[i] = (tstwo_dptr)g_m.d[i].leaf;
    for (int i = 0; i < G; i++)
        if (!trace_shards[i]) return fail(TSTWO_ERR_BAD_ARG, "trace shard %d is null", i);
    rc = g_fan.run(G, [&](int i) -> int {
        DeviceScope sc(i);
        Context &c = ctx();
        // my stores go into every device's leaf buffer: the previous step's hashing of those must be over
        if (g_m.ran) for (int e = 0; e < G; e++) TS_CUDA(cudaStreamWaitEvent(c.stream, g_m.d[e].sub_done, 0));
        int r = tstwo_lde_columns_to_peers_overlapped(trace_shards[i], lc, log_size, log_blowup, g_m.d[i].tw, 0, (tstwo_dptr)g_m.d[i].work,
                                                      bases, G, i * lc, flags, groups, peer_sms);
        if (r) return r;
        TS_CUDA(cudaEventRecord(g_m.d[i].lde_done, c.stream));
        return 0;
    });
    if (rc) return rc;
    return subtrees_and_top(root);
}

// The host's call: host_trace = [n_cols][2^log_size] words in HOST memory (pinned for full PCIe speed; pageable works).  Every
// device pulls its column shard over its own PCIe link in column chunks on its copy stream, and each chunk's LDE + peer stores
// start as soon as it has landed.
int tstwo_lde_commit_multi(const uint32_t *host_trace, int n_cols, int log_size, int log_blowup, uint8_t root[32], uint32_t flags)
{
    int rc = check(n_cols, log_size, log_blowup);
    if (rc) return rc;
    if (!host_trace) return fail(TSTWO_ERR_BAD_ARG, "null host trace");
    if ((rc = ensure(n_cols, log_size, log_blowup, true))) return rc;
    const int G = g_m.n_dev, lc = n_cols / G;
    const uint64_t n = 1ull << log_size, ne = 1ull << (log_size + log_blowup);
    int chunks = lc < 4 ? lc : 4;
    while (lc % chunks) chunks--;
    const int per = lc / chunks;
    // Column-interleaved ownership (the library picks which columns a device pulls): device i takes its share of the trace's FIRST
    // half of the columns, then its share of the second, so that after everyone's first chunks/2 chunks every device holds columns
    // [0, n_cols/2) of its rows and hashes that half of every leaf while the second half is still crossing PCIe (as
    // PeerShardedCommit(ownership="halves"), tstwo_b200/distributed.py).  Needs whole BLAKE2s blocks per half and an even chunk count.
    const bool halves = n_cols % 32 == 0 && chunks % 2 == 0;
    auto first_col = [&](int i, int k) -> uint64_t {           // trace column of local column k * per of device i
        if (!halves) return (uint64_t)i * lc + (uint64_t)k * per;
        const int h = lc / 2, c = k * per;
        return c < h ? (uint64_t)i * h + c : (uint64_t)n_cols / 2 + (uint64_t)i * h + (c - h);
    };
    tstwo_dptr bases[MAX_SLOTS];
    for (int i = 0; i < G; i++) bases[i] = (tstwo_dptr)g_m.d[i].leaf;
    // all copies first, so that every PCIe link starts at once
    for (int i = 0; i < G; i++) {
        DeviceScope sc(i);
        Context &c = ctx();
        Dev &v = g_m.d[i];
        while ((int)v.chunk_ev.size() < chunks) { cudaEvent_t e; TS_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); v.chunk_ev.push_back(e); }
        if (g_m.ran) TS_CUDA(cudaStreamWaitEvent(c.copy_stream, v.stage_free, 0));   // the previous step's transforms have read the staging buffer
        for (int k = 0; k < chunks; k++) {
            TS_CUDA(cudaMemcpyAsync(v.stage + (uint64_t)k * per * n, host_trace + first_col(i, k) * n, (uint64_t)per * n * 4,
                                    cudaMemcpyHostToDevice, c.copy_stream));
            TS_CUDA(cudaEventRecord(v.chunk_ev[k], c.copy_stream));
        }
    }
    // chunks [k0, k1) of every device; `after_half`: first wait for every device's first half, hash it, then go on
    auto run_chunks = [&](int k0, int k1, bool after_half) -> int {
        return g_fan.run(G, [&](int i) -> int {
            DeviceScope sc(i);
            Context &c = ctx();
            Dev &v = g_m.d[i];
            if (k0 == 0 && g_m.ran) for (int e = 0; e < G; e++) TS_CUDA(cudaStreamWaitEvent(c.stream, g_m.d[e].sub_done, 0));
            if (after_half) {
                int sl = 0;
                while ((1 << sl) < G) sl++;
                const uint64_t rows = ne >> sl;
                for (int e = 0; e < G; e++) TS_CUDA(cudaStreamWaitEvent(c.stream, g_m.d[e].half_done, 0));
                int r = tstwo_merkle_leaf_absorb((tstwo_dptr)v.leaf, rows, 0, n_cols / 2, n_cols, log_size + log_blowup - sl, (tstwo_dptr)v.layers);
                if (r) return r;
            }
            for (int k = k0; k < k1; k++) {
                TS_CUDA(cudaStreamWaitEvent(c.stream, v.chunk_ev[k], 0));
                int r = tstwo_lde_columns_to_peers((tstwo_dptr)(v.stage + (uint64_t)k * per * n), per, log_size, log_blowup, v.tw, 0,
                                                   (tstwo_dptr)(v.work + (uint64_t)k * per * ne), bases, G, (int)first_col(i, k), flags);
                if (r) return r;
            }
            if (k1 == chunks) {
                TS_CUDA(cudaEventRecord(v.stage_free, c.stream));
                TS_CUDA(cudaEventRecord(v.lde_done, c.stream));
            } else TS_CUDA(cudaEventRecord(v.half_done, c.stream));
            return 0;
        });
    };
    // an event must be RECORDED (host side) before another device's stream is told to wait for it: one fan-out per phase
    if (halves) {
        if ((rc = run_chunks(0, chunks / 2, false))) return rc;
        if ((rc = run_chunks(chunks / 2, chunks, true))) return rc;
    } else if ((rc = run_chunks(0, chunks, false))) return rc;
    return subtrees_and_top(root, halves);
}

}  // extern "C"
