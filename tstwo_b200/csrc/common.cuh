// Shared host-side plumbing for libtstwo_cuda.so: context, error slot, launch accounting.
#pragma once
#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>   // header-only NVTX 3: ranges cost nothing unless a profiler is attached

#include <cstdint>
#include <cstdio>
#include <string>

#include "../../include/tstwo_cuda.h"

namespace tstwo {

constexpr int MAX_SLOTS = 16;   // devices one process can drive (tstwo_init_devices)

// Everything the library keeps per DEVICE.  One process may drive several devices from one host thread (the reference host
// is a single process: SURVEY 8b "Conventions"): slot k holds device ids[k] of tstwo_init_devices, slot 0 is the device of
// tstwo_init.  Entry points work on the CURRENT slot (tstwo_set_device; default 0).
struct Context {
    bool ready = false;
    int slot = 0;
    int device = -1;
    int sm_count = 0;
    cudaStream_t own_stream = nullptr;   // library stream
    cudaStream_t stream = nullptr;       // stream in use (own_stream or caller's)
    cudaStream_t copy_stream = nullptr;  // H2D pipeline for tstwo_lde_commit_host
    uint64_t launches = 0;
    int grid_limit = 0;                  // > 0: persistent FFT passes launch at most this many CTAs (SM partitioning between
                                         // two streams, tstwo_lde_columns_to_peers_overlapped); 0 = one CTA per SM
    // small device scratch for pointer tables / parameters
    uint64_t *d_ptr_table = nullptr;
    size_t ptr_table_cap = 0;
    // per-device one-time kernel attributes (cudaFuncSetAttribute applies to the current device only)
    bool attr_fft = false, attr_stream = false, attr_tma = false, attr_leaf = false, attr_leaf_chunk = false;
    // pipeline.cu: coefficient scratch, side stream + events of the overlapped peer-store path
    uint32_t *coeff_scratch = nullptr;
    size_t coeff_scratch_words = 0;
    cudaStream_t peer_stream = nullptr;
    cudaEvent_t peer_ev[2] = { nullptr, nullptr };
    // elementwise.cu scratch
    uint32_t *ew_scratch = nullptr;
    size_t ew_scratch_bytes = 0;
};

Context &ctx();                 // context of the calling thread's current slot
Context &ctx_at(int slot);
int ctx_slots();                // number of initialised slots
// makes `slot` current for the calling thread (cudaSetDevice + ctx()); returns the previous slot
int ctx_select(int slot);
struct DeviceScope {            // RAII form
    int prev;
    explicit DeviceScope(int slot) : prev(ctx_select(slot)) {}
    ~DeviceScope() { ctx_select(prev); }
};
int fail(int code, const char *fmt, ...);

// NVTX range for the lifetime of the object; the names are Stwo's tracing spans (Rust text quoted in the reference at
// packages/core/src/pcs/prover.ts:44,92-97,180,222-226): "Commitment", "Interpolation for commitment", "Extension", "Merkle",
// "Evaluate columns out of domain", "Grind" -- so that an Nsight timeline of a TStwo host reads like Stwo's own trace
// TSTWO_TRACE=1 (env): every span also prints "tstwo: <span> <ms>" to stderr when it closes -- host time of the call
// (enqueue + whatever the call waits for); TSTWO_TRACE=2 drains the library's stream first, so the figure is the device's.
int trace_level();                                  // abi_core.cu: getenv once
void trace_span_end(const char *name, double t0);   // prints (and drains the stream for level 2)
double trace_now();
struct Span {
    const char *name_;
    double t0_;
    explicit Span(const char *name) : name_(name), t0_(trace_level() ? trace_now() : 0.0) { nvtxRangePushA(name); }
    ~Span() { nvtxRangePop(); if (trace_level()) trace_span_end(name_, t0_); }
};

#define TS_CUDA(expr)                                                                              \
    do {                                                                                           \
        cudaError_t _e = (expr);                                                                   \
        if (_e != cudaSuccess)                                                                     \
            return ::tstwo::fail(TSTWO_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                                 __FILE__, __LINE__);                                              \
    } while (0)

#define TS_REQUIRE_INIT()                                                                          \
    do {                                                                                           \
        if (!::tstwo::ctx().ready) return ::tstwo::fail(TSTWO_ERR_NOT_INIT, "tstwo_init() has not been called"); \
    } while (0)

// every kernel launch goes through this so tstwo_launch_count() is exact
#define TS_LAUNCH(kernel, grid, block, smem, ...)                                                  \
    do {                                                                                           \
        kernel<<<(grid), (block), (smem), ::tstwo::ctx().stream>>>(__VA_ARGS__);                   \
        ::tstwo::ctx().launches++;                                                                 \
        cudaError_t _e = cudaGetLastError();                                                       \
        if (_e != cudaSuccess)                                                                     \
            return ::tstwo::fail(TSTWO_ERR_CUDA, "launch %s failed: %s (%s:%d)", #kernel,          \
                                 cudaGetErrorString(_e), __FILE__, __LINE__);                      \
    } while (0)

// Column address list: either uniform stride (base + c*stride words) or a device table of addresses.
struct ColSet {
    const uint32_t *base;       // used when table == nullptr
    uint64_t stride;            // words
    const uint64_t *table;      // device array of addresses, or nullptr
    // Optional row-sharded layout [shard][column][rows]: element idx of column c lives at
    // col(c) + idx + (idx >> shard_log) * shard_adj  (shard_log = 31 disables it: idx < 2^31).
    int shard_log = 31;
    uint64_t shard_adj = 0;
    bool aligned16 = true;      // every column starts on a 16-byte boundary (what the bulk copies of fft_tma.cu need)
    // Optional per-shard base addresses (device array of 2^k pointers): shard s = idx >> shard_log of column c starts at
    // shard_bases[s] + c * stride words.  The bases may be PEER memory (another GPU's buffer mapped over NVLink): this
    // is how the last FFT pass of the multi-GPU path stores every row range straight into its owner's leaf buffer.
    const uint64_t *shard_bases = nullptr;
    // address of element idx of column c; runs of a tile (<= 2^shard_log aligned indices) are contiguous from it
    __device__ __forceinline__ uint32_t *ptr(int c, uint32_t idx) const
    {
        if (shard_bases)
            return reinterpret_cast<uint32_t *>(shard_bases[idx >> shard_log]) + (uint64_t)c * stride + (idx & ((1u << shard_log) - 1));
        return col(c) + off(idx);
    }
    __host__ __device__ __forceinline__ uint32_t *col(int c) const
    {
        return table ? reinterpret_cast<uint32_t *>(table[c]) : const_cast<uint32_t *>(base) + (uint64_t)c * stride;
    }
    __host__ __device__ __forceinline__ uint64_t off(uint32_t idx) const { return idx + (uint64_t)(idx >> shard_log) * shard_adj; }
};
// Builds a ColSet from a host array of device addresses; uploads a table only when the stride is not uniform.
// `slot` selects one of several independent table regions (so src and dst sets can coexist in one launch).
int make_colset(const tstwo_dptr *cols, int n_cols, int slot, ColSet *out);

// ---- twiddle trees ----
struct TwiddleTree {
    int slot;                // device slot the tables live on
    uint32_t root_init;
    int root_log;
    uint32_t *tw, *itw;      // reference layout (backend/cpu/circle.ts:210-239), 2^root_log words each
    uint32_t *tw2, *itw2;    // the same tables pre-doubled (2*t), consumed by the butterfly kernels
};
TwiddleTree *get_tree(tstwo_twiddles h);
// is `half coset (init, log)` a repeated doubling of the tree root?  (circle.ts:264-266)
bool is_doubling_of(uint32_t init, int log, const TwiddleTree *t);

// kernels implemented in the other translation units (host launchers)
// out2[i] = 2 / coord(init + bitrev(i, m) * subgroup_gen(step_log)), i < 2^m; which: 0 = x, 1 = y
int launch_coset_points(uint32_t *out2, int which, uint32_t init, int step_log, int m);
cudaError_t elementwise_init();   // per-TU constant tables of elementwise.cu
int fft_interpolate(const ColSet &src, const ColSet &dst, int n_cols, int log_size, const TwiddleTree *t, uint32_t flags);
int fft_evaluate(const ColSet &src, int poly_log, const ColSet &dst, int n_cols, int domain_log, const TwiddleTree *t,
                 uint32_t flags, const ColSet *work = nullptr, int pass_select = 0);
int merkle_commit_layer(int log_size, const uint8_t *prev, const ColSet &cols, int n_cols, uint8_t *out);
// node layers start_log .. 0 of the tree buffer `layers` (root first) in one launch; layer start_log + 1 must be there.
// Returns the layer it started from (start_log clamped to what the kernel takes), or -1 when the fused form is switched off.
int merkle_tail(uint8_t *layers, int start_log);
// leaf layer fed in column chunks: absorbs columns [c0, c0 + nc) (given as `cols`) of leaves of total_cols words
int merkle_leaf_absorb(int log_size, const ColSet &cols, int c0, int nc, int total_cols, uint8_t *leaf_layer);
// the same through the bulk-copy leaf kernel where its tensor map applies (evenly spaced aligned columns, >= 256 rows)
int merkle_leaf_absorb_fast(int log_size, const ColSet &cols, int c0, int nc, int total_cols, uint8_t *leaf_layer);
// tstwo_shutdown: frees the scratch and the side stream / events pipeline.cu keeps between calls
void pipeline_release();

}  // namespace tstwo
