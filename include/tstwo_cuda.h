/*
 * tstwo_cuda.h -- C ABI of libtstwo_cuda.so, the B200 (sm_100a) CudaBackend for TStwo.
 *
 * Drop-in boundary: these entry points are what a `bun:ffi` (or ctypes / cgo) binding of the reference's
 * backend interfaces binds for the LDE + Blake2s-commit hot path.  Each function cites the reference interface
 * it replaces (paths relative to packages/core/src of teddyjfpender/tstwo).  See INTEGRATION.md for the
 * TypeScript-side stub.
 *
 * Conventions
 *  - Every function returns 0 on success or a negative tstwo_status; tstwo_last_error() gives the message of
 *    the last failure on the calling thread (the reference throws `Error(msg)`).
 *  - Device memory is addressed by plain 64-bit device addresses (tstwo_dptr).  They may come from
 *    tstwo_malloc() or from any other allocator in the same process (e.g. a torch tensor's data_ptr()).
 *  - Field elements are canonical little-endian u32 in [0, 2^31-1).  A SecureField (QM31) column is four
 *    coordinate columns (SecureColumnByCoords, fields/secure_columns.ts:124-217); a QM31 scalar is u32[4] in
 *    to_m31_array order (fields/qm31.ts:139-141).  Hashes are 32 raw bytes.
 *  - Calls are issued from one host thread and are asynchronous on the library stream unless they return
 *    host-visible data (those synchronise before returning).  tstwo_sync() waits for everything.
 *  - Arrays of column addresses (`const tstwo_dptr *cols`) are HOST arrays of device addresses.
 */
#ifndef TSTWO_CUDA_H
#define TSTWO_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef uint64_t tstwo_dptr;      /* device address */
typedef uint64_t tstwo_twiddles;  /* opaque TwiddleTree handle (poly/twiddles.ts:14-29) */

enum tstwo_status {
    TSTWO_OK = 0,
    TSTWO_ERR_BAD_ARG = -1,
    TSTWO_ERR_SIZE_MISMATCH = -2,     /* e.g. "fold_circle_into_line: Length mismatch ..." (fri.ts:167-169) */
    TSTWO_ERR_TWIDDLE_MISMATCH = -3,  /* "twiddle tree mismatch" (backend/cpu/circle.ts:90,141) */
    TSTWO_ERR_CUDA = -4,
    TSTWO_ERR_OOM = -5,
    TSTWO_ERR_NOT_POW2 = -6,          /* "length is not power of two" (backend/cpu/index.ts:65) */
    TSTWO_ERR_NOT_INIT = -7
};

/* flags */
#define TSTWO_FLAG_QUIRKS 1u /* reproduce CpuBackend's log_size==3 swap of values[5]<->values[7]
                                (backend/cpu/circle.ts:127-131,147-151) */

/* ---- lifecycle ------------------------------------------------------------------------------------------- */
int tstwo_init(int device);                 /* select device, create the library stream; idempotent */
/* One process, several devices.  The reference host is ONE single-threaded process (MerkleProver.commit is called inline from
 * prover/index.ts:582-662; no worker / IPC layer), so the multi-GPU fan-out lives inside the library: slot k drives CUDA device
 * ids[k], all pairs get NVLink peer access, and ONE host thread issues everything (per-device streams, events as the
 * cross-device fences).  Slot 0 is current afterwards; tstwo_set_device(slot) makes another slot the target of every
 * following call of this thread (columns, twiddles, transforms ... are per device). */
int tstwo_init_devices(int n_devices, const int *ids);
/* Environment: TSTWO_CUDA_DEVICES="0,1,2,3" is the device list of tstwo_init_devices(n, NULL) (n = 0: all of them);
 * TSTWO_TRACE=1 prints one line per span ("Interpolation for commitment", "Extension", "Merkle", ...) with its host time to
 * stderr, TSTWO_TRACE=2 drains the stream first (device time); TSTWO_FFT_TC / TSTWO_OVERLAP_TAPER: see tstwo_cuda_debug.h. */
int tstwo_device_count(void);               /* slots initialised (1 after tstwo_init) */
int tstwo_set_device(int slot);
int tstwo_shutdown(void);
const char *tstwo_last_error(void);
int tstwo_set_stream(void *cuda_stream);    /* run on a caller-owned cudaStream_t (NULL = library stream) */
int tstwo_sync(void);
int tstwo_device_info(int *sm_count, int *cc_major, int *cc_minor, uint64_t *total_mem);
uint64_t tstwo_launch_count(void);          /* number of kernels this library has launched (for bench.py) */

/* ---- Column<T> / ColumnOps  (backend/index.ts:53-74, :12-14; impl backend/cpu/index.ts:18-53,85-157) ------ */
int tstwo_malloc(uint64_t bytes, tstwo_dptr *out);                       /* Column.zeros/uninitialized */
int tstwo_free(tstwo_dptr p);                                            /* returns the block to the library's free lists */
int tstwo_mem_trim(void);                                                /* gives cached free blocks back to the driver */
int tstwo_memset_zero(tstwo_dptr p, uint64_t bytes);
int tstwo_upload(tstwo_dptr dst, const void *host, uint64_t bytes);      /* createBaseFieldColumn(data) */
int tstwo_download(void *host, tstwo_dptr src, uint64_t bytes);          /* Column.toCpu() / at(i) */
int tstwo_copy(tstwo_dptr dst, tstwo_dptr src, uint64_t bytes);
int tstwo_bit_reverse(tstwo_dptr col, uint64_t len);                     /* ColumnOps.bitReverseColumn; len = #u32 */
/* MerkleProver.decommit (vcs/prover.ts:32-109) and FRI's witness evaluations (fri.ts:346-384) read hashes and
 * column values one index at a time; here the caller lists the device addresses (layer + 32 j, column + 4 i) and one
 * launch gathers words_each consecutive u32 from each into host_out[n * words_each]. */
int tstwo_gather_words(const tstwo_dptr *addrs, int n, int words_each, uint32_t *host_out);
/* MerkleProver.decommit (vcs/prover.ts:32-109) on a tree written by tstwo_merkle_commit (root first: layer k at byte
 * 32 (2^k - 1) of `tree`, max_log = its deepest layer).  cols / col_logs: the committed columns in the caller's order (taken
 * largest first, stable, as the reference sorts them).  Queries: sorted positions of the columns of log size l at
 * query_positions[query_offsets[l] .. query_offsets[l+1]), l = 0..max_log (max_log + 2 offsets).  counts = {queried values,
 * witness hashes, witness column values}.  Called with the three output pointers NULL it only sizes; otherwise the outputs
 * (host) must hold counts[0] words, 32 counts[1] bytes and counts[2] words. */
int tstwo_merkle_decommit(tstwo_dptr tree, int max_log, const tstwo_dptr *cols, const int *col_logs, int n_cols,
                          const int *query_offsets, const uint32_t *query_positions, uint32_t *queried_values,
                          uint8_t *hash_witness, uint32_t *column_witness, uint64_t counts[3]);
/* FriProver.decommitOnQueries' loop over the inner layers (fri.ts:769-785; FriInnerLayerProver.decommit :533-541;
 * computeDecommitmentPositionsAndWitnessEvals :346-384 with fold step 1) in ONE call with one synchronisation.
 *   trees[l], evals[l][4], logs[l] : layer l's tree buffer (root first), its 4 coordinate columns and their log size
 *   queries[n_queries]            : sorted distinct positions on layer 0's domain; folded by 1 from layer to layer
 *   fri_witness / hash_witness / column_witness : host outputs, layer after layer (capacities in evaluations of 4 words / hashes of
 *                                   32 bytes / words); counts[3 l ..] = {FRI witness evaluations, witness hashes, column-witness
 *                                   words} of layer l.  Too small a capacity is an error (nothing is written). */
int tstwo_fri_decommit_layers(int n_layers, const tstwo_dptr *trees, const tstwo_dptr *evals, const int *logs, const uint32_t *queries,
                              int n_queries, uint32_t *fri_witness, uint64_t cap_fri_witness, uint8_t *hash_witness, uint64_t cap_hashes,
                              uint32_t *column_witness, uint64_t cap_column_witness, uint64_t *counts);

/* ---- PolyOps  (poly/circle/ops.ts:20-66; impl backend/cpu/circle.ts:17-239) -------------------------------- */
/* precomputeTwiddles(coset): backend/cpu/circle.ts:210-239.  The tree holds 2^log words of twiddles+itwiddles. */
int tstwo_twiddles_precompute(uint32_t coset_initial_index, int coset_log_size, tstwo_twiddles *out);
int tstwo_twiddles_free(tstwo_twiddles tw);
int tstwo_twiddles_info(tstwo_twiddles tw, uint32_t *root_initial_index, int *root_log_size,
                        tstwo_dptr *twiddles, tstwo_dptr *itwiddles);
/* interpolateColumns / CpuCirclePoly.interpolate (circle.ts:136-207): n_cols evaluations (bit-reversed order) on
 * the circle domain with half coset (half_coset_initial_index, log_size-1) -> coefficients.  evals[i] may equal
 * coeffs[i] (in place). */
int tstwo_interpolate_columns(const tstwo_dptr *evals, const tstwo_dptr *coeffs, int n_cols, int log_size,
                              uint32_t half_coset_initial_index, tstwo_twiddles tw, uint32_t flags);
/* evaluatePolynomials / CpuCirclePoly.evaluate with extend (circle.ts:71-134): coefficient vectors of 2^poly_log
 * words -> 2^domain_log evaluations on the domain with half coset (half_coset_initial_index, domain_log-1). */
int tstwo_evaluate_columns(const tstwo_dptr *coeffs, int poly_log, const tstwo_dptr *evals, int n_cols,
                           int domain_log, uint32_t half_coset_initial_index, tstwo_twiddles tw, uint32_t flags);
/* CpuCirclePoly.eval_at_point (circle.ts:52-69): point = CirclePoint<SecureField>. */
int tstwo_eval_at_point(tstwo_dptr coeffs, int log_size, const uint32_t px[4], const uint32_t py[4], uint32_t out[4]);
/* The same for n_polys polynomials of one size at one point (the sampled values of a commitment tree): one batch of launches,
 * out = n_polys x 4 words. */
int tstwo_eval_at_point_many(const tstwo_dptr *coeffs, int n_polys, int log_size, const uint32_t px[4], const uint32_t py[4],
                             uint32_t *out);
/* tstwo_eval_at_point for callers that must not allocate (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same result as
 * tstwo_eval_at_point, bit for bit, but the partial sums live in a fixed 1.5 MiB file-scope device array (held by every GPU that
 * loads the library until exit) instead of the library's grow-on-demand scratch, and the call performs no cudaMalloc / cudaFree /
 * allocating library routine.  Limit: log_size <= tstwo_eval_at_point_noalloc_max_log() = 20 (2^20 coefficients); larger sizes
 * are rejected with TSTWO_ERR_BAD_ARG and a message naming the limit.  coeffs must be 16-byte aligned for log_size >= 4.
 * All calls share the array, which keeps the last call's contents (nothing in it needs to start at zero: every word read was
 * written earlier in the same call); calls that overlap on the GPU -- two streams, two host threads -- race on it. */
int tstwo_eval_at_point_noalloc(tstwo_dptr coeffs, int log_size, const uint32_t px[4], const uint32_t py[4], uint32_t out[4]);
int tstwo_eval_at_point_noalloc_max_log(void);

/* ---- MerkleOps<Blake2sHash>  (vcs/ops.ts:16-26; vcs/blake2_merkle.ts:9-24; backend/cpu/blake2.ts:98-127) --- */
/* commitOnLayer(logSize, prevLayer, columns): out_layer = 2^log_size hashes (32 B each); prev_layer = 0 or the
 * 2^(log_size+1) hashes of the layer below.  node_i = blake2s(prev[2i] || prev[2i+1] || LE32(col_0[i]) || ...). */
int tstwo_merkle_commit_layer(int log_size, tstwo_dptr prev_layer, const tstwo_dptr *cols, int n_cols,
                              tstwo_dptr out_layer);
/* MerkleProver.commit (vcs/prover.ts:13-30): all layers, root first: layer k (2^k hashes) at byte offset
 * 32*(2^k-1) of `layers` ((2^(max_log+1)-1)*32 bytes).  root[32] is copied to the host (synchronises). */
int tstwo_merkle_commit(const tstwo_dptr *cols, const int *col_logs, int n_cols, tstwo_dptr layers, uint8_t root[32]);
/* The leaf layer of MerkleProver.commit (vcs/prover.ts:13-30, all columns of one length) hashed in COLUMN CHUNKS, for callers whose
 * columns become available in groups (the multi-GPU exchange: the first half of the columns of a rank's rows arrives long before the
 * last).  `layers` is the tree buffer of tstwo_merkle_commit for max_log = log_size; between chunks the BLAKE2s chaining value of
 * row i lives in leaf slot i.  A call absorbs columns [c0, c0 + nc) of leaves of total_cols words, given as nc evenly spaced columns
 * (column k at cols + 4 k col_stride_words); c0 and nc are multiples of 16 (one 64-byte block) except that the last chunk ends at
 * total_cols and finalises.  Chunks must be given in order, c0 = 0 first.  Same digests as hashNode over the whole row
 * (vcs/blake2_merkle.ts:9-24).  tstwo_merkle_commit_nodes then builds layers log_size - 1 .. 0 over the finished leaf layer;
 * root[32] (may be NULL) is copied to the host, which synchronises. */
int tstwo_merkle_leaf_absorb(tstwo_dptr cols, uint64_t col_stride_words, int c0, int nc, int total_cols, int log_size, tstwo_dptr layers);
int tstwo_merkle_commit_nodes(int log_size, tstwo_dptr layers, uint8_t root[32]);
/* GrindOps.grind (backend/cpu/grind.ts:31-42; Backend.grind, backend/index.ts): smallest nonce such that the Blake2sChannel
 * digest after mix_u64(nonce) has >= pow_bits trailing zero bits (channel/blake2.ts:94-111,139-151). */
int tstwo_grind(const uint8_t digest[32], uint32_t pow_bits, uint64_t *nonce);

/* ---- FriOps  (fri.ts:93-110, :120-192; impl backend/cpu/fri.ts:23-164) -------------------------------------- */
/* tw may be 0: the twiddles are then derived from the coset on the fly. */
int tstwo_fold_line(const tstwo_dptr src[4], int log_size, uint32_t coset_initial_index, const uint32_t alpha[4],
                    const tstwo_dptr dst[4], tstwo_twiddles tw);
int tstwo_fold_circle_into_line(const tstwo_dptr dst[4], const tstwo_dptr src[4], int log_size,
                                uint32_t half_coset_initial_index, const uint32_t alpha[4], tstwo_twiddles tw);
int tstwo_fri_decompose(const tstwo_dptr src[4], int log_size, const tstwo_dptr g[4], uint32_t lambda[4]);
/* FriProver.commitInnerLayers (fri.ts:676-716) in ONE call with the Fiat-Shamir transcript kept on the device: per layer
 * MerkleProver.commit of the 4 coordinate columns (fri.ts:519-531) -> Blake2sMerkleChannel.mix_root (vcs/blake2_merkle.ts:27-32)
 * -> channel.draw_felt (channel/blake2.ts:172-181) -> fold_line, then fold_circle_into_line of the circle column whose folded
 * size equals the new layer's (fri.ts:704-711), with the same alpha.  Everything is enqueued back to back; the host waits once.
 *   first[4], first_log, first_coset_init : the layer evaluation after the first circle fold and its LineDomain coset
 *   last_log                             : stop when the layer has 2^last_log elements (config.last_layer_domain_size())
 *   circle_cols[n_circle][4], circle_logs, circle_half_inits : the remaining circle columns, descending sizes
 *   channel_state[19]                    : in/out  digest[8], n_challenges, n_sent, queue_len, queue[8] (Blake2sChannel state)
 *   layer_evals[n_layers][4]             : out, evaluation columns of layers 1 .. n_layers; the four coordinates of a layer are
 *                                          the quarters of ONE tstwo_malloc block (evenly spaced columns take the bulk-copy leaf
 *                                          kernel): the caller frees layer_evals[l][0] only
 *   layer_trees[n_layers]                : out, Merkle tree buffer of layers 0 .. n_layers-1 (root first; caller frees)
 *   roots[n_layers][32], alphas[n_layers][4], *n_layers : out, n_layers = first_log - last_log */
int tstwo_fri_commit_layers(const tstwo_dptr first[4], int first_log, uint32_t first_coset_init, int last_log,
                            const tstwo_dptr *circle_cols, const int *circle_logs, const uint32_t *circle_half_inits, int n_circle,
                            tstwo_twiddles tw, uint32_t channel_state[19], tstwo_dptr *layer_evals, tstwo_dptr *layer_trees,
                            uint8_t *roots, uint32_t *alphas, int *n_layers);


/* ---- AccumulationOps / QuotientOps  (backend/cpu/accumulation.ts:38-63, backend/cpu/quotients.ts:52-178) ---- */
int tstwo_accumulate(const tstwo_dptr dst[4], const tstwo_dptr src[4], uint64_t len);
/* Sample batches are flattened: batch b has point (px[4], py[4]) at points[8b..8b+8) and entries
 * [batch_offsets[b], batch_offsets[b+1]) of (col_index[j], values[4j..4j+4)).  Host arrays. */
int tstwo_accumulate_quotients(int log_size, uint32_t half_coset_initial_index, const tstwo_dptr *cols, int n_cols,
                               const uint32_t random_coeff[4], int n_batches, const uint32_t *points,
                               const int *batch_offsets, const int *col_index, const uint32_t *values,
                               const tstwo_dptr out[4]);
/* Same with flags: TSTWO_FLAG_QUIRKS = the denominators exactly as backend/cpu/quotients.ts:160-178 computes them (from the M31
 * parts of the sample point only; what the function above does), 0 = Stwo's CM31 denominators (pcs/quotients.ts Rust text,
 * `denominator_inverses`), the ones for which the quotients are low degree and FRI accepts them. */
int tstwo_accumulate_quotients_ex(int log_size, uint32_t half_coset_initial_index, const tstwo_dptr *cols, int n_cols,
                                  const uint32_t random_coeff[4], int n_batches, const uint32_t *points,
                                  const int *batch_offsets, const int *col_index, const uint32_t *values,
                                  const tstwo_dptr out[4], uint32_t flags);

/* Wide-Fibonacci constraint quotients over the evaluation domain (examples/fibonacci.ts:72-91: constraint i is
 * c - (a^2 + b^2) on columns i, i+1, i+2; the domain loop is Stwo's FrameworkComponent::evaluate_constraint_quotients_on_domain,
 * a placeholder at constraint_framework/index.ts:131-152).  cols: n_cols evaluations of the trace on
 * CanonicCoset(eval_log_size).circleDomain(), bit-reversed.  coeffs: n_cols-2 QM31, the random-coefficient power that multiplies
 * constraint i (ColumnAccumulator.randomCoeffPowers reversed, air/accumulator.ts:159-176).  denom_inv: 2^(eval_log_size -
 * trace_log_size) M31, 1 / coset_vanishing(CanonicCoset(trace_log_size).coset, domain.at(i)) bit-reversed (constraints.ts:32-47).
 * out[row] += denom_inv[row >> trace_log_size] * sum_i coeffs[i] * constraint_i(row).  Host arrays except cols / out. */
int tstwo_widefib_constraints(int trace_log_size, int eval_log_size, const tstwo_dptr *cols, int n_cols, const uint32_t *coeffs,
                              const uint32_t *denom_inv, const tstwo_dptr out[4]);

/* ---- the metric path in one call: Stwo's CommitmentTreeProver::new (quoted at pcs/prover.ts:215-235) -------- */
/* interpolate n_cols trace columns on CanonicCoset(log_size), evaluate on CanonicCoset(log_size+log_blowup),
 * Blake2s Merkle commit.  trace/coeffs/evals are contiguous column-major blocks (column c at base + c*len).
 * coeffs may be 0 (coefficients are then not kept).  tw must cover half_odds(log_size+log_blowup-1). */
int tstwo_lde_commit(tstwo_dptr trace, int n_cols, int log_size, int log_blowup, tstwo_twiddles tw,
                     tstwo_dptr coeffs, tstwo_dptr evals, tstwo_dptr layers, uint8_t root[32], uint32_t flags);
/* Multi-GPU building block: LDE of this rank's n_cols columns with the evaluations written grouped by destination
 * rank, [n_shards][n_cols][2^(log_size+log_blowup)/n_shards] (contiguous row ranges = complete Merkle subtrees), ready
 * for one all-to-all.  The receiver then calls tstwo_merkle_commit on its [all columns][rows] block. */
int tstwo_lde_columns_sharded(tstwo_dptr trace, int n_cols, int log_size, int log_blowup, tstwo_twiddles tw,
                              tstwo_dptr coeffs, tstwo_dptr evals, int n_shards, uint32_t flags);
/* Multi-GPU, exchange fused into the last FFT pass: the rows of shard d are stored straight into shard_bases[d], the
 * [total columns][2^(log_size+log_blowup)/n_shards] leaf buffer of rank d (its own, or a peer's mapped with
 * tstwo_ipc_import), at column col_offset + c.  work: n_cols * 2^(log_size+log_blowup) words of local scratch.
 * Replaces tstwo_lde_columns_sharded + all-to-all; the caller synchronises the ranks before the receivers hash. */
int tstwo_lde_columns_to_peers(tstwo_dptr trace, int n_cols, int log_size, int log_blowup, tstwo_twiddles tw,
                               tstwo_dptr coeffs, tstwo_dptr work, const tstwo_dptr *shard_bases, int n_shards,
                               int col_offset, uint32_t flags);
/* The same call with the exchange overlapped with ALL of the LDE's arithmetic: the columns are transformed in col_groups
 * groups; the last pass of group k (the one whose stores cross NVLink) runs on a side stream on at most peer_sms SMs while the
 * other SMs interpolate and extend group k + 1.  Identical results for every (col_groups, peer_sms); col_groups = 1 or
 * peer_sms = 0 is tstwo_lde_columns_to_peers.  On return the library's stream is ordered after both streams' work. */
int tstwo_lde_columns_to_peers_overlapped(tstwo_dptr trace, int n_cols, int log_size, int log_blowup, tstwo_twiddles tw,
                                          tstwo_dptr coeffs, tstwo_dptr work, const tstwo_dptr *shard_bases, int n_shards,
                                          int col_offset, uint32_t flags, int col_groups, int peer_sms);
/* The overlapped call for COLUMN-INTERLEAVED ownership, which lets the receivers hash half of every leaf while the exchange is still
 * running: the rank's first n_cols / 2 columns are columns col_offset_lo.. of the trace, the others col_offset_hi.. (rank r of G
 * holding c columns: r c / 2 and total / 2 + r c / 2), so that once every rank's first col_groups / 2 groups have arrived the
 * receivers hold columns [0, total / 2) of their rows -- a prefix of every leaf's message.  col_groups must be even, >= 2 and divide
 * n_cols; peer_sms > 0.  The side stream stores `value` (as tstwo_peer_signal) to the n_flags words half_flag_addrs once the first
 * half's stores are complete and to arrive_flag_addrs once all are.  The LAST group's storing pass also keeps to peer_sms SMs and,
 * unlike the call above, the side stream is NOT joined on return: what the caller enqueues next on the library's stream
 * (tstwo_peer_wait on the half flags, tstwo_merkle_leaf_absorb of columns [0, total / 2)) runs beside that pass on the other SMs.
 * tstwo_peer_join then makes the library's stream wait for the side stream. */
int tstwo_lde_columns_to_peers_halves(tstwo_dptr trace, int n_cols, int log_size, int log_blowup, tstwo_twiddles tw,
                                      tstwo_dptr coeffs, tstwo_dptr work, const tstwo_dptr *shard_bases, int n_shards,
                                      int col_offset_lo, int col_offset_hi, uint32_t flags, int col_groups, int peer_sms,
                                      const tstwo_dptr *half_flag_addrs, const tstwo_dptr *arrive_flag_addrs, int n_flags, uint32_t value);
int tstwo_peer_join(void);
/* Multi-GPU, exchange on the copy engines: shard d of a tstwo_lde_columns_sharded block ([n_shards][n_cols][rows_per_shard]) is
 * copied into shard_bases[d] at column col_offset (destination layout as above) on side streams, after the work the library's
 * stream holds now and concurrently with what it is given next (the next column group's LDE).  tstwo_scatter_join makes the
 * library's stream wait for all copies issued so far. */
int tstwo_scatter_shards(tstwo_dptr evals_sharded, int n_cols, uint64_t rows_per_shard, const tstwo_dptr *shard_bases,
                         int n_shards, int col_offset);
int tstwo_scatter_join(void);
/* CUDA IPC for the above (one process per GPU): 64-byte handle of a tstwo_malloc'ed buffer / mapping of a peer's. */
int tstwo_ipc_export(tstwo_dptr ptr, uint8_t handle[64]);
int tstwo_ipc_import(const uint8_t handle[64], tstwo_dptr *out);
int tstwo_ipc_close(tstwo_dptr ptr);
/* Cross-device fences WITHOUT a collective: 32-bit flags in peer-mapped device memory (CUDA IPC or peer access over NVLink),
 * written with st.release.sys and polled with ld.acquire.sys by one-warp kernels on the library's stream.  They replace the
 * all-reduce / all-gather that only ordered the steps of the exchange above (vcs/prover.ts:13-30 is one tree: the ranks must agree
 * on "all rows have arrived" and "all leaves are hashed").  Values are step counters compared in wrap-around order.
 *   tstwo_peer_signal  once everything enqueued so far on the library's stream has completed, store `value` to the n words
 *                      flag_addrs[i] (n <= 64; typically one word in every peer's flag block);
 *   tstwo_peer_wait    the library's stream waits until each of the n words at `flags` (n <= 64, LOCAL device memory) has reached
 *                      `value`; traps after timeout_ms (0 = 4000) instead of hanging the device when a peer never arrives;
 *   tstwo_peer_put     copy n_words (<= 1024) from src to dst (a peer address), then signal flag_addr = value. */
int tstwo_peer_signal(const tstwo_dptr *flag_addrs, int n, uint32_t value);
int tstwo_peer_wait(tstwo_dptr flags, int n, uint32_t value, uint32_t timeout_ms);
int tstwo_peer_put(tstwo_dptr dst, tstwo_dptr src, int n_words, tstwo_dptr flag_addr, uint32_t value);
/* Same, but the trace lives in HOST memory (pinned or pageable): host->device copies are pipelined with the
 * kernels inside the call; only the root comes back. */
int tstwo_lde_commit_host(const uint32_t *host_trace, int n_cols, int log_size, int log_blowup, tstwo_twiddles tw,
                          tstwo_dptr trace_staging, tstwo_dptr coeffs, tstwo_dptr evals, tstwo_dptr layers,
                          uint8_t root[32], uint32_t flags);

/* The metric path on ALL devices of tstwo_init_devices in one call from one host thread (BASELINE config 3; CommitmentTreeProver
 * ::new over the whole trace, pcs/prover.ts:215-235): columns are sharded by column for the LDE, the last FFT pass stores every
 * row range into its owner's leaf buffer over NVLink peer memory, every device hashes its rows and builds its subtree, device 0
 * finishes the top log2 G layers.  The root equals tstwo_lde_commit's on one device, bit for bit.  The library owns all device
 * buffers (allocated at the first call for a shape).  G = tstwo_device_count() must be a power of two dividing n_cols, and every
 * device needs >= 2^14 rows.
 *   tstwo_lde_commit_multi      host_trace = [n_cols][2^log_size] words in HOST memory (pinned for full PCIe speed): every device
 *                               pulls its shard over its own PCIe link, chunk by chunk, overlapped with the transforms (which
 *                               columns a device pulls is the library's choice: for n_cols a multiple of 32, its share of the first
 *                               half of the columns and then of the second, so that half of every leaf is hashed during the copy);
 *   tstwo_lde_commit_multi_dev  trace_shards[d] = device address ON slot d of columns [d n_cols/G, (d+1) n_cols/G). */
int tstwo_lde_commit_multi(const uint32_t *host_trace, int n_cols, int log_size, int log_blowup, uint8_t root[32], uint32_t flags);
int tstwo_lde_commit_multi_dev(const tstwo_dptr *trace_shards, int n_cols, int log_size, int log_blowup, uint8_t root[32],
                               uint32_t flags);

#ifdef __cplusplus
}
#endif
#endif /* TSTWO_CUDA_H */
