/*
 * tstwo_cuda_debug.h -- test / measurement hooks of libtstwo_cuda.so.  NOT part of the drop-in boundary
 * (include/tstwo_cuda.h): nothing a TStwo host calls.  They exist so that tests/ can force every kernel family and
 * replay the reference's golden field vectors on the device arithmetic, and so that bench.py can name what ran.
 */
#ifndef TSTWO_CUDA_DEBUG_H
#define TSTWO_CUDA_DEBUG_H

#include "tstwo_cuda.h"

#ifdef __cplusplus
extern "C" {
#endif

/* FFT pass family: 0 = bulk-copy (TMA) persistent passes where they apply (default), 1 = cp.async persistent passes
 * (fft_stream.cu), 2 = one CTA per tile (fft.cu) */
int tstwo_debug_set_fft_stream(int mode);
/* per-layer kernels only (the on-device cross-check of the tiled passes) */
int tstwo_debug_force_naive_fft(int on);
/* the 512-row (9-layer) strided pass of the bulk-copy family, planned when exactly nine layers remain above the contiguous pass
 * (log size 23); 0 = plan an 8-layer window plus a one-layer pass as before (default 1) */
int tstwo_debug_set_fft_rows9(int on);
/* rows per thread of the fft.cu family: 16 (default) or 8 */
int tstwo_debug_set_fft_radix(int rows_per_thread);
/* what the last transform ran: bit 0 tiled, bits 1.. family */
int tstwo_debug_last_fft_mode(void);
/* top stage of the bulk-copy passes on the tensor cores (tcgen05 kind::i8, tc_stage.cuh): bit 0 = forward strided passes
 * (default), bit 1 = contiguous passes, bit 2 = inverse strided passes; 0 = integer pipes everywhere */
int tstwo_debug_set_fft_tc(int mask);
/* tstwo_lde_columns_to_peers_overlapped: percent of an equal share that the first and the last column group get (100 = equal
 * groups).  Smaller edge groups shorten the two phases that have no partner (the first group's own passes, the last group's
 * NVLink-bound storing pass); every phase then splits the SMs in proportion to the columns on its two sides. */
int tstwo_debug_set_overlap_taper(int percent);

/* CUDA events around every bulk-copy FFT pass launch from now on (on = 1; 0 stops and forgets).  tstwo_debug_pass_times
 * waits for the recorded passes and returns how many there were (<= max): tags[i] bit 0 = inverse, bit 1 = strided pass,
 * bit 2 = tensor-core top stage; ms[i] = duration of that launch on the library stream. */
int tstwo_debug_pass_timing(int on);
int tstwo_debug_pass_times(int max, int *tags, float *ms);

/* Merkle: bit 0 = leaf layer by the bulk-copy-fed kernel where it applies (default 1; 0 = merkle_layer_kernel always);
 * bit 1 set = one launch per node layer instead of the fused bottom layers (merkle_tail_kernel) */
int tstwo_debug_set_leaf_mode(int mode);
int tstwo_debug_last_leaf_mode(void);

/* DEEP quotients / wide-Fibonacci: 0 auto, 1 integer-pipe kernel, 2 INT8 tensor-core kernel */
int tstwo_debug_set_quot_mode(int mode);
int tstwo_debug_last_quot_mode(void);

/* Device field arithmetic (csrc/m31.cuh), one operation per operand set.  a, b, out: HOST arrays of n x 4 words.
 * field 0 = M31 (word 0), 1 = CM31 (words 0..1), 2 = QM31.  op: 0 add, 1 sub, 2 mul, 3 neg, 4 inverse,
 * 5 QM31 x CM31 (fields/qm31.ts mul_cm31), 6 x M31 (b word 0), 7 M31 reduce of the u64 (a word 0 | word 1 << 32)
 * (fields/m31.ts:60-101), 8 M31 partial_reduce (fields/m31.ts:39-44), 9 M31 square. */
int tstwo_debug_field_op(int field, int op, const uint32_t *a, const uint32_t *b, uint32_t *out, int n);

#ifdef __cplusplus
}
#endif
#endif
