// Parameters shared by the circle-FFT pass kernels (fft.cu: one CTA per tile; fft_stream.cu: persistent, pipelined).
#pragma once
#include "common.cuh"

namespace tstwo {

struct PassParams {
    ColSet src, dst;
    int n_cols;
    int n;                 // log size of the index space of dst
    uint32_t src_mask;     // src index = idx & src_mask (coset replication of zero-extended coefficients)
    int lo;                // window = index bits [lo, lo+RHO)
    int j_min;             // relative layers j in [j_min, RHO) are applied
    const uint32_t *tw2;   // pre-doubled twiddle (or inverse twiddle) tree
    uint32_t tree_len;     // 2^root_log
    uint32_t scale2;       // 2/2^n after the last inverse layer, or 0
    uint32_t zero;         // run-time 0 (see m31::butterfly_z)
    int win = 0;           // fft_tma.cu: width of a strided window, 8 (default when 0) or 9 (the 512-row pass)
};

// fft_stream.cu: same pass semantics (contiguous 12-layer window at lo == 0, strided 10-layer windows above), run by
// one persistent CTA per SM with the next tiles prefetched into shared memory while the current ones are transformed.
int stream_pass(bool inverse, const PassParams &p);
int stream_pass_init();

// fft_tma.cu: persistent CTAs fed by bulk asynchronous copies (TMA).  Window sizes differ from the kernels above:
// the contiguous pass covers TMA_W0 = 14 layers (one column x 16384 indices per tile), strided passes TMA_W1 = 8.
constexpr int TMA_W0 = 14, TMA_W1 = 8, TMA_W9 = 9;   // TMA_W9: the 512-row strided pass, used when exactly 9 layers remain above the contiguous pass
int tma_pass(bool inverse, const PassParams &p);
int tma_pass_init();
bool tma_pass_supported(const PassParams &p);
// cuTensorMapEncodeTiled of the loaded driver (resolved once by tma_pass_init), or nullptr; also used by blake2s.cu
void *tma_encode_tiled_fn();

}  // namespace tstwo
