// Device-side circle-group helpers: CirclePointIndex.to_point (circle.ts:172-174) via a table of G*2^k.
// Each translation unit that includes this header gets its own __constant__ copy; it must be filled once by
// calling upload_gen_table() (done from tstwo_init through the per-TU init hooks).
#pragma once
#include "host_m31.h"
#include "m31.cuh"

namespace tstwo {

static __constant__ m31::Pt c_gen_pow[31];   // G * 2^k, k = 0..30

static inline cudaError_t upload_gen_table()
{
    host::Pt tab[31];
    host::Pt cur = { 2u, 1268011823u };      // circle.ts:137 M31_CIRCLE_GEN
    for (int k = 0; k < 31; k++) { tab[k] = cur; cur = host::padd(cur, cur); }
    return cudaMemcpyToSymbol(c_gen_pow, tab, sizeof tab);
}

__device__ __forceinline__ m31::Pt index_to_point(uint32_t index)
{
    m31::Pt r = { 1u, 0u };
#pragma unroll 1
    for (int k = 0; k < 31; k++)
        if ((index >> k) & 1) r = m31::padd(r, c_gen_pow[k]);
    return r;
}

// CircleDomain.indexAt (poly/circle/domain.ts:76-88): i < half -> half_coset.index_at(i), else -index_at(i-half)
__device__ __forceinline__ uint32_t domain_index_at(uint32_t half_init, int half_log, uint32_t i)
{
    uint32_t half = 1u << half_log;
    uint32_t step = half_log == 0 ? 0u : (1u << (31 - half_log));
    if (i < half) return (half_init + step * i) & 0x7fffffffu;
    return (0u - (half_init + step * (i - half))) & 0x7fffffffu;
}

}  // namespace tstwo
