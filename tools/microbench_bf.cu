// Register-resident radix-16 stage microbenchmark: which butterfly instruction mix and how many warps per SM
// sub-partition saturate the integer pipes of sm_100a.  Mirrors a stage of fft_stream.cu (16 units x 4 lanes per
// thread, 4 layers, 8/4/2/1 twiddles) without any memory traffic.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/microbench_bf tools/microbench_bf.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>
#include "../tstwo_b200/csrc/m31.cuh"
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

// V: 0 plain (compiler picks IMAD.IADD for both adds), 1 sum as IADD3 (a+m+z), 2 alternate 0/1 by unit parity,
//    3 both adds IADD3, 4 alternate 0/1 by lane parity, 5: 3 of 4 plain, 1 of 4 z
// fold of the 62-bit product on the FMA pipe: hi + (lo >> 1) = mad.hi(lo, 2^31, hi)
__device__ __forceinline__ uint32_t mul2_madhi(uint32_t a, uint32_t t2, uint32_t half)   // half = 2^31, run-time so that ptxas keeps the IMAD.HI
{
    uint64_t p = (uint64_t)a * t2;
    uint32_t lo = (uint32_t)p, hi = (uint32_t)(p >> 32), s;
    asm("mad.hi.u32 %0, %1, %2, %3;" : "=r"(s) : "r"(lo), "r"(half), "r"(hi));
    return m31::csub(s);
}
template <int V> __device__ __forceinline__ void bf(uint32_t &a, uint32_t &b, uint32_t t, uint32_t z, int d, int e)
{
    if (V >= 6 && V <= 8) {       // 6: every butterfly folds with mad.hi; 7: lanes 1 of 4; 8: lanes 2 of 4
        const bool hi = V == 6 || (V == 7 && e == 3) || (V == 8 && (e & 1));
        uint32_t m = hi ? mul2_madhi(b, t, z) : m31::mul2(b, t), x = a;
        a = m31::add(x, m); b = m31::sub(x, m);
        return;
    }
    bool usez = V == 1 || (V == 2 && (d & 1)) || (V == 4 && (e & 1)) || (V == 5 && (e == 3));
    if (V == 3) {
        uint32_t m = m31::mul2(b, t), x = a;
        a = m31::csub(x + m + z);
        uint32_t dd = x - m + z; b = __viaddmin_u32(dd, m31::P, dd);
    } else if (usez) m31::butterfly_z(a, b, t, z);
    else m31::butterfly(a, b, t);
}
template <int V, int THREADS>
__global__ void __launch_bounds__(THREADS) k_stage(uint32_t *out, const uint32_t *tw, int iters, uint32_t z)
{
    uint32_t v[16][4];
#pragma unroll
    for (int d = 0; d < 16; d++)
#pragma unroll
        for (int e = 0; e < 4; e++) v[d][e] = (threadIdx.x * 7 + d * 13 + e * 5 + blockIdx.x) % m31::P;
    uint32_t t[15];
#pragma unroll
    for (int i = 0; i < 15; i++) t[i] = tw[(threadIdx.x + i) & 1023];
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int k = 3; k >= 0; k--) {
#pragma unroll
            for (int d = 0; d < 16; d++) {
                if (d & (1 << k)) continue;
                const uint32_t w = t[(8 >> k) - 1 + (d >> (k + 1))];
#pragma unroll
                for (int e = 0; e < 4; e++) bf<V>(v[d][e], v[d | (1 << k)][e], w, z, d, e);
            }
        }
    }
    uint32_t s = 0;
#pragma unroll
    for (int d = 0; d < 16; d++)
#pragma unroll
        for (int e = 0; e < 4; e++) s ^= v[d][e];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int V, int THREADS> int run(const char *name, int sms, int bps, uint32_t *out, uint32_t *tw)
{
    int iters = 2048, grid = sms * bps;
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    const uint32_t z = (V >= 6 && V <= 8) ? 0x80000000u : 0u;
    k_stage<V, THREADS><<<grid, THREADS>>>(out, tw, 16, z); CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0)); k_stage<V, THREADS><<<grid, THREADS>>>(out, tw, iters, z); CK(cudaEventRecord(e1)); CK(cudaDeviceSynchronize());
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    double bfs = (double)grid * THREADS * 128.0 * iters;       // butterflies
    double per_smsp_clk = bfs / 32 / (sms * 4) / (ms * 1e-3 * 1.965e9);
    printf("%-34s threads/SM=%4d (%2d warps/SMSP): %7.1f G butterflies/s, %.2f clk per warp-butterfly per SMSP\n", name, THREADS * bps,
           THREADS * bps / 128, bfs / ms / 1e6, 1.0 / per_smsp_clk);
    return 0;
}
int main()
{
    cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
    int sms = prop.multiProcessorCount;
    uint32_t *out, *tw; CK(cudaMalloc(&out, (size_t)sms * 8 * 512 * 4)); CK(cudaMalloc(&tw, 4096)); CK(cudaMemset(tw, 0x55, 4096));
    run<0, 128>("plain, 1 warp/SMSP", sms, 1, out, tw);
    run<0, 256>("plain", sms, 1, out, tw);
    run<0, 256>("plain", sms, 2, out, tw);
    run<0, 512>("plain 512x1", sms, 1, out, tw);
    run<1, 256>("sum IADD3", sms, 2, out, tw);
    run<2, 256>("alternate by unit", sms, 2, out, tw);
    run<4, 256>("alternate by lane", sms, 2, out, tw);
    run<5, 256>("1 of 4 lanes IADD3", sms, 2, out, tw);
    run<3, 256>("both IADD3", sms, 2, out, tw);
    run<6, 256>("fold by mad.hi (all)", sms, 2, out, tw);
    run<7, 256>("fold by mad.hi (1 of 4 lanes)", sms, 2, out, tw);
    run<8, 256>("fold by mad.hi (2 of 4 lanes)", sms, 2, out, tw);
    run<2, 256>("alternate by unit, 2 warps/SMSP", sms, 1, out, tw);
    run<4, 256>("alternate by lane, 2 warps/SMSP", sms, 1, out, tw);
    return 0;
}
