// Microbenchmarks that set the integer-pipe rooflines for the M31 butterfly and Blake2s kernels on B200.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/microbench tools/microbench.cu
// Prints one line per test: name, ops per SM per clock (from clock64), and chip-wide Gops/s (from CUDA events).
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <cuda_runtime.h>
#include "../tstwo_b200/csrc/m31.cuh"

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

constexpr int ILP = 8;

template <int V>
__device__ __forceinline__ void op(uint32_t &a, uint32_t &b, uint32_t t)
{
    if (V == 0) { m31::butterfly(a, b, t); }                                   // product butterfly (7 instr)
    else if (V == 1) { m31::ibutterfly(a, b, t); }
    else if (V == 2) {                                                         // butterfly without VIADDMNMX
        uint64_t p = (uint64_t)b * t; uint32_t s = (uint32_t)(p >> 32) + ((uint32_t)p >> 1);
        uint32_t m = min(s, s - m31::P); uint32_t x = a + m, y = a - m;
        a = min(x, x - m31::P); b = min(y, y + m31::P);
    }
    else if (V == 3) { a = __viaddmin_u32(a, t, b); b = __viaddmin_u32(b, t, a); }   // VIADDMNMX only (2 per op)
    else if (V == 4) { a = a + b + t; b = b + a + t; }                               // IADD3 only (2 per op)
    else if (V == 5) { a = a * t + b; b = b * t + a; }                               // IMAD only (2 per op)
    else if (V == 6) { a = (a ^ b) & t; b = (b | a) ^ t; }                           // LOP3 only (2 per op)
    else if (V == 7) { a = __funnelshift_r(a, b, 7); b = __funnelshift_r(b, a, 12); } // SHF only (2 per op)
    else if (V == 8) { a = __byte_perm(a, b, 0x1032); b = __byte_perm(b, a, 0x0321); } // PRMT only (2 per op)
    else if (V == 9) { uint64_t p = (uint64_t)a * t + b; a = (uint32_t)p; b = (uint32_t)(p >> 32) ^ t; } // IMAD.WIDE + LOP
    else if (V == 11) { uint64_t p = (uint64_t)a * t; a = (uint32_t)p ^ b; b = (uint32_t)(p >> 32) ^ a; }   // IMAD.WIDE + 2 LOP
    else if (V == 12) { a = __umulhi(a, t) ^ b; b = __umulhi(b, t) ^ a; }                                   // IMAD.HI + LOP (x2)
    else if (V == 13) { a = a * t + b; b = (b ^ a) & t; }                                                   // 1 IMAD + 1 LOP3
    else if (V == 14) { uint32_t s = a + b; a = m31::csub(s); uint32_t d = a - b; b = __viaddmin_u32(d, m31::P, d); } // add+sub mod P (4)
    else if (V == 15) { a = (a >> 1) + b; b = (b >> 1) + a; }                                               // LEA.HI-like x2
    else if (V == 10) {                                                              // Shoup-style: 3 IMAD-pipe + 1 ALU mul
        uint32_t q = __umulhi(b, t); uint32_t r = b * (t >> 1) - q * m31::P; uint32_t m = m31::csub(r);
        uint32_t x = a; a = m31::add(x, m); b = m31::sub(x, m);
    }
}

template <int V>
__global__ void __launch_bounds__(256) k_ops(uint32_t *out, int iters, long long *cycles)
{
    uint32_t a[ILP], b[ILP];
    uint32_t t = 2 * (threadIdx.x * 2654435761u % m31::P);
#pragma unroll
    for (int i = 0; i < ILP; i++) { a[i] = (threadIdx.x * 7 + i * 13 + blockIdx.x) % m31::P; b[i] = (threadIdx.x * 11 + i * 17 + 5) % m31::P; }
    long long c0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < ILP; i++) op<V>(a[i], b[i], t);
    }
    long long c1 = clock64();
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) s ^= a[i] ^ b[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cycles[blockIdx.x] = c1 - c0;
}

// ---- blake2s compress chain
__device__ __forceinline__ uint32_t rotr16(uint32_t x) { return __byte_perm(x, x, 0x1032); }
__device__ __forceinline__ uint32_t rotr8(uint32_t x) { return __byte_perm(x, x, 0x0321); }
__device__ __forceinline__ uint32_t rotr12(uint32_t x) { return __funnelshift_r(x, x, 12); }
__device__ __forceinline__ uint32_t rotr7(uint32_t x) { return __funnelshift_r(x, x, 7); }
#define G(a, b, c, d, x, y) do { a = a + b + (x); d = rotr16(d ^ a); c = c + d; b = rotr12(b ^ c); a = a + b + (y); d = rotr8(d ^ a); c = c + d; b = rotr7(b ^ c); } while (0)
#define RND(s0, s1, s2, s3, s4, s5, s6, s7, s8, s9, s10, s11, s12, s13, s14, s15) do { \
    G(v0, v4, v8, v12, m[s0], m[s1]); G(v1, v5, v9, v13, m[s2], m[s3]); G(v2, v6, v10, v14, m[s4], m[s5]); G(v3, v7, v11, v15, m[s6], m[s7]); \
    G(v0, v5, v10, v15, m[s8], m[s9]); G(v1, v6, v11, v12, m[s10], m[s11]); G(v2, v7, v8, v13, m[s12], m[s13]); G(v3, v4, v9, v14, m[s14], m[s15]); } while (0)
__device__ __forceinline__ void compress(uint32_t (&h)[8], const uint32_t (&m)[16], uint32_t t)
{
    uint32_t v0 = h[0], v1 = h[1], v2 = h[2], v3 = h[3], v4 = h[4], v5 = h[5], v6 = h[6], v7 = h[7];
    uint32_t v8 = 0x6A09E667u, v9 = 0xBB67AE85u, v10 = 0x3C6EF372u, v11 = 0xA54FF53Au, v12 = 0x510E527Fu ^ t, v13 = 0x9B05688Cu, v14 = 0x1F83D9ABu, v15 = 0x5BE0CD19u;
    RND(0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15); RND(14, 10, 4, 8, 9, 15, 13, 6, 1, 12, 0, 2, 11, 7, 5, 3);
    RND(11, 8, 12, 0, 5, 2, 15, 13, 10, 14, 3, 6, 7, 1, 9, 4); RND(7, 9, 3, 1, 13, 12, 11, 14, 2, 6, 5, 10, 4, 0, 15, 8);
    RND(9, 0, 5, 7, 2, 4, 10, 15, 14, 1, 11, 12, 6, 8, 3, 13); RND(2, 12, 6, 10, 0, 11, 8, 3, 4, 13, 7, 5, 15, 14, 1, 9);
    RND(12, 5, 1, 15, 14, 13, 4, 10, 0, 7, 6, 3, 9, 2, 8, 11); RND(13, 11, 7, 14, 12, 1, 3, 9, 5, 0, 15, 4, 8, 6, 2, 10);
    RND(6, 15, 14, 9, 11, 3, 0, 8, 12, 2, 13, 7, 1, 4, 10, 5); RND(10, 2, 8, 4, 7, 6, 1, 5, 15, 11, 9, 14, 3, 12, 13, 0);
    h[0] ^= v0 ^ v8; h[1] ^= v1 ^ v9; h[2] ^= v2 ^ v10; h[3] ^= v3 ^ v11; h[4] ^= v4 ^ v12; h[5] ^= v5 ^ v13; h[6] ^= v6 ^ v14; h[7] ^= v7 ^ v15;
}
// same compression with every addition forced onto the FMA pipe (x * one + y)
#define GA(x, y) ((x) * one + (y))
#define G2(a, b, c, d, x, y) do { a = GA(GA(a, b), (x)); d = rotr16(d ^ a); c = GA(c, d); b = rotr12(b ^ c); a = GA(GA(a, b), (y)); d = rotr8(d ^ a); c = GA(c, d); b = rotr7(b ^ c); } while (0)
// only the two-input adds on the FMA pipe, a + b + m stays IADD3
#define G3(a, b, c, d, x, y) do { a = a + b + (x); d = rotr16(d ^ a); c = GA(c, d); b = rotr12(b ^ c); a = a + b + (y); d = rotr8(d ^ a); c = GA(c, d); b = rotr7(b ^ c); } while (0)
#define RNDX(GX, s0, s1, s2, s3, s4, s5, s6, s7, s8, s9, s10, s11, s12, s13, s14, s15) do { \
    GX(v0, v4, v8, v12, m[s0], m[s1]); GX(v1, v5, v9, v13, m[s2], m[s3]); GX(v2, v6, v10, v14, m[s4], m[s5]); GX(v3, v7, v11, v15, m[s6], m[s7]); \
    GX(v0, v5, v10, v15, m[s8], m[s9]); GX(v1, v6, v11, v12, m[s10], m[s11]); GX(v2, v7, v8, v13, m[s12], m[s13]); GX(v3, v4, v9, v14, m[s14], m[s15]); } while (0)
template <int VAR>
__device__ __forceinline__ void compress_v(uint32_t (&h)[8], const uint32_t (&m)[16], uint32_t t, uint32_t one)
{
    uint32_t v0 = h[0], v1 = h[1], v2 = h[2], v3 = h[3], v4 = h[4], v5 = h[5], v6 = h[6], v7 = h[7];
    uint32_t v8 = 0x6A09E667u, v9 = 0xBB67AE85u, v10 = 0x3C6EF372u, v11 = 0xA54FF53Au, v12 = 0x510E527Fu ^ t, v13 = 0x9B05688Cu, v14 = 0x1F83D9ABu, v15 = 0x5BE0CD19u;
#define ALLR(GX) RNDX(GX, 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15); RNDX(GX, 14, 10, 4, 8, 9, 15, 13, 6, 1, 12, 0, 2, 11, 7, 5, 3); \
    RNDX(GX, 11, 8, 12, 0, 5, 2, 15, 13, 10, 14, 3, 6, 7, 1, 9, 4); RNDX(GX, 7, 9, 3, 1, 13, 12, 11, 14, 2, 6, 5, 10, 4, 0, 15, 8); \
    RNDX(GX, 9, 0, 5, 7, 2, 4, 10, 15, 14, 1, 11, 12, 6, 8, 3, 13); RNDX(GX, 2, 12, 6, 10, 0, 11, 8, 3, 4, 13, 7, 5, 15, 14, 1, 9); \
    RNDX(GX, 12, 5, 1, 15, 14, 13, 4, 10, 0, 7, 6, 3, 9, 2, 8, 11); RNDX(GX, 13, 11, 7, 14, 12, 1, 3, 9, 5, 0, 15, 4, 8, 6, 2, 10); \
    RNDX(GX, 6, 15, 14, 9, 11, 3, 0, 8, 12, 2, 13, 7, 1, 4, 10, 5); RNDX(GX, 10, 2, 8, 4, 7, 6, 1, 5, 15, 11, 9, 14, 3, 12, 13, 0);
    if (VAR == 0) { ALLR(G) } else if (VAR == 1) { ALLR(G2) } else { ALLR(G3) }
    h[0] ^= v0 ^ v8; h[1] ^= v1 ^ v9; h[2] ^= v2 ^ v10; h[3] ^= v3 ^ v11; h[4] ^= v4 ^ v12; h[5] ^= v5 ^ v13; h[6] ^= v6 ^ v14; h[7] ^= v7 ^ v15;
}
// ---- rotations on the FMA pipe: rotr(x, r) = lo32(x * 2^(32-r)) + hi32(x * 2^(32-r)) = IMAD.HI + IMAD (the multiplier is a
// run-time value so that ptxas cannot turn the product back into a shift).  ROTM = 0 none, 1 = rotr7 of every second G,
// 2 = rotr7 of every G, 3 = rotr7 and rotr12 of every G, 4 = rotr7 of one G in four.
__device__ __forceinline__ uint32_t rotm(uint32_t x, uint32_t mul) { return x * mul + __umulhi(x, mul); }
#define G4(a, b, c, d, x, y, U7, U12) do { a = GA(GA(a, b), (x)); d = rotr16(d ^ a); c = GA(c, d); b = (U12) ? rotm(b ^ c, c20) : rotr12(b ^ c); \
    a = GA(GA(a, b), (y)); d = rotr8(d ^ a); c = GA(c, d); b = (U7) ? rotm(b ^ c, c25) : rotr7(b ^ c); } while (0)
#define RND4(s0, s1, s2, s3, s4, s5, s6, s7, s8, s9, s10, s11, s12, s13, s14, s15) do { \
    G4(v0, v4, v8, v12, m[s0], m[s1], ROTM >= 1, ROTM == 3); G4(v1, v5, v9, v13, m[s2], m[s3], ROTM == 2 || ROTM == 3, ROTM == 3); \
    G4(v2, v6, v10, v14, m[s4], m[s5], (ROTM >= 1 && ROTM <= 3), ROTM == 3); G4(v3, v7, v11, v15, m[s6], m[s7], ROTM == 2 || ROTM == 3, ROTM == 3); \
    G4(v0, v5, v10, v15, m[s8], m[s9], (ROTM >= 1 && ROTM <= 3), ROTM == 3); G4(v1, v6, v11, v12, m[s10], m[s11], ROTM == 2 || ROTM == 3, ROTM == 3); \
    G4(v2, v7, v8, v13, m[s12], m[s13], ROTM >= 1, ROTM == 3); G4(v3, v4, v9, v14, m[s14], m[s15], ROTM == 2 || ROTM == 3, ROTM == 3); } while (0)
template <int ROTM>
__device__ __forceinline__ void compress_r(uint32_t (&h)[8], const uint32_t (&m)[16], uint32_t t, uint32_t one, uint32_t c25, uint32_t c20)
{
    uint32_t v0 = h[0], v1 = h[1], v2 = h[2], v3 = h[3], v4 = h[4], v5 = h[5], v6 = h[6], v7 = h[7];
    uint32_t v8 = 0x6A09E667u, v9 = 0xBB67AE85u, v10 = 0x3C6EF372u, v11 = 0xA54FF53Au, v12 = 0x510E527Fu ^ t, v13 = 0x9B05688Cu, v14 = 0x1F83D9ABu, v15 = 0x5BE0CD19u;
    RND4(0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15); RND4(14, 10, 4, 8, 9, 15, 13, 6, 1, 12, 0, 2, 11, 7, 5, 3);
    RND4(11, 8, 12, 0, 5, 2, 15, 13, 10, 14, 3, 6, 7, 1, 9, 4); RND4(7, 9, 3, 1, 13, 12, 11, 14, 2, 6, 5, 10, 4, 0, 15, 8);
    RND4(9, 0, 5, 7, 2, 4, 10, 15, 14, 1, 11, 12, 6, 8, 3, 13); RND4(2, 12, 6, 10, 0, 11, 8, 3, 4, 13, 7, 5, 15, 14, 1, 9);
    RND4(12, 5, 1, 15, 14, 13, 4, 10, 0, 7, 6, 3, 9, 2, 8, 11); RND4(13, 11, 7, 14, 12, 1, 3, 9, 5, 0, 15, 4, 8, 6, 2, 10);
    RND4(6, 15, 14, 9, 11, 3, 0, 8, 12, 2, 13, 7, 1, 4, 10, 5); RND4(10, 2, 8, 4, 7, 6, 1, 5, 15, 11, 9, 14, 3, 12, 13, 0);
    h[0] ^= v0 ^ v8; h[1] ^= v1 ^ v9; h[2] ^= v2 ^ v10; h[3] ^= v3 ^ v11; h[4] ^= v4 ^ v12; h[5] ^= v5 ^ v13; h[6] ^= v6 ^ v14; h[7] ^= v7 ^ v15;
}
template <int ROTM>
__global__ void __launch_bounds__(256) k_blake_rot(uint32_t *out, unsigned long long n_rows, int blocks, uint32_t one, uint32_t c25, uint32_t c20)
{
    for (unsigned long long row0 = (unsigned long long)blockIdx.x * 256; row0 < n_rows; row0 += (unsigned long long)gridDim.x * 256) {
        uint32_t h[8], m[16];
        const unsigned long long i = row0 + threadIdx.x;
        for (int k = 0; k < 8; k++) h[k] = threadIdx.x * 31 + k;
        for (int k = 0; k < 16; k++) m[k] = threadIdx.x * 17 + k * 3 + blockIdx.x;
        for (int b = 0; b < blocks; b++) compress_r<ROTM>(h, m, b, one, c25, c20);
        uint32_t s = 0;
        for (int k = 0; k < 8; k++) s ^= h[k];
        out[i] = s;
    }
}
template <int ROTM> int run_rot(const char *name, uint32_t *out, unsigned long long n_rows, int blocks, int sms, uint32_t *check)
{
    const int grid = (int)(n_rows / 256);
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    k_blake_rot<ROTM><<<grid, 256>>>(out, n_rows, blocks, 1u, 1u << 25, 1u << 20); CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0)); k_blake_rot<ROTM><<<grid, 256>>>(out, n_rows, blocks, 1u, 1u << 25, 1u << 20); CK(cudaEventRecord(e1)); CK(cudaDeviceSynchronize());
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    uint32_t first[4]; CK(cudaMemcpy(first, out, 16, cudaMemcpyDeviceToHost));
    if (ROTM == 0) memcpy(check, first, 16);
    double comp = (double)n_rows * blocks;
    printf("blake2s rotations on the FMA pipe: %-34s chip %7.2f Gcompress/s  => %.0f clk per warp-compress per SMSP (%.3f ms) %s\n",
           name, comp / ms / 1e6, 1.965e9 / (comp / (ms * 1e-3) / 32 / (sms * 4)), ms, memcmp(check, first, 16) ? "RESULT DIFFERS" : "same result");
    return 0;
}
template <int VAR>
__global__ void __launch_bounds__(256) k_blake_v(uint32_t *out, int iters, uint32_t one)
{
    uint32_t h[8], m[16];
    for (int i = 0; i < 8; i++) h[i] = threadIdx.x * 31 + i;
    for (int i = 0; i < 16; i++) m[i] = threadIdx.x * 17 + i * 3 + blockIdx.x;
    for (int it = 0; it < iters; it++) { compress_v<VAR>(h, m, it, one); m[it & 15] ^= h[0]; }
    uint32_t s = 0;
    for (int i = 0; i < 8; i++) s ^= h[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// ---- what separates the leaf kernel from the pure compress loop: launch shape and message loads
// LOADS: 0 = message stays in registers, 1 = 16 column-major loads per compression (row i of column c at in[c * n_rows + i]),
// each CTA handles row tiles of 256 rows: tile = blockIdx.x, blockIdx.x + gridDim.x, ... ; `blocks` compressions per row.
template <int LOADS>
__global__ void __launch_bounds__(256) k_blake_shape(uint32_t *out, const uint32_t *__restrict__ in, unsigned long long n_rows, int blocks, uint32_t one)
{
    for (unsigned long long row0 = (unsigned long long)blockIdx.x * 256; row0 < n_rows; row0 += (unsigned long long)gridDim.x * 256) {
        uint32_t h[8], m[16];
        const unsigned long long i = row0 + threadIdx.x;
        for (int k = 0; k < 8; k++) h[k] = threadIdx.x * 31 + k;
        for (int k = 0; k < 16; k++) m[k] = threadIdx.x * 17 + k * 3 + blockIdx.x;
        for (int b = 0; b < blocks; b++) {
            if (LOADS) {
#pragma unroll
                for (int k = 0; k < 16; k++) m[k] = __ldcs(in + (unsigned long long)(16 * b + k) * n_rows + i);
            }
            compress_v<1>(h, m, b, one);
        }
        uint32_t s = 0;
        for (int k = 0; k < 8; k++) s ^= h[k];
        out[i] = s;
    }
}
template <int LOADS> int run_shape(const char *name, int grid, uint32_t *out, const uint32_t *in, unsigned long long n_rows, int blocks, int sms)
{
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    k_blake_shape<LOADS><<<grid, 256>>>(out, in, n_rows, blocks, 1u); CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0)); k_blake_shape<LOADS><<<grid, 256>>>(out, in, n_rows, blocks, 1u); CK(cudaEventRecord(e1)); CK(cudaDeviceSynchronize());
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    double comp = (double)n_rows * blocks;
    printf("blake2s shape %-44s grid=%6d  chip %7.2f Gcompress/s  => %.0f clk per warp-compress per SMSP @1965MHz (%.3f ms)\n",
           name, grid, comp / ms / 1e6, 1.965e9 / (comp / (ms * 1e-3) / 32 / (sms * 4)), ms);
    return 0;
}
template <int VAR> int run_blake(const char *name, int sms, uint32_t *out)
{
    for (int bps : {2, 4, 6}) {
        int iters = 4096, grid = sms * bps;
        cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
        k_blake_v<VAR><<<grid, 256>>>(out, 512, 1u); CK(cudaDeviceSynchronize());
        CK(cudaEventRecord(e0)); k_blake_v<VAR><<<grid, 256>>>(out, iters, 1u); CK(cudaEventRecord(e1)); CK(cudaDeviceSynchronize());
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        double comp = (double)grid * 256 * iters;
        printf("blake2s %-24s blocks/SM=%d  chip %7.2f Gcompress/s  => %.0f clk per warp-compress per SMSP @1965MHz (%.2f ms)\n",
               name, bps, comp / ms / 1e6, 1.965e9 / (comp / (ms * 1e-3) / 32 / (sms * 4)), ms);
    }
    return 0;
}

__global__ void __launch_bounds__(256) k_blake(uint32_t *out, int iters, long long *cycles)
{
    uint32_t h[8], m[16];
    for (int i = 0; i < 8; i++) h[i] = threadIdx.x * 31 + i;
    for (int i = 0; i < 16; i++) m[i] = threadIdx.x * 17 + i * 3 + blockIdx.x;
    long long c0 = clock64();
    for (int it = 0; it < iters; it++) { compress(h, m, it); m[it & 15] ^= h[0]; }
    long long c1 = clock64();
    uint32_t s = 0;
    for (int i = 0; i < 8; i++) s ^= h[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cycles[blockIdx.x] = c1 - c0;
}

__global__ void k_copy(const uint4 *__restrict__ src, uint4 *__restrict__ dst, size_t n)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
    for (; i + 3 * stride < n; i += 4 * stride) {
        uint4 a = __ldcs(src + i), b = __ldcs(src + i + stride), c = __ldcs(src + i + 2 * stride), d = __ldcs(src + i + 3 * stride);
        __stcs(dst + i, a); __stcs(dst + i + stride, b); __stcs(dst + i + 2 * stride, c); __stcs(dst + i + 3 * stride, d);
    }
    for (; i < n; i += stride) dst[i] = src[i];
}

template <int V> int run_ops(const char *name, int ops_per_call, int blocks_per_sm, int sms, uint32_t *out, long long *cyc)
{
    int iters = 65536, grid = sms * blocks_per_sm;
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    k_ops<V><<<grid, 256>>>(out, 64, cyc); CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0)); k_ops<V><<<grid, 256>>>(out, iters, cyc); CK(cudaEventRecord(e1)); CK(cudaDeviceSynchronize());
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    long long hc[1]; CK(cudaMemcpy(hc, cyc, 8, cudaMemcpyDeviceToHost));
    double calls = (double)grid * 256 * ILP * iters;
    double per_sm_clk = (double)blocks_per_sm * 256 * ILP * iters / (double)hc[0];
    (void)per_sm_clk;
    double calls_per_sm_clk = calls / (ms * 1e-3) / sms / 1.965e9;   // wall-clock based, SM clock 1965 MHz (nvidia-smi)
    printf("%-28s blocks/SM=%d  chip %8.1f Gcalls/s => %6.2f calls/SM/clk, x%d instr = %6.1f instr/SM/clk, %.2f clk per warp-call per SMSP (%.2f ms)\n",
           name, blocks_per_sm, calls / ms / 1e6, calls_per_sm_clk, ops_per_call, calls_per_sm_clk * ops_per_call, 128.0 / calls_per_sm_clk, ms);
    return 0;
}

int main()
{
    cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
    int sms = prop.multiProcessorCount;
    printf("device %s, %d SMs, clock %d kHz\n", prop.name, sms, prop.clockRate);
    uint32_t *out; long long *cyc;
    CK(cudaMalloc(&out, (size_t)sms * 8 * 256 * 4)); CK(cudaMalloc(&cyc, (size_t)sms * 8 * 8));
    for (int bps : {2, 6}) {
        run_ops<0>("butterfly(7)", 7, bps, sms, out, cyc);
        run_ops<1>("ibutterfly(7)", 7, bps, sms, out, cyc);
        run_ops<2>("butterfly_min(10)", 10, bps, sms, out, cyc);
        run_ops<10>("butterfly_shoup(9)", 9, bps, sms, out, cyc);
        run_ops<3>("viaddmnmx x2", 2, bps, sms, out, cyc);
        run_ops<4>("iadd3 x2", 2, bps, sms, out, cyc);
        run_ops<5>("imad x2", 2, bps, sms, out, cyc);
        run_ops<6>("lop3 x2", 2, bps, sms, out, cyc);
        run_ops<7>("shf x2", 2, bps, sms, out, cyc);
        run_ops<8>("prmt x2", 2, bps, sms, out, cyc);
        run_ops<9>("imad.wide(acc)+lop", 2, bps, sms, out, cyc);
        run_ops<11>("imad.wide+2lop", 3, bps, sms, out, cyc);
        run_ops<12>("(imad.hi+lop) x2", 4, bps, sms, out, cyc);
        run_ops<13>("imad+lop3", 2, bps, sms, out, cyc);
        run_ops<14>("add+sub modP (4)", 4, bps, sms, out, cyc);
        run_ops<15>("lea.hi x2", 2, bps, sms, out, cyc);
    }
    run_blake<0>("iadd3 adds (as compiled)", sms, out);
    run_blake<1>("all adds IMAD*one", sms, out);
    run_blake<2>("2-input adds IMAD*one", sms, out);
    {
        const unsigned long long n_rows = 1ull << 23;
        const int cols = 256;
        uint32_t *in, *o2; CK(cudaMalloc(&in, n_rows * cols * 4)); CK(cudaMalloc(&o2, n_rows * 4)); CK(cudaMemset(in, 5, n_rows * cols * 4));
        run_shape<0>("registers, one CTA per row tile", (int)(n_rows / 256), o2, in, n_rows, cols / 16, sms);
        {
            uint32_t chk[4];
            run_rot<0>("none (SHF for 12 and 7)", o2, n_rows, cols / 16, sms, chk);
            run_rot<4>("rotr7 of 1 G in 4 (20 of 320)", o2, n_rows, cols / 16, sms, chk);
            run_rot<1>("rotr7 of every second G (40)", o2, n_rows, cols / 16, sms, chk);
            run_rot<2>("rotr7 of every G (80)", o2, n_rows, cols / 16, sms, chk);
            run_rot<3>("rotr7 and rotr12 of every G (160)", o2, n_rows, cols / 16, sms, chk);
        }
        run_shape<0>("registers, persistent (4 CTAs/SM)", sms * 4, o2, in, n_rows, cols / 16, sms);
        run_shape<1>("column loads, one CTA per row tile", (int)(n_rows / 256), o2, in, n_rows, cols / 16, sms);
        run_shape<1>("column loads, persistent (4 CTAs/SM)", sms * 4, o2, in, n_rows, cols / 16, sms);
        run_shape<1>("column loads, persistent (3 CTAs/SM)", sms * 3, o2, in, n_rows, cols / 16, sms);
        CK(cudaFree(in)); CK(cudaFree(o2));
    }
    for (int bps : {8}) {
        int iters = 2048, grid = sms * bps;
        cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
        k_blake<<<grid, 256>>>(out, 16, cyc); CK(cudaDeviceSynchronize());
        CK(cudaEventRecord(e0)); k_blake<<<grid, 256>>>(out, iters, cyc); CK(cudaEventRecord(e1)); CK(cudaDeviceSynchronize());
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        long long hc; CK(cudaMemcpy(&hc, cyc, 8, cudaMemcpyDeviceToHost));
        double comp = (double)grid * 256 * iters;
        printf("blake2s compress            blocks/SM=%d  %7.3f compress/SM/clk (%.1f clk per warp-compress per SMSP)  chip %7.2f Gcompress/s (%.3f ms)\n",
               bps, (double)bps * 256 * iters / hc, (double)hc / iters / (bps * 2), comp / ms / 1e6, ms);
    }
    size_t bytes = 2ull << 30;
    uint4 *a, *b; CK(cudaMalloc(&a, bytes)); CK(cudaMalloc(&b, bytes)); CK(cudaMemset(a, 1, bytes));
    for (int rep = 0; rep < 2; rep++) {
        cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
        CK(cudaEventRecord(e0)); k_copy<<<sms * 16, 512>>>(a, b, bytes / 16); CK(cudaEventRecord(e1)); CK(cudaDeviceSynchronize());
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        printf("copy 2 GiB: %.3f ms  => %.1f GB/s (read+write)\n", ms, 2.0 * bytes / ms / 1e6);
    }
    return 0;
}
