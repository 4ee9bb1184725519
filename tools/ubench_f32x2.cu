// Throughput of packed fp32x2 arithmetic on B200 (FADD2 / FFMA2) against scalar FFMA, per SM.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_f32x2 ubench_f32x2.cu && ./ubench_f32x2
// Patterns: NACC independent accumulators per thread, ITER iterations, 8 / 16 warps per SM.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint64_t sub2(uint64_t a, uint64_t b) { uint64_t r; asm volatile("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) { uint64_t r; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }

template <int MODE, int NACC>
__global__ void k(uint64_t* out, uint64_t seed, int iters) {
    uint64_t acc[NACC], a[8];
    for (int i = 0; i < NACC; ++i) acc[i] = seed + i + threadIdx.x;
    for (int i = 0; i < 8; ++i) a[i] = seed * (i + 3) + threadIdx.x;
    float facc[NACC], fa[8];
    for (int i = 0; i < NACC; ++i) facc[i] = float(i + threadIdx.x);
    for (int i = 0; i < 8; ++i) fa[i] = float(seed) * 1e-9f + i;
    for (int it = 0; it < iters; ++it) {
        if (MODE == 0) {          // FFMA2 only, independent accumulators
#pragma unroll
            for (int i = 0; i < NACC; ++i) acc[i] = fma2(a[i & 7], a[(i + 1) & 7], acc[i]);
        } else if (MODE == 1) {   // FADD2 batch then FFMA2 batch (the sqdist inner loop)
            uint64_t dd[NACC];
#pragma unroll
            for (int i = 0; i < NACC; ++i) dd[i] = sub2(a[i & 7], a[(i + 3) & 7]);
#pragma unroll
            for (int i = 0; i < NACC; ++i) acc[i] = fma2(dd[i], dd[i], acc[i]);
        } else if (MODE == 2) {   // scalar FFMA, independent accumulators
#pragma unroll
            for (int i = 0; i < NACC; ++i) facc[i] = fmaf(fa[i & 7], fa[(i + 1) & 7], facc[i]);
        } else {                  // scalar FADD batch then FFMA batch
            float dd[NACC];
#pragma unroll
            for (int i = 0; i < NACC; ++i) dd[i] = fa[i & 7] - fa[(i + 3) & 7];
#pragma unroll
            for (int i = 0; i < NACC; ++i) facc[i] = fmaf(dd[i], dd[i], facc[i]);
            fa[it & 7] += 1e-7f;
        }
    }
    uint64_t s = 0;
    for (int i = 0; i < NACC; ++i) s += acc[i] + uint64_t(__float_as_uint(facc[i]));
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// the sqdist inner loop as it is: 64 packed accumulators, a[8] resident, b streamed from shared memory
__global__ void k64(uint64_t* out, uint64_t seed, int iters) {
    __shared__ uint64_t sb[8 * 32 * 8];
    for (int i = threadIdx.x; i < 8 * 32 * 8; i += blockDim.x) sb[i] = seed * i;
    __syncthreads();
    uint64_t acc[64], a[8];
    for (int i = 0; i < 64; ++i) acc[i] = 0;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) a[i] = sb[((it + i) & 7) * 256 + (warp & 7) * 32 + lane];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const uint64_t b = sb[((it + j + 3) & 7) * 256 + (warp & 7) * 32 + lane];
            uint64_t dd[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) dd[i] = sub2(a[i], b);
#pragma unroll
            for (int i = 0; i < 8; ++i) acc[i * 8 + j] = fma2(dd[i], dd[i], acc[i * 8 + j]);
        }
    }
    uint64_t s = 0;
    for (int i = 0; i < 64; ++i) s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

void run64(int warps) {
    uint64_t* out;
    cudaMalloc(&out, 148 * 1024 * 8);
    const int iters = 4000;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k64<<<148, warps * 32>>>(out, 12345, 10);
    cudaEventRecord(e0);
    k64<<<148, warps * 32>>>(out, 12345, iters);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double instr_per_thread = double(iters) * 128;
    const double clk = ms * 1e-3 * 1.965e9;
    printf("sqdist inner loop (64 acc, 16 LDS.64 per 128 packed)  warps/SM %2d : %6.1f fp32 lane-ops/clk/SM, %5.2f clk per packed instruction per scheduler\n",
           warps, instr_per_thread * warps * 64 / clk, clk / (instr_per_thread * warps / 4.0));
    cudaFree(out);
}

template <int MODE, int NACC>
void run(const char* name, int warps) {
    uint64_t* out;
    cudaMalloc(&out, 148 * 1024 * 8);
    const int iters = 20000;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE, NACC><<<148, warps * 32>>>(out, 12345, 100);
    cudaEventRecord(e0);
    k<MODE, NACC><<<148, warps * 32>>>(out, 12345, iters);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double instr_per_thread = double(iters) * NACC * ((MODE == 1 || MODE == 3) ? 2 : 1);
    const double lane_ops = instr_per_thread * warps * 32 * ((MODE <= 1) ? 2 : 1);   // fp32 operations (an fma counts once)
    const double clk = ms * 1e-3 * 1.965e9;
    printf("%-34s warps/SM %2d  NACC %2d : %6.1f fp32 lane-ops/clk/SM, %5.2f clk per warp instruction per scheduler\n", name, warps, NACC,
           lane_ops / clk, clk / (instr_per_thread * warps / 4.0));
    cudaFree(out);
}

int main() {
    for (int w : {4, 8, 16}) {
        run<0, 16>("FFMA2 independent", w);
        run<1, 8>("FADD2 x8 then FFMA2 x8", w);
        run<1, 16>("FADD2 x16 then FFMA2 x16", w);
        run<2, 16>("FFMA scalar independent", w);
        run<3, 8>("FADD x8 then FFMA x8 scalar", w);
    }
    for (int w : {4, 8}) run64(w);
    return 0;
}
