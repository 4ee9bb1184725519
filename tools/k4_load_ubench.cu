// K4 operand-feed microbenchmark (GPU box only): 144 CTAs x 256 threads stream the same global
// access pattern as the Gram producers (CTA = tile (bi,bj) x K split; thread = 2 A rows + 2 B
// rows x 8 columns per 32-column k-block) with DEPTH k-blocks of loads in flight per thread and
// nothing else.  Answers: what load latency / throughput does the pattern see, and how many
// k-blocks must be in flight for one k-block per 768 clk (the tf32x3 MMA time)?
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o /tmp/k4_load_ubench tools/k4_load_ubench.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

__device__ __forceinline__ void ld8(const float* p, float (&v)[8]) {
    asm volatile("ld.global.nc.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7]) : "l"(p));
}
__device__ __forceinline__ void ld44(const float* p, float (&v)[8]) {
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]) : "l"(p));
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7]) : "l"(p + 4));
}

template <int DEPTH, bool V8, bool PF>
__global__ void __launch_bounds__(256, 1) feed(const float* X, int n, int64_t d, int nb, int64_t cols_per_split, float* out, int pf_dist) {
    int t = blockIdx.x, bi = 0;
    while (t >= nb - bi) { t -= nb - bi; ++bi; }
    const int bj = bi + t;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rloc = 16 * warp + (lane >> 2), q = lane & 3;
    const float* rows[4];
    for (int h = 0; h < 2; ++h) {
        rows[h] = X + int64_t(min(bi * 128 + rloc + 8 * h, n - 1)) * d;
        rows[2 + h] = X + int64_t(min(bj * 128 + rloc + 8 * h, n - 1)) * d;
    }
    const int64_t c0 = int64_t(blockIdx.y) * cols_per_split, c1 = min(d, c0 + cols_per_split);
    const int nkb = int((c1 - c0) / 32);
    float buf[DEPTH][4][8];
    float acc = 0.f;
    // one row per thread for the L2 prefetch: 256 threads <-> 128 A rows + 128 B rows
    const float* pfrow = X + int64_t(min((threadIdx.x < 128 ? bi * 128 + threadIdx.x : bj * 128 + threadIdx.x - 128), n - 1)) * d;
    auto fetch = [&](int slot, int kb) {
        if (kb < nkb) {
            const int64_t col = c0 + int64_t(kb) * 32 + q * 8;
#pragma unroll
            for (int r = 0; r < 4; ++r) { if (V8) ld8(rows[r] + col, buf[slot][r]); else ld44(rows[r] + col, buf[slot][r]); }
            if (PF && kb + pf_dist < nkb) asm volatile("prefetch.global.L2 [%0];" ::"l"(pfrow + c0 + int64_t(kb + pf_dist) * 32));
        }
    };
#pragma unroll
    for (int s = 0; s < DEPTH; ++s) fetch(s, s);
    for (int kb = 0; kb < nkb; kb += DEPTH) {
#pragma unroll
        for (int s = 0; s < DEPTH; ++s) {
            if (kb + s < nkb) {
#pragma unroll
                for (int r = 0; r < 4; ++r)
#pragma unroll
                    for (int j = 0; j < 8; ++j) acc += buf[s][r][j];
                fetch(s, kb + s + DEPTH);
            }
        }
    }
    if (acc == 12345.678f) out[0] = acc;
}

template <int DEPTH, bool V8, bool PF>
void run(const float* X, int n, int64_t d, float* out, int pf_dist, const char* tag) {
    const int nb = (n + 127) / 128, tiles = nb * (nb + 1) / 2, splits = 148 / tiles > 0 ? 148 / tiles : 1;
    const int64_t cps = ((d / splits) + 127) / 128 * 128;
    dim3 grid(tiles, splits);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    feed<DEPTH, V8, PF><<<grid, 256>>>(X, n, d, nb, cps, out, pf_dist);
    cudaEventRecord(e0);
    for (int i = 0; i < 3; ++i) feed<DEPTH, V8, PF><<<grid, 256>>>(X, n, d, nb, cps, out, pf_dist);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= 3;
    const double kb_per_cta = double(cps) / 32, clk_per_kb = ms * 1e-3 * 1.9e9 / kb_per_cta;
    printf("%-28s n=%d d=%lld: %.2f ms  %.0f clk/k-block (target 768)  L1<-L2 %.2f TB/s  (%s)\n", tag, n, (long long)d, ms, clk_per_kb,
           double(tiles) * splits * kb_per_cta * 32768 / ms / 1e9, cudaGetErrorString(cudaGetLastError()));
}

int main(int argc, char** argv) {
    const int n = 1000;
    for (int64_t d : {int64_t(100000), int64_t(3125000)}) {
        float *X, *out;
        cudaMalloc(&X, sizeof(float) * n * d); cudaMalloc(&out, 4);
        cudaMemset(X, 0, sizeof(float) * n * d);
        run<1, true, false>(X, n, d, out, 0, "depth1 v8");
        run<2, true, false>(X, n, d, out, 0, "depth2 v8");
        run<3, true, false>(X, n, d, out, 0, "depth3 v8");
        run<4, true, false>(X, n, d, out, 0, "depth4 v8");
        run<2, false, false>(X, n, d, out, 0, "depth2 v4x2");
        run<4, false, false>(X, n, d, out, 0, "depth4 v4x2");
        run<2, true, true>(X, n, d, out, 8, "depth2 v8 +L2 prefetch 8");
        run<2, true, true>(X, n, d, out, 32, "depth2 v8 +L2 prefetch 32");
        run<1, true, true>(X, n, d, out, 16, "depth1 v8 +L2 prefetch 16");
        cudaFree(X); cudaFree(out);
    }
    return 0;
}
