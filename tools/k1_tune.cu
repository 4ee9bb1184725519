// Tuning harness for K1 (column-wise median / trimmed mean): times flagg_colwise_select
// of one or more libflagg builds through the C ABI on a synthetic n x d matrix.
//   k1_tune <n> <d> <mode 0|1> <k> <lib.so> [<lib.so> ...]
// Prints ms per launch (CUDA events, 5 launches after 2 warm-ups) and a checksum of the
// output, which must agree between builds.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

__device__ __forceinline__ unsigned hash32(unsigned x) {
    x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
    return x;
}
// rows = theta + delta_i like run_experiments.py:238 (full weights): theta ~ U(-0.1,0.1), delta ~ U(-1e-3,1e-3)
__global__ void fill(float* X, int n, int64_t d) {
    for (int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; c < d; c += (int64_t)gridDim.x * blockDim.x) {
        const float theta = (hash32((unsigned)c * 2654435761u + 17u) >> 8) * (0.2f / 16777216.f) - 0.1f;
        for (int i = 0; i < n; ++i) {
            const float dl = (hash32((unsigned)c * 40503u + (unsigned)i * 2246822519u + 3u) >> 8) * (2e-3f / 16777216.f) - 1e-3f;
            X[(size_t)i * d + c] = theta + (i < n / 10 ? 10.f * dl : dl);
        }
    }
}
__global__ void checksum(const float* o, int64_t d, unsigned long long* s) {
    unsigned long long a = 0;
    for (int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; c < d; c += (int64_t)gridDim.x * blockDim.x)
        a += (unsigned long long)__float_as_uint(o[c]) * (unsigned long long)(c % 1000003 + 1);
    atomicAdd(s, a);
}

typedef int (*view_create_t)(void**, int, int, const int64_t*, const void* const*, void*);
typedef int (*view_destroy_t)(void*);
typedef int (*select_t)(const void*, int, int, float*, void*);
typedef const char* (*err_t)(void);

int main(int argc, char** argv) {
    if (argc < 6) { printf("usage: k1_tune n d mode k lib...\n"); return 2; }
    const int n = atoi(argv[1]);
    const int64_t d = atoll(argv[2]);
    const int mode = atoi(argv[3]), k = atoi(argv[4]);
    float *X, *out;
    unsigned long long* cs;
    CK(cudaMalloc(&X, sizeof(float) * (size_t)n * d));
    CK(cudaMalloc(&out, sizeof(float) * d));
    CK(cudaMalloc(&cs, 8));
    fill<<<148 * 8, 256>>>(X, n, d);
    CK(cudaDeviceSynchronize());
    std::vector<const void*> ptrs(n);
    for (int i = 0; i < n; ++i) ptrs[i] = X + (size_t)i * d;
    for (int a = 5; a < argc; ++a) {
        void* h = dlopen(argv[a], RTLD_NOW | RTLD_LOCAL);
        if (!h) { printf("%s: %s\n", argv[a], dlerror()); continue; }
        view_create_t vc = (view_create_t)dlsym(h, "flagg_view_create");
        view_destroy_t vd = (view_destroy_t)dlsym(h, "flagg_view_destroy");
        select_t sel = (select_t)dlsym(h, "flagg_colwise_select");
        err_t le = (err_t)dlsym(h, "flagg_last_error");
        void* view = nullptr;
        if (vc(&view, n, 1, &d, ptrs.data(), nullptr) != 0) { printf("view: %s\n", le()); return 1; }
        CK(cudaMemset(out, 0, sizeof(float) * d));
        for (int w = 0; w < 2; ++w)
            if (sel(view, mode, k, out, nullptr) != 0) { printf("select: %s\n", le()); return 1; }
        CK(cudaDeviceSynchronize());
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0); cudaEventCreate(&e1);
        cudaEventRecord(e0);
        for (int it = 0; it < 5; ++it) sel(view, mode, k, out, nullptr);
        cudaEventRecord(e1);
        CK(cudaEventSynchronize(e1));
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        ms /= 5;
        CK(cudaMemset(cs, 0, 8));
        checksum<<<148 * 4, 256>>>(out, d, cs);
        unsigned long long hcs;
        CK(cudaMemcpy(&hcs, cs, 8, cudaMemcpyDeviceToHost));
        printf("%-44s n=%d d=%lld mode=%d k=%d  %8.3f ms  %7.1f GB/s (alg 4d(n+1))  checksum=%016llx\n", argv[a], n,
               (long long)d, mode, k, ms, 4.0 * d * (n + 1) / ms / 1e6, hcs);
        fflush(stdout);
        vd(view);
        dlclose(h);
    }
    return 0;
}
