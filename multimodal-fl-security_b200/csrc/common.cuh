// Shared declarations for libflagg.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include <vector>

#include "flagg.h"

namespace flagg {

void set_error(const char* fmt, ...);

#define FLAGG_CUDA(call)                                                              \
    do {                                                                              \
        cudaError_t e__ = (call);                                                     \
        if (e__ != cudaSuccess) {                                                     \
            flagg::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call,            \
                             cudaGetErrorString(e__));                                \
            return FLAGG_ECUDA;                                                       \
        }                                                                             \
    } while (0)

#define FLAGG_REQUIRE(cond, ...)                                                      \
    do {                                                                              \
        if (!(cond)) {                                                                \
            flagg::set_error(__VA_ARGS__);                                            \
            return FLAGG_EINVAL;                                                      \
        }                                                                             \
    } while (0)

// Device-side description of the n x P segmented client-update matrix.
struct ViewDev {
    const float* const* seg_ptrs;  // [n*P]  seg_ptrs[i*P+p]
    const int64_t* seg_start;      // [P+1]  column prefix; seg_start[P] == d
    const int64_t* unit_start[4];  // [P+1] each: prefix of UNIT_COLS[g]-wide column units per segment
    int n, P;
    int64_t d;
    // CUtensorMap[P] (64 B each, device memory) when every tensor is one dense [n x len] matrix
    // with a 16-byte-multiple row pitch -- box {128 columns, n rows}; nullptr otherwise.
    const void* tmaps;
    int dense_cols;                // columns per tmaps box: 256 (n <= 110) or 128
    // Separately allocated rows (the reference's list-of-lists, run_experiments.py:238): ONE CUtensorMap
    // over the address range [min pointer, max end) seen as a 2-D fp32 matrix of 128-column (512-byte)
    // rows, and seg_rows[p][i] = (seg_ptrs[i*P+p] - min pointer) / 512 ([P][n rounded up to 4], padded with client n-1).  A 128-column unit of any
    // tensor of any client is then one row of that matrix, and cp.async.bulk.tensor.2d.tile::gather4
    // brings four clients' units per TMA request.  nullptr unless every tensor base is 512-byte aligned
    // (what PyTorch's allocator returns for separately allocated tensors) and the range fits 2^31 rows.
    const void* gmap;
    const int32_t* seg_rows;
    int gather_cols;               // columns per gmap row box: 256 (rows overlap: pitch 512 B, extent 1 KB) or 128
    // Round-epilogue form (flagg_view_set_outputs): when non-NULL, kernels that produce a [d] result
    // write tensor p of it to out_ptrs[p] (the global model's own parameter tensors) instead of
    // out + seg_start[p] -- the inverse of seg_ptrs; replaces the per-tensor param.copy_(new_val)
    // loop at experiments/run_experiments.py:255-259.
    float* const* out_ptrs;
};

// Where column c of tensor p of a [d] result goes.
__device__ __forceinline__ float* out_addr(const ViewDev& v, float* out, int p, int64_t c) {
    return v.out_ptrs ? v.out_ptrs[p] + c : out + v.seg_start[p] + c;
}

// Work is indexed in "units": runs of UNIT_COLS[g] columns that never cross a segment
// (tensor) boundary.  Three granularities serve the three kernel families.
constexpr int NUM_GRAN = 4;
constexpr int GRAN_128 = 0, GRAN_1K = 1, GRAN_2K = 2, GRAN_256 = 3;
__host__ __device__ constexpr int unit_cols(int g) { return g == 0 ? 128 : (g == 1 ? 1024 : (g == 2 ? 2048 : 256)); }

int device_sm_count();

}  // namespace flagg

struct flagg_view {
    flagg::ViewDev dev;            // device pointers inside
    std::vector<int64_t> seg_numel, seg_start;              // host copies
    std::vector<int64_t> unit_start[flagg::NUM_GRAN];
    std::vector<const void*> seg_ptrs;                      // host copy [n*P]
    std::vector<char> upload;                               // host image of d_tables (one H2D copy at creation)
    void* d_tables = nullptr;      // one allocation holding all device tables
    void* d_out_ptrs = nullptr;    // [P] slot inside d_tables for flagg_view_set_outputs
    std::vector<void*> out_ptrs;   // host copy of the output table (empty = flat output)
    void* d_scratch = nullptr;     // weights / row indices for row_combine, partials
    size_t scratch_bytes = 0;
    int align_bytes = 4;           // min alignment over all segment base addresses (4/8/16)
    bool dense = false;            // dev.tmaps is valid (see ViewDev::tmaps)
    bool gather = false;           // dev.gmap / dev.seg_rows are valid (see ViewDev::gmap)
    int64_t total_units[flagg::NUM_GRAN] = {0, 0, 0, 0};
    int device = 0;
    void* last_stream = nullptr;   // stream of the latest call on this view: frees are ordered on it
    int ensure_scratch(size_t bytes);
};

// ---- small device helpers -------------------------------------------------------
namespace flagg {

// Locate the segment holding unit `u` (unit_start is a prefix array of length P+1).
__device__ __forceinline__ int find_segment(const int64_t* __restrict__ unit_start, int P, int64_t u) {
    int lo = 0, hi = P;  // invariant: unit_start[lo] <= u < unit_start[hi]
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (unit_start[mid] <= u) lo = mid; else hi = mid;
    }
    return lo;
}

// Order-preserving float -> int32 key: signed compare of keys == IEEE order with
// -0 < +0; NaN (any sign/payload) -> INT_MAX so it sorts last like torch.sort.
__device__ __forceinline__ int f2key(float x) {
    int b = __float_as_int(x);
    int k = b ^ ((b >> 31) & 0x7fffffff);
    return (x != x) ? 0x7fffffff : k;
}
__device__ __forceinline__ float key2f(int k) {
    return __int_as_float(k ^ ((k >> 31) & 0x7fffffff));
}

__device__ __forceinline__ float ldg_stream(const float* p) {
    float v;
    asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ float4 ldg_stream4(const float4* p) {
    float4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}

}  // namespace flagg
