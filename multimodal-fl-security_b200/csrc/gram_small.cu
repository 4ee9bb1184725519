// K4 for small cohorts (n <= 32): the reference's default regime (10-20 clients, krum.py:92-97).
//
// With so few rows the Gram matrix is a tall-skinny contraction: 4n bytes and n(n+1)/2 products per
// column, i.e. HBM-bound up to n ~ 20 and near the FP32 FMA rate at n = 32 -- the tensor-core
// kernels (128-row tiles) and the 64x64 FFMA kernel both waste most of their tile here.
//
// Layout of the work: rows are cut into blocks of 8; a warp owns one pair of row blocks
// (bi <= bj: an 8x8 register tile of fp32 accumulators per column, 36 of them on the diagonal) and
// one column slot.  The CTA streams 128-column units through a 4-stage cp.async ring in shared
// memory (n rows + the centre row per unit, `cs` units per stage; bytes beyond a tensor's end are
// zero-filled by the copy), so the loads in flight are bounded by shared memory, not registers.
// A warp reads its 8 + 8 rows of one unit as 128-bit shared loads (a lane owns 4 consecutive
// columns), centres them and issues 256 FMAs.
// A lane accumulates at most GS_MAX_LANE_COLS columns in fp32 (the split-K count is sized for
// that); the per-warp tile is then reduced across lanes in FP64 and written as a partial, and
// gram_small_reduce_kernel sums the partials in a fixed order (deterministic).
#include <algorithm>

#include "common.cuh"

namespace flagg {

constexpr int GS_RB = 8;                  // rows per block
constexpr int GS_MAX_N = 32;
constexpr int GS_MAX_LANE_COLS = 2048;    // fp32 accumulation length per lane
constexpr int GS_MAX_WARPS = 12;

struct GsShape {
    int nblk, tp, cs, warps;              // row blocks, block pairs, column slots, warps per CTA
};
__host__ __device__ inline GsShape gs_shape(int n) {
    GsShape s;
    s.nblk = (n + GS_RB - 1) / GS_RB;
    s.tp = s.nblk * (s.nblk + 1) / 2;                       // 1, 3, 6, 10
    // column slots: CTAs of 4-6 warps run two to an SM (independent barrier domains hide each other's
    // stalls: a diagonal warp has 144 FMAs per unit, an off-diagonal one 256), the 10-warp shape runs alone
    s.cs = s.tp == 1 ? 4 : s.tp == 3 ? 2 : 1;
    s.warps = s.tp * s.cs;                                  // 4, 6, 6, 10
    return s;
}

constexpr int GS_STAGES = 4;
constexpr int GS_VEC_ITEMS = 4;            // 16-byte copies per thread per unit: ceil(32 n / threads) <= 4 for every shape
__host__ __device__ inline int gs_rows_pad(int n) { return (n + GS_RB - 1) / GS_RB * GS_RB; }
__host__ __device__ inline size_t gs_stage_floats(int n, int cs) { return size_t(cs) * (gs_rows_pad(n) + 1) * 128; }

template <int BYTES>
__device__ __forceinline__ void gs_cp_async(uint32_t dst, const void* src, int valid_bytes) {
    if constexpr (BYTES == 16)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(valid_bytes) : "memory");
    else
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst), "l"(src), "r"(valid_bytes) : "memory");
}

// grid = splits; block = 32 * warps.  part[(cta * warps + warp) * 64 + i * 8 + j]
// VEC: every row pointer is 16-byte aligned (128-bit copies); otherwise 4-byte copies.
template <bool VEC>
__global__ void __launch_bounds__(32 * GS_MAX_WARPS, 1)
gram_small_kernel(ViewDev v, const float* __restrict__ center, int64_t total_units, int64_t units_per_cta, double* __restrict__ part) {
    extern __shared__ __align__(16) float gs_smem[];
    const GsShape sh = gs_shape(v.n);
    const int rows_pad = gs_rows_pad(v.n);
    const int unit_floats = (rows_pad + 1) * 128;                 // rows [0, rows_pad) + the centre row
    const int stage_floats = sh.cs * unit_floats;
    const int nthreads = 32 * sh.warps;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tp = warp % sh.tp, slot = warp / sh.tp;
    int t = tp, bi = 0;
    while (t >= sh.nblk - bi) { t -= sh.nblk - bi; ++bi; }
    const int bj = bi + t;
    const bool diag = bi == bj;
    const int64_t u0 = int64_t(blockIdx.x) * units_per_cta;
    const int64_t u1 = min(total_units, u0 + units_per_cta);
    const int iters = int((max(u1 - u0, int64_t(0)) + sh.cs - 1) / sh.cs);
    const int64_t* __restrict__ us = v.unit_start[GRAN_128];
    const uint32_t smem0 = uint32_t(__cvta_generic_to_shared(gs_smem));

    // loader state: segment of the first unit of the next stage to issue (advances monotonically)
    int lp = (u0 < u1) ? find_segment(us, v.P, u0) : 0;
    constexpr int CH = VEC ? 4 : 1;                               // floats per copy
    constexpr int CPR = 128 / CH;                                 // copies per row
    const float* cptr[GS_VEC_ITEMS];
    int cp_seg = -1;
    bool c_ok[GS_VEC_ITEMS];
    int c_row[GS_VEC_ITEMS], c_q4[GS_VEC_ITEMS];
    uint32_t c_dst[GS_VEC_ITEMS];
#pragma unroll
    for (int i = 0; i < GS_VEC_ITEMS; ++i) {
        const int item = threadIdx.x + i * nthreads;
        c_ok[i] = VEC && item < v.n * 32;
        c_row[i] = item / 32;
        c_q4[i] = (item % 32) * 4;
        c_dst[i] = uint32_t(c_row[i] * 128 + c_q4[i]) * 4u;
    }
    auto issue = [&](int k) {                                     // stage k -> ring slot k % GS_STAGES
        if (k < iters) {
            const uint32_t sbase = smem0 + uint32_t((k % GS_STAGES) * stage_floats) * 4u;
            int pj = lp;
            for (int j = 0; j < sh.cs; ++j) {
                const int64_t u = u0 + int64_t(k) * sh.cs + j;
                if (u >= u1) break;
                while (us[pj + 1] <= u) ++pj;
                if (j == 0) lp = pj;
                const int64_t c0 = (u - us[pj]) * 128;
                const int len = int(min((long long)128, (long long)(v.seg_start[pj + 1] - v.seg_start[pj] - c0)));
                const uint32_t ubase = sbase + uint32_t(j * unit_floats) * 4u;
                if constexpr (VEC) {
                    // a thread copies the same (row, chunk) items of every unit: row pointers cached per tensor,
                    // shared-memory offsets and chunk columns fixed for the whole kernel
                    if (pj != cp_seg) {
#pragma unroll
                        for (int i = 0; i < GS_VEC_ITEMS; ++i)
                            if (c_ok[i]) cptr[i] = v.seg_ptrs[size_t(c_row[i]) * v.P + pj] + c_q4[i];
                        cp_seg = pj;
                    }
                    if (len == 128) {
#pragma unroll
                        for (int i = 0; i < GS_VEC_ITEMS; ++i)
                            if (c_ok[i]) gs_cp_async<16>(ubase + c_dst[i], cptr[i] + c0, 16);
                    } else {
#pragma unroll
                        for (int i = 0; i < GS_VEC_ITEMS; ++i)
                            if (c_ok[i]) {
                                const int nvalid = max(0, min(4, len - c_q4[i]));
                                gs_cp_async<16>(ubase + c_dst[i], cptr[i] + c0 - (nvalid > 0 ? 0 : c_q4[i]), 4 * nvalid);
                            }
                    }
                } else {
                    for (int item = threadIdx.x; item < v.n * CPR; item += nthreads) {
                        const int q = item % CPR, rr = item / CPR;
                        const float* rowptr = v.seg_ptrs[size_t(rr) * v.P + pj] + c0;
                        const int nvalid = max(0, min(CH, len - q * CH));
                        gs_cp_async<4>(ubase + uint32_t(rr * 128 + q * CH) * 4u, rowptr + (nvalid > 0 ? q * CH : 0), 4 * nvalid);
                    }
                }
                // the centre row: 4-byte copies (center + seg_start[p] is only 4-byte aligned in general); zero-filled without a centre
                if (threadIdx.x < 128) {
                    const int q = threadIdx.x;
                    const bool ok = center != nullptr && q < len;
                    const float* src = ok ? center + v.seg_start[pj] + c0 + q : v.seg_ptrs[0];
                    gs_cp_async<4>(ubase + uint32_t(rows_pad * 128 + q) * 4u, src, ok ? 4 : 0);
                }
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };

    float acc[GS_RB][GS_RB];
#pragma unroll
    for (int i = 0; i < GS_RB; ++i)
#pragma unroll
        for (int j = 0; j < GS_RB; ++j) acc[i][j] = 0.f;

#pragma unroll 1
    for (int k = 0; k < GS_STAGES - 1; ++k) issue(k);
#pragma unroll 1
    for (int k = 0; k < iters; ++k) {
        asm volatile("cp.async.wait_group %0;" ::"n"(GS_STAGES - 2) : "memory");
        __syncthreads();                                          // stage k has landed; everyone is done with stage k - 1
        issue(k + GS_STAGES - 1);                                 // refills the slot stage k - 1 used
        if (u0 + int64_t(k) * sh.cs + slot < u1) {
            const float* ub = gs_smem + size_t(k % GS_STAGES) * stage_floats + size_t(slot) * unit_floats;
            const float4 c4 = *reinterpret_cast<const float4*>(ub + rows_pad * 128 + 4 * lane);
            float4 a[GS_RB], b[GS_RB];
#pragma unroll
            for (int i = 0; i < GS_RB; ++i) {
                a[i] = *reinterpret_cast<const float4*>(ub + (bi * GS_RB + i) * 128 + 4 * lane);
                a[i].x -= c4.x; a[i].y -= c4.y; a[i].z -= c4.z; a[i].w -= c4.w;
            }
            if (diag) {
#pragma unroll
                for (int i = 0; i < GS_RB; ++i)
#pragma unroll
                    for (int j = i; j < GS_RB; ++j) {
                        acc[i][j] = fmaf(a[i].x, a[j].x, acc[i][j]);
                        acc[i][j] = fmaf(a[i].y, a[j].y, acc[i][j]);
                        acc[i][j] = fmaf(a[i].z, a[j].z, acc[i][j]);
                        acc[i][j] = fmaf(a[i].w, a[j].w, acc[i][j]);
                    }
            } else {
#pragma unroll
                for (int i = 0; i < GS_RB; ++i) {
                    b[i] = *reinterpret_cast<const float4*>(ub + (bj * GS_RB + i) * 128 + 4 * lane);
                    b[i].x -= c4.x; b[i].y -= c4.y; b[i].z -= c4.z; b[i].w -= c4.w;
                }
#pragma unroll
                for (int i = 0; i < GS_RB; ++i)
#pragma unroll
                    for (int j = 0; j < GS_RB; ++j) {
                        acc[i][j] = fmaf(a[i].x, b[j].x, acc[i][j]);
                        acc[i][j] = fmaf(a[i].y, b[j].y, acc[i][j]);
                        acc[i][j] = fmaf(a[i].z, b[j].z, acc[i][j]);
                        acc[i][j] = fmaf(a[i].w, b[j].w, acc[i][j]);
                    }
            }
        }
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    // lanes -> one FP64 tile per warp (butterfly: every lane ends with the same sum)
    double* dst = part + (size_t(blockIdx.x) * sh.warps + warp) * (GS_RB * GS_RB);
#pragma unroll
    for (int i = 0; i < GS_RB; ++i)
#pragma unroll
        for (int j = 0; j < GS_RB; ++j) {
            double s = double(acc[i][j]);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            if (lane == ((i * GS_RB + j) & 31)) dst[i * GS_RB + j] = s;
        }
}

// One warp per output element (r <= c): lanes stride over the (cta, slot) partials, then a
// fixed-shape shuffle tree -- the same order every run.
__global__ void __launch_bounds__(256)
gram_small_reduce_kernel(const double* __restrict__ part, int ctas, int n, int accumulate, double* __restrict__ G) {
    const GsShape sh = gs_shape(n);
    const int lane = threadIdx.x & 31;
    const int e = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (e >= n * n) return;
    const int r = e / n, c = e % n;
    if (c < r) return;
    const int bi = r / GS_RB, bj = c / GS_RB;
    int tp = 0;
    for (int b = 0; b < bi; ++b) tp += sh.nblk - b;
    tp += bj - bi;
    const int idx = (r % GS_RB) * GS_RB + (c % GS_RB);
    const int terms = ctas * sh.cs;
    double s = 0.0;
    for (int k = lane; k < terms; k += 32) {
        const int cta = k / sh.cs, slot = k % sh.cs;
        s += part[(size_t(cta) * sh.warps + slot * sh.tp + tp) * (GS_RB * GS_RB) + idx];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) {
        if (accumulate) {
            G[size_t(r) * n + c] += s;
            if (r != c) G[size_t(c) * n + r] += s;
        } else {
            G[size_t(r) * n + c] = s;
            if (r != c) G[size_t(c) * n + r] = s;
        }
    }
}

// Host launch; called from flagg_gram (gram.cu) for n <= GS_MAX_N.  Any alignment.
int gram_small_launch(flagg_view* v, const float* center, int accumulate, double* G, cudaStream_t st) {
    const int n = v->dev.n;
    const int64_t units = v->total_units[GRAN_128];
    FLAGG_REQUIRE(n >= 1 && n <= GS_MAX_N, "flagg_gram: the small-cohort kernel serves n <= %d (n=%d)", GS_MAX_N, n);
    const GsShape sh = gs_shape(n);
    const int sms = device_sm_count();
    FLAGG_REQUIRE(sms > 0, "cannot query the SM count of the current device");
    // CTAs: enough that a lane never accumulates more than GS_MAX_LANE_COLS columns in fp32, and a
    // multiple of the SM count (one CTA per SM: 64 accumulators + two operand sets per thread; one wave
    // when the cap allows, so the pipeline fill and the FP64 butterfly are paid once per SM)
    const int64_t max_units = std::max<int64_t>(1, int64_t(GS_MAX_LANE_COLS) * 32 * sh.cs / 128);
    int64_t ctas = (units + max_units - 1) / max_units;
    const int per_sm = sh.warps <= 6 ? 2 : 1;
    ctas = std::max<int64_t>(ctas, std::min<int64_t>(units, int64_t(sms) * per_sm));
    ctas = (ctas + sms * per_sm - 1) / (sms * per_sm) * (sms * per_sm);
    ctas = std::min<int64_t>(ctas, units);
    const int64_t upc = (units + ctas - 1) / ctas;
    ctas = (units + upc - 1) / upc;
    FLAGG_REQUIRE(upc <= max_units, "flagg_gram: internal split error");
    const size_t need = sizeof(double) * size_t(ctas) * sh.warps * GS_RB * GS_RB;
    int rc = v->ensure_scratch(need);
    if (rc != FLAGG_OK) return rc;
    double* part = static_cast<double*>(v->d_scratch);
    const size_t smem = sizeof(float) * GS_STAGES * gs_stage_floats(n, sh.cs);
    static bool attr_set = false;
    if (!attr_set) {
        FLAGG_CUDA(cudaFuncSetAttribute(gram_small_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        FLAGG_CUDA(cudaFuncSetAttribute(gram_small_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        attr_set = true;
    }
    if (v->align_bytes >= 16)
        gram_small_kernel<true><<<unsigned(ctas), 32 * sh.warps, smem, st>>>(v->dev, center, units, upc, part);
    else
        gram_small_kernel<false><<<unsigned(ctas), 32 * sh.warps, smem, st>>>(v->dev, center, units, upc, part);
    FLAGG_CUDA(cudaGetLastError());
    const int warps_per_block = 8;
    gram_small_reduce_kernel<<<(n * n + warps_per_block - 1) / warps_per_block, 32 * warps_per_block, 0, st>>>(part, int(ctas), n, accumulate, G);
    FLAGG_CUDA(cudaGetLastError());
    return FLAGG_OK;
}

}  // namespace flagg
