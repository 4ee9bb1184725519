// K4 for small cohorts (n <= 32), pairwise-difference form: D2[i][j] = ||x_i - x_j||^2.
//
// This is the reference's default regime (10-20 clients) and its own arithmetic: krum.py:92-97 forms
// torch.norm(f_i - f_j) pair by pair.  Subtracting first means there is no cancellation to manage -- no
// centre vector, hence no centre pass over X (the Gram form reads X twice: a K1 median of the rows, then
// the products) -- and NaN / Inf rows give NaN / inf distances by IEEE arithmetic exactly where the
// reference's are.  The byte floor is one read of X: 4 n d.
//
// Layout of the work: rows are cut into blocks of 8; a warp owns a "job" of at most 64 row pairs whose fp32
// accumulators stay in registers for the whole kernel, and one column slot; lane = column (32 columns per
// quarter unit):
//     O(bi, bj): rows of block bi against (a half of) the rows of block bj;  D(b..): the i<j pairs inside a block
//     n <=  8: { D(0) }                                                        x 8 column slots
//     n <= 16: { O(0,1), D(0,1) }                                              x 4
//     n <= 20: { O(0,1)a, O(0,1)b, O(0,2)a, O(1,2)a, D(0), D(1), D(2), - }     x 1   (at most 32 pairs each)
//     n <= 24: { six half O jobs, D(0,1), D(2) }                               x 1
//     n <= 32: { six O jobs, D(0,1), D(2,3) }                                  x 1
// always 8 warps; rows that do not exist are skipped at even granularity (n = 10 computes 45 pairs, not 120).
// 128 registers per thread -> two CTAs per SM, four warps per scheduler.  The CTA streams 128-column units
// through a cp.async ring in shared memory (n rows per unit; bytes past a tensor's end are zero-filled, so they
// add nothing); units are dealt round-robin over the grid so the CTAs read neighbouring columns of every row.
// A lane accumulates at most SD_MAX_ACC_COLS columns in fp32; jobs are then reduced across lanes in FP64, written
// as per-CTA partials and summed in a fixed order by sqdist_reduce_kernel (deterministic).
//
// What bounds it (profiles/r02_sqdist_*): 2 instructions per pair and column, i.e. n (n-1) = 380 thread
// instructions per column at n = 20 against ~456 issue slots per column at the HBM rate: the kernel is issue /
// FP32-pipe bound from n ~ 16 (HBM-bound below).  Packed fp32x2 arithmetic (sub.rn.f32x2 + fma.rn.f32x2, one
// instruction per pair and TWO columns) was built and measured first: with 64 live accumulators it is
// register-file bound at 2.3 clk per packed instruction per scheduler (tools/ubench_f32x2.cu), i.e. no faster per
// column than scalar code, and its 128 accumulator registers allow only one 8-warp CTA per SM (FMA pipe 47 % busy,
// stalled on latency); scalar accumulators halve the registers and measured faster at every n.
#include <algorithm>

#include "common.cuh"

namespace flagg {

constexpr int SD_RB = 8;
constexpr int SD_MAX_N = 32;
constexpr int SD_WARPS = 8;
constexpr int SD_MAX_ACC_COLS = 1024;      // fp32 accumulation length per packed lane
constexpr int SD_ACC = 64;                 // accumulators per job: 64 (O) or 2 x 28 (D)

__host__ __device__ constexpr int sd_tri(int i, int j) { return i * SD_RB - i * (i + 1) / 2 + (j - i - 1); }   // i < j < 8 -> 0..27
__host__ __device__ inline int sd_nblk(int n) { return (n + SD_RB - 1) / SD_RB; }
__host__ __device__ inline int sd_noff(int nblk) { return nblk * (nblk - 1) / 2; }
__host__ __device__ inline int sd_njobs(int nblk) { return nblk == 1 ? 1 : nblk == 2 ? 2 : 8; }
// index among the pairs of different blocks (bi < bj), enumerated (0,1), (0,2), ..., (1,2), ...
__host__ __device__ inline int sd_off_pair(int nblk, int bi, int bj) {
    int j = 0;
    for (int b = 0; b < bi; ++b) j += nblk - 1 - b;
    return j + (bj - bi - 1);
}
// Job table.  kind 0 = O: rows of block b0 against rows [j0, j0 + jn) of block b1; kind 1 = D: blocks b0 .. b0 + nd - 1.
struct SdJob { int kind, b0, b1, j0, jn, nd; };
__host__ __device__ inline SdJob sd_job(int nblk, int job, int n) {
    SdJob d = {1, 0, 0, 0, SD_RB, 1};
    if (nblk == 3 && n <= 20) {
        // 17..20 clients: the third block has at most four rows, so the "b" halves of O(0,2) and O(1,2) are empty and
        // D(0,1) can be split instead: { O(0,1)a, O(0,1)b, O(0,2)a, O(1,2)a, D(0), D(1), D(2), idle }, at most 32 pairs each
        if (job < 2) return SdJob{0, 0, 1, 4 * job, 4, 0};
        if (job < 4) return SdJob{0, job - 2, 2, 0, 4, 0};
        if (job < 7) { d.b0 = job - 4; return d; }
        return SdJob{2, 0, 0, 0, 0, 0};
    }
    if (nblk == 1) return d;                                              // D(0)
    if (nblk == 2) {
        if (job == 0) d = SdJob{0, 0, 1, 0, SD_RB, 0};                    // O(0,1)
        else d.nd = 2;                                                    // D(0,1)
        return d;
    }
    if (job < 6) {
        const int pair = nblk == 3 ? job / 2 : job;                       // three pairs in halves, or six whole pairs
        int t = pair, bi = 0;
        while (t >= nblk - 1 - bi) { t -= nblk - 1 - bi; ++bi; }
        d = SdJob{0, bi, bi + 1 + t, nblk == 3 ? 4 * (job % 2) : 0, nblk == 3 ? 4 : SD_RB, 0};
        return d;
    }
    if (nblk == 3) { d.b0 = job == 6 ? 0 : 2; d.nd = job == 6 ? 2 : 1; }  // D(0,1), D(2)
    else { d.b0 = 2 * (job - 6); d.nd = 2; }                              // D(0,1), D(2,3)
    return d;
}
// where the pair (r < c) is accumulated: job and accumulator index
__host__ __device__ inline void sd_locate(int nblk, int n, int r, int c, int& job, int& e) {
    const int bi = r / SD_RB, bj = c / SD_RB, ri = r % SD_RB, cj = c % SD_RB;
    if (nblk == 3 && n <= 20) {
        if (bi != bj) {
            job = bj == 1 ? cj / 4 : 2 + bi;
            e = ri * SD_RB + cj % 4;
        } else {
            job = 4 + bi;
            e = sd_tri(ri, cj);
        }
        return;
    }
    if (bi != bj) {
        const int pair = sd_off_pair(nblk, bi, bj);
        job = nblk == 2 ? 0 : nblk == 3 ? 2 * pair + cj / 4 : pair;
        e = ri * SD_RB + (nblk == 3 ? cj % 4 : cj);          // half jobs index their four rows 0..3
    } else {
        int t = bi;
        if (nblk <= 2) job = nblk - 1;
        else if (nblk == 3) { job = bi < 2 ? 6 : 7; t = bi < 2 ? bi : 0; }
        else { job = 6 + bi / 2; t = bi % 2; }
        e = t * 28 + sd_tri(ri, cj);
    }
}

template <int BYTES>
__device__ __forceinline__ void sd_cp_async(uint32_t dst, const void* src, int valid_bytes) {
    if constexpr (BYTES == 16)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(valid_bytes) : "memory");
    else
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst), "l"(src), "r"(valid_bytes) : "memory");
}

// One quarter unit (32 columns, lane = column) of an O job: rows a[0..8) of block b0 against JR rows of block b1
// (JR = 2, 4, 6, 8: the rows that exist, rounded up to even -- padding rows are zero and their accumulators are
// never read).
template <int JR>
__device__ __forceinline__ void sd_o_step(float (&acc)[SD_ACC], const float* pa, const float* pb) {
    float a[SD_RB];
#pragma unroll
    for (int i = 0; i < SD_RB; ++i) a[i] = pa[i * 128];
#pragma unroll
    for (int j = 0; j < JR; ++j) {
        const float b = pb[j * 128];
#pragma unroll
        for (int i = 0; i < SD_RB; ++i) {
            const float dd = a[i] - b;
            acc[i * SD_RB + j] = fmaf(dd, dd, acc[i * SD_RB + j]);
        }
    }
}
// One quarter unit of one block of a D job: the i < j pairs among its first R rows; accumulators T * 28 + tri(i, j).
template <int R, int T>
__device__ __forceinline__ void sd_d_step(float (&acc)[SD_ACC], const float* pa) {
    float a[R];
#pragma unroll
    for (int i = 0; i < R; ++i) a[i] = pa[i * 128];
#pragma unroll
    for (int i = 0; i < R - 1; ++i)
#pragma unroll
        for (int j = i + 1; j < R; ++j) {
            const float dd = a[i] - a[j];
            acc[T * 28 + sd_tri(i, j)] = fmaf(dd, dd, acc[T * 28 + sd_tri(i, j)]);
        }
}
template <int T>
__device__ __forceinline__ void sd_d_block(float (&acc)[SD_ACC], const float* pa, int r) {   // r: warp-uniform, even
    if (r >= 8) sd_d_step<8, T>(acc, pa);
    else if (r == 6) sd_d_step<6, T>(acc, pa);
    else if (r == 4) sd_d_step<4, T>(acc, pa);
    else sd_d_step<2, T>(acc, pa);
}

// grid = column splits; block = 256.  part[((cta * 8) + warp) * SD_ACC + e]
// VEC: every row pointer is 16-byte aligned (128-bit copies); otherwise 4-byte copies.
// SD_STAGES: depth of the cp.async ring: 5 stages of 32 KB for n <= 16, 10 stages of 12-16 KB above (a stage is one
// unit there; with 4 stages only 48 KB per SM were in flight and the kernel ran at the memory latency, not the rate).
// UG: consecutive units per slot and stage (2 for n > 16, where a stage would otherwise be a single unit and the
// barrier / loader bookkeeping per stage is as large as the arithmetic).
template <bool VEC, int SD_STAGES, int UG>
__global__ void __launch_bounds__(32 * SD_WARPS, 2)
sqdist_small_kernel(ViewDev v, int total_units, int units_per_cta, double* __restrict__ part) {
    extern __shared__ __align__(16) float sd_smem[];
    const int n = v.n;
    const int nblk = sd_nblk(n), njobs = sd_njobs(nblk), cs = SD_WARPS / njobs;
    const int unit_floats = nblk * SD_RB * 128;
    const int stage_floats = cs * UG * unit_floats;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // three row blocks: pair the jobs on a scheduler (warps w and w + 4) so that heavy meets light --
    // { D(0,1) + O(1,2)b, O(0,1)a + O(0,2)b, O(0,1)b + D(2), O(0,2)a + O(1,2)a }; the "b" halves are empty up to n = 20
    const int job = (nblk == 3 && n > 20) ? ((0x60125374 >> (4 * (7 - warp))) & 7) : warp % njobs;
    const int slot = warp / njobs;
    const SdJob jd = sd_job(nblk, job, n);
    const bool offdiag = jd.kind == 0;
    const int b0 = jd.b0, b1 = jd.b1, nd = jd.nd;
    // rows that exist: an O job stops at the last real row of its b1 range, a D job at the last real row of each block
    // (rounded up to even: a padding row is all zeros and the accumulators of its pairs are never read)
    const int jreal = offdiag ? (max(0, min(jd.jn, n - (b1 * SD_RB + jd.j0))) + 1) & ~1 : 0;
    const int dreal0 = (min(SD_RB, n - b0 * SD_RB) + 1) & ~1, dreal1 = (max(0, min(SD_RB, n - (b0 + 1) * SD_RB)) + 1) & ~1;
    // Units are dealt round-robin in groups of cs: stage k of CTA b covers units (k * grid + b) * cs + [0, cs), so at any
    // moment the CTAs of the grid read neighbouring columns of every row (DRAM page locality; a CTA streaming its own
    // far-apart column range opens a new page per 512-byte request: 1.7-2.3 TB/s measured against 4-5 TB/s this way).
    const int u0 = blockIdx.x * cs * UG, ustep = gridDim.x * cs * UG, u1 = total_units;
    const int iters = u0 < u1 ? (u1 - u0 + ustep - 1) / ustep : 0;
    (void)units_per_cta;
    const int64_t* __restrict__ us = v.unit_start[GRAN_128];
    const uint32_t smem0 = uint32_t(__cvta_generic_to_shared(sd_smem));

    // rows n .. 8 nblk - 1 of every unit are never copied: clear them once so the unused pairs see zeros
    for (int i = threadIdx.x; i < SD_STAGES * cs * UG * (nblk * SD_RB - n) * 128; i += blockDim.x) {
        const int per = (nblk * SD_RB - n) * 128;
        sd_smem[(i / per) * unit_floats + n * 128 + (i % per)] = 0.f;
    }

    // Loader: the warps of column slot s bring in unit s of every stage -- warp `job` of the slot copies rows
    // job, job + njobs, ... (at most 8), lane = 16-byte chunk -- so a warp issues at most 8 copies per unit it
    // computes on.  The segment (tensor) of the slot's next unit is tracked in registers and advances monotonically.
    const int lrow = warp % njobs;           // this warp copies rows lrow, lrow + njobs, ... of its slot's unit
    int lp = 0, lu0 = 0, lu1 = 0;            // segment of this slot's next unit: index, first unit, one past its last unit
    int64_t llen = 0;
    if (u0 + slot * UG < u1) {
        lp = find_segment(us, v.P, u0 + slot * UG);
        lu0 = int(us[lp]);
        lu1 = int(us[lp + 1]);
        llen = v.seg_start[lp + 1] - v.seg_start[lp];
    }
    auto issue = [&](int k) {                                     // stage k -> ring slot k % SD_STAGES
#pragma unroll
        for (int g = 0; g < UG; ++g) {
        const int u = u0 + k * ustep + slot * UG + g;
        if (k < iters && u < u1) {
            while (u >= lu1) {
                ++lp;
                lu0 = lu1;
                lu1 = int(us[lp + 1]);
                llen = v.seg_start[lp + 1] - v.seg_start[lp];
            }
            const int64_t c0 = int64_t(u - lu0) * 128;
            const int len = int(min((long long)128, (long long)(llen - c0)));
            const uint32_t ubase = smem0 + uint32_t((k % SD_STAGES) * stage_floats + (slot * UG + g) * unit_floats) * 4u;
            if constexpr (VEC) {
                const int q4 = lane * 4;
                const int nvalid = max(0, min(4, len - q4));
                const int64_t off = c0 + (nvalid > 0 ? q4 : 0);
                const float* const* pp = v.seg_ptrs + lp;
                const uint32_t dst = ubase + uint32_t(q4) * 4u;
#pragma unroll 2
                for (int rr = lrow; rr < n; rr += njobs)       // 1-8 rows: a real loop, predicated-off copies still cost issue slots
                    sd_cp_async<16>(dst + uint32_t(rr) * 512u, pp[size_t(rr) * v.P] + off, 4 * nvalid);
            } else {
                for (int rr = lrow; rr < n; rr += njobs) {
                    const float* src = v.seg_ptrs[size_t(rr) * v.P + lp] + c0;
#pragma unroll
                    for (int q = lane; q < 128; q += 32)
                        sd_cp_async<4>(ubase + uint32_t(rr * 128 + q) * 4u, src + (q < len ? q : 0), q < len ? 4 : 0);
                }
            }
        }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };

    float acc[SD_ACC];
#pragma unroll
    for (int e = 0; e < SD_ACC; ++e) acc[e] = 0.f;

#pragma unroll 1
    for (int k = 0; k < SD_STAGES - 1; ++k) issue(k);
#pragma unroll 1
    for (int k = 0; k < iters; ++k) {
        asm volatile("cp.async.wait_group %0;" ::"n"(SD_STAGES - 2) : "memory");
        __syncthreads();                                          // stage k has landed; everyone is done with stage k - 1
        issue(k + SD_STAGES - 1);                                 // refills the slot stage k - 1 used
#pragma unroll 1
        for (int g = 0; g < UG; ++g)
        if (u0 + k * ustep + slot * UG + g < u1) {
            const float* ub = sd_smem + size_t(k % SD_STAGES) * stage_floats + size_t(slot * UG + g) * unit_floats + lane;
            if (offdiag) {
                if (jreal > 0) {
#pragma unroll
                    for (int h = 0; h < 4; ++h) {
                        const float* pa = ub + (b0 * SD_RB) * 128 + 32 * h;
                        const float* pb = ub + (b1 * SD_RB + jd.j0) * 128 + 32 * h;
                        if (jreal >= 8) sd_o_step<8>(acc, pa, pb);
                        else if (jreal == 6) sd_o_step<6>(acc, pa, pb);
                        else if (jreal == 4) sd_o_step<4>(acc, pa, pb);
                        else sd_o_step<2>(acc, pa, pb);
                    }
                }
            } else if (jd.kind == 1) {
#pragma unroll
                for (int h = 0; h < 4; ++h) {
                    sd_d_block<0>(acc, ub + (b0 * SD_RB) * 128 + 32 * h, dreal0);
                    if (nd > 1) sd_d_block<1>(acc, ub + ((b0 + 1) * SD_RB) * 128 + 32 * h, dreal1);
                }
            }
        }
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    // lanes -> one FP64 value per accumulator (butterfly: every lane ends with the same sum)
    double* dst = part + (size_t(blockIdx.x) * SD_WARPS + slot * njobs + job) * SD_ACC;
#pragma unroll
    for (int e = 0; e < SD_ACC; ++e) {
        double s = double(acc[e]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == (e & 31)) dst[e] = s;
    }
}

// One CTA per pair (r < c): 128 threads stride over the (cta, slot) partials, then a fixed-shape shuffle tree and a
// fixed-order sum over the four warps -- the same order every run.  The last CTA writes the exact-zero diagonal.
constexpr int SD_REDUCE_THREADS = 128;
__global__ void __launch_bounds__(SD_REDUCE_THREADS)
sqdist_reduce_kernel(const double* __restrict__ part, int ctas, int n, int accumulate, double* __restrict__ D2) {
    __shared__ double wsum[SD_REDUCE_THREADS / 32];
    const int nblk = sd_nblk(n), njobs = sd_njobs(nblk), cs = SD_WARPS / njobs;
    const int tid = threadIdx.x, lane = tid & 31;
    const int el = blockIdx.x, np = n * (n - 1) / 2;
    if (el >= np) {
        if (!accumulate)
            for (int i = tid; i < n; i += SD_REDUCE_THREADS) D2[size_t(i) * n + i] = 0.0;
        return;
    }
    int r = 0, rem = el;                                           // row-major pair index -> (r, c)
    while (rem >= n - 1 - r) {
        rem -= n - 1 - r;
        ++r;
    }
    const int c = r + 1 + rem;
    int job, e;
    sd_locate(nblk, n, r, c, job, e);
    const int terms = ctas * cs;
    double s = 0.0;
#pragma unroll 4
    for (int k = tid; k < terms; k += SD_REDUCE_THREADS) {
        const int cta = k / cs, slot = k % cs;
        s += part[(size_t(cta) * SD_WARPS + slot * njobs + job) * SD_ACC + e];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) wsum[tid >> 5] = s;
    __syncthreads();
    if (tid == 0) {
        double t = wsum[0];
#pragma unroll
        for (int w = 1; w < SD_REDUCE_THREADS / 32; ++w) t += wsum[w];
        if (accumulate) {
            D2[size_t(r) * n + c] += t;
            D2[size_t(c) * n + r] += t;
        } else {
            D2[size_t(r) * n + c] = t;
            D2[size_t(c) * n + r] = t;
        }
    }
}

// Host launch; called from flagg_sqdist (gram.cu).  Any alignment.
int sqdist_small_launch(flagg_view* v, int accumulate, double* D2, cudaStream_t st) {
    const int n = v->dev.n;
    const int64_t units = v->total_units[GRAN_128];
    FLAGG_REQUIRE(n >= 1 && n <= SD_MAX_N, "flagg_sqdist: serves cohorts of up to %d clients (n=%d)", SD_MAX_N, n);
    FLAGG_REQUIRE(units < (int64_t(1) << 31) - 65536, "flagg_sqdist: too many columns");
    if (units == 0 || n == 1) {
        if (!accumulate) FLAGG_CUDA(cudaMemsetAsync(D2, 0, sizeof(double) * size_t(n) * n, st));
        return FLAGG_OK;
    }
    const int nblk = sd_nblk(n), cs = SD_WARPS / sd_njobs(nblk);
    const int sms = device_sm_count();
    FLAGG_REQUIRE(sms > 0, "cannot query the SM count of the current device");
    // CTAs: a lane sees 4 columns of every unit of its slot; keep that under SD_MAX_ACC_COLS, and use whole waves
    // of two CTAs per SM so the pipeline fill is paid once per SM
    const int64_t max_units = int64_t(SD_MAX_ACC_COLS / 4) * cs;
    int64_t ctas = (units + max_units - 1) / max_units;
    ctas = std::max<int64_t>(ctas, std::min<int64_t>(units, 2 * sms));
    ctas = (ctas + 2 * sms - 1) / (2 * sms) * (2 * sms);
    ctas = std::min<int64_t>(ctas, units);
    const int64_t upc = (units + ctas - 1) / ctas;
    ctas = (units + upc - 1) / upc;
    FLAGG_REQUIRE(upc <= max_units, "flagg_sqdist: internal split error");
    const size_t need = sizeof(double) * size_t(ctas) * SD_WARPS * SD_ACC;
    int rc = v->ensure_scratch(need);
    if (rc != FLAGG_OK) return rc;
    double* part = static_cast<double*>(v->d_scratch);
    // ring (two CTAs per SM share 227 KB): 3 stages of cs = 8 / 4 units (32 KB each) up to n = 16; above, stages of
    // two units: 4 x 24 KB (n <= 24), 3 x 32 KB (n <= 32)
    const int ug = nblk <= 2 ? 1 : 2, stages = nblk == 3 ? 4 : 3;
    const size_t smem = sizeof(float) * stages * size_t(cs) * ug * nblk * SD_RB * 128;
    const bool vec = v->align_bytes >= 16;
    const unsigned g = unsigned(ctas), b = 32 * SD_WARPS;
    auto go = [&](auto kern) -> int {
        FLAGG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024));
        kern<<<g, b, smem, st>>>(v->dev, int(units), int(upc), part);
        return FLAGG_OK;
    };
    int lrc;
    if (nblk <= 2) lrc = vec ? go(sqdist_small_kernel<true, 3, 1>) : go(sqdist_small_kernel<false, 3, 1>);
    else if (nblk == 3) lrc = vec ? go(sqdist_small_kernel<true, 4, 2>) : go(sqdist_small_kernel<false, 4, 2>);
    else lrc = vec ? go(sqdist_small_kernel<true, 3, 2>) : go(sqdist_small_kernel<false, 3, 2>);
    if (lrc != FLAGG_OK) return lrc;
    FLAGG_CUDA(cudaGetLastError());
    sqdist_reduce_kernel<<<n * (n - 1) / 2 + 1, SD_REDUCE_THREADS, 0, st>>>(part, int(ctas), n, accumulate, D2);
    FLAGG_CUDA(cudaGetLastError());
    return FLAGG_OK;
}

}  // namespace flagg
