// K4 for small cohorts, "all pairs in one warp" form of flagg_sqdist: D2[i][j] = ||x_i - x_j||^2 by direct
// differences (krum.py:92-97), for 16-byte aligned rows.
//
// sqdist_small.cu cuts the rows into blocks of 8 and gives every warp a job of at most 64 pairs; from n = 17 the jobs
// are uneven (32 / 28 / 6 pairs at n = 20), every stage ends in a CTA barrier that waits for the heaviest job, and the
// kernel ran at 2.3 TB/s with the FP32 pipe a third busy (profiles/r02_ncu_sqdist_n20_summary.txt).  Here the
// n (n - 1) / 2 pairs, enumerated row-major, are cut into H equal groups; a warp owns ONE group (at most 124 fp32
// accumulators in registers, lane = column) and computes it for a whole 32-column quarter at a time from N
// shared-memory loads.  Every warp of the kernel executes the same instruction count per quarter, so the barrier that
// closes a stage finds the warps together, and there are N loads per 2 G arithmetic instructions instead of 12 per 64.
//     H = 1, 2 or 4 groups; CTA = 2 or 4 warps, one 128-column unit per stage, CTA-private ring filled by the TMA
//     engine (2-D box, tile::gather4 or per-row bulk copies, see the loader); registers capped at 168 -> 12 warps per
//     SM; units dealt round-robin over the grid (neighbouring columns are read at the same time); fp32 accumulation
//     over at most 1024 columns per lane, FP64 across lanes, per-warp partials, fixed-order reduce (deterministic).
// Arithmetic per pair and column is the same as in sqdist_small.cu: one fp32 subtraction, one fma.
#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "common.cuh"

namespace flagg {

constexpr int SP_MIN_N = 2, SP_MAX_N = 32;
constexpr int SP_MAX_ACC_COLS = 1024;     // fp32 accumulation length per lane
#ifndef FLAGG_SP_REG_CAP
#define FLAGG_SP_REG_CAP 168              // tuning builds: 128 (16 warps per SM, smaller groups)
#endif
constexpr int SP_REG_CAP = FLAGG_SP_REG_CAP;
constexpr int SP_MAX_STAGES = 8;
constexpr int SP_MAX_GROUPS = 4;

__host__ __device__ constexpr int sp_pairs(int n) { return n * (n - 1) / 2; }
// Groups: the accumulators of a group, the N row values and ~36 registers of addressing fit the register cap.  1, 2 or 4
// groups only: a CTA of 2 or 4 warps puts the same number of warps on each of the SM's four schedulers (warp id mod 4;
// five warps per CTA leave them 3 / 3 / 2 / 2) and the warps a scheduler holds all run the code of ONE group.  Builds
// with 3, 5 and 6 groups measured 0.9-2.3 ms at n = 32 against 0.66 for the row-block kernel -- measured with the old
// 2400-instruction epilogue per group (see the end of sp_run), so part of that was code size, not the group count.
// Registers per thread the kernel is built for: 168 (12 warps per SM) by default.  Fewer, larger groups won wherever
// they fit the register file -- 22-25 clients: 2 groups of up to 150 accumulators in 224 registers (8 warps per SM;
// n = 22: 0.267 -> 0.240 ms, n = 24: 0.320 -> 0.299 against 4 groups), 26-27: 2 groups of up to 176 in 255 (n = 26:
// 0.356 -> 0.336); 30-32 clients would need 8 groups at 168, so they get 4 groups of up to 124 accumulators in ~190
// registers.  One group of up to 190 for 17-20 clients was measured too: better at n = 17, worse at 18 and 20 -- left.
__host__ __device__ constexpr int sp_reg_cap(int n) {
    return n >= 30 ? 200 : (n >= 26 && n <= 27 ? 255 : (n >= 22 && n <= 25 ? 224 : SP_REG_CAP));
}
__host__ __device__ constexpr int sp_groups(int n) {
    const int gmax = sp_reg_cap(n) - n - 36;
    const int h = sp_pairs(n) <= gmax ? 1 : (sp_pairs(n) + gmax - 1) / gmax;
    return h <= 2 ? h : (h <= 4 ? 4 : 8);
}
__host__ __device__ constexpr bool sp_serves(int n) { return n >= SP_MIN_N && n <= SP_MAX_N && sp_groups(n) <= SP_MAX_GROUPS; }
__host__ __device__ constexpr int sp_group_size(int n) { return (sp_pairs(n) + sp_groups(n) - 1) / sp_groups(n); }
// Column slots per CTA.  Small CTAs win: the per-stage bookkeeping (mbarrier wait, CTA barrier, ring indices, ~95
// instructions) is paid per warp and stage, and the CTA barrier waits for the slowest of its warps.  One group: two warps
// of two quarters each (with four warps of one quarter the bookkeeping was as large as the arithmetic: n = 14 ran at
// 3.4 TB/s, 4.2 this way); two and four groups: one warp per group, four quarters each (n = 20: 0.270 -> 0.238 ms
// against two slots per group).
__host__ __device__ constexpr int sp_cslots(int h) { return h == 1 ? 2 : 1; }
__host__ __device__ constexpr int sp_pair_index(int n, int i, int j) { return i * n - i * (i + 1) / 2 + (j - i - 1); }   // i < j
__host__ __device__ constexpr int sp_min_blocks(int threads, int cap = SP_REG_CAP) { return 65536 / (threads * cap) > 0 ? 65536 / (threads * cap) : 1; }
__host__ __device__ constexpr int sp_threads(int n) { return 32 * sp_groups(n) * sp_cslots(sp_groups(n)); }
__host__ __device__ constexpr int sp_npad(int n) { return (n + 3) & ~3; }     // tile rows: gather4 brings four rows per request
// ring depth: the CTAs the register cap allows per SM share 216 KB of shared memory (1 KB per CTA is the system's)
__host__ __device__ constexpr int sp_stages(int n) {
    const int s = (216 * 1024 / sp_min_blocks(sp_threads(n), sp_reg_cap(n)) - 1024) / (sp_npad(n) * 512);
    return s > SP_MAX_STAGES ? SP_MAX_STAGES : (s < 2 ? 2 : s);
}

__device__ __forceinline__ void sp_mbar_init(uint32_t bar) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
}
__device__ __forceinline__ void sp_mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void sp_mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void sp_mbar_wait(uint32_t bar, uint32_t phase) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "SP_WAIT: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra SP_DONE;\n\tbra SP_WAIT;\n\tSP_DONE:\n\t}" ::"r"(bar), "r"(phase) : "memory");
}
// 1-D bulk copy (TMA engine) of `bytes` (a multiple of 16, both addresses 16-byte aligned) global -> shared, completion
// counted on the mbarrier
__device__ __forceinline__ void sp_bulk_copy(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src),
                 "r"(bytes), "r"(bar) : "memory");
}

// 2-D tensor-map box {128 columns, n rows} at column c0 of a dense tensor (ViewDev::tmaps; columns past the end are zeros)
__device__ __forceinline__ void sp_box_copy(uint32_t dst, const void* tmap, int c0, uint32_t bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(dst),
                 "l"(tmap), "r"(c0), "r"(0), "r"(bar) : "memory");
}
// tile::gather4 through the view-wide 512-byte-pitch map (ViewDev::gmap[1]): rows r.x .. r.w of that map = one 128-column
// unit of four clients
__device__ __forceinline__ void sp_gather4(uint32_t dst, const void* tmap, int4 r, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5, %6}], [%7];" ::"r"(dst),
        "l"(tmap), "r"(0), "r"(r.x), "r"(r.y), "r"(r.z), "r"(r.w), "r"(bar) : "memory");
}

// One quarter (32 columns, lane = column) of pair group HI: p = unit base + 32 * quarter + lane, rows 128 floats apart.
// Fully unrolled: the pair index of (i, j) is a constant, pairs of other groups and rows no pair of this group touches
// disappear at compile time.
template <int N, int HI>
__device__ __forceinline__ void sp_step(float (&acc)[sp_group_size(N)], const float* p) {
    constexpr int G = sp_group_size(N), LO = HI * G, END = (LO + G < sp_pairs(N)) ? LO + G : sp_pairs(N);
    float a[N];
#pragma unroll
    for (int i = 0; i < N; ++i) a[i] = p[i * 128];
#pragma unroll
    for (int i = 0; i < N - 1; ++i)
#pragma unroll
        for (int j = i + 1; j < N; ++j) {
            const int e = sp_pair_index(N, i, j);
            if (e >= LO && e < END) {
                const float dd = a[i] - a[j];
                acc[e - LO] = fmaf(dd, dd, acc[e - LO]);
            }
        }
}

template <int N, int HI>
__device__ __forceinline__ void sp_run(const ViewDev& v, int total_units, double* __restrict__ part, float* smem, uint32_t bars) {
    constexpr int H = sp_groups(N), CS = sp_cslots(H), WP = H * CS, G = sp_group_size(N), QW = 4 / CS, STAGES = sp_stages(N);
    constexpr int QU = QW <= 2 ? QW : 1;                           // four quarters: a loop, not four copies (instruction cache)
    constexpr int NPAD = sp_npad(N), UNIT = NPAD * 128;            // floats per stage: one 128-column unit of every row
    constexpr int ROWS = (N + WP - 1) / WP;                        // rows a warp copies
    static_assert(STAGES * UNIT >= WP * 32 * 33, "the ring must hold the epilogue's [32][33] tile of every warp");
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, cs = warp / H;
    // stage k of CTA b is unit k * grid + b
    const int u0 = blockIdx.x, ustep = gridDim.x;
    const int iters = u0 < total_units ? (total_units - u0 + ustep - 1) / ustep : 0;
    const int64_t* __restrict__ us = v.unit_start[GRAN_128];
    const uint32_t smem0 = uint32_t(__cvta_generic_to_shared(smem));

    // Loader.  Every copy is made by the TMA engine (no registers, completion counted on the stage's mbarrier):
    //   FEED_BOX     the rows are one dense matrix per tensor: ONE 2-D box {128 columns, n rows} per unit (thread 0);
    //   FEED_GATHER  separately allocated 512-byte aligned rows (the reference's list of lists): tile::gather4, four
    //                clients per request, lane j of warp w issues group w + j WP;
    //   FEED_ROWS    any other 16-byte aligned rows: a 512-byte bulk copy per row, issued by the lane that owns the row
    //                (lane j of warp w: row w + j WP) -- the copy instruction takes uniform operands, so this costs a
    //                ~10-instruction turn per row (a third of the stall samples at n = 20 when it was the only feed).
    // A tensor's ragged last unit (fewer than 128 columns) arrives zero-filled in a box; the other two feeds copy it
    // with plain loads and stores, zeros past the end.  The tensor of the next unit advances monotonically; each lane
    // keeps the base address of its row in that tensor.
    enum { FEED_BOX = 0, FEED_GATHER = 1, FEED_ROWS = 2 };
    const int feed = (v.tmaps != nullptr && v.dense_cols == 128) ? FEED_BOX : (v.gmap != nullptr ? FEED_GATHER : FEED_ROWS);
    const int myrow = warp + lane * WP;
    const float* rowptr = nullptr;
    const int mygroup = warp + lane * WP;                          // gather4: lane j of warp w owns clients 4 g .. 4 g + 3, g = w + j WP
    int4 grow = make_int4(0, 0, 0, 0);                             // their rows in the view-wide map, for the current tensor
    int lp = 0, lu0 = 0, lu1 = 0;
    int64_t llen = 0;
    auto enter_tensor = [&]() {
        lu0 = int(us[lp]);
        lu1 = int(us[lp + 1]);
        llen = v.seg_start[lp + 1] - v.seg_start[lp];
        if (myrow < N) rowptr = v.seg_ptrs[size_t(myrow) * v.P + lp];
        if (feed == FEED_GATHER && mygroup < NPAD / 4) grow = reinterpret_cast<const int4*>(v.seg_rows + size_t(lp) * NPAD)[mygroup];
    };
    if (iters > 0) {
        lp = find_segment(us, v.P, u0);
        enter_tensor();
    }
    int fill = 0;                                                  // ring slot the next stage goes to
    int c0 = 0, len = 0;                                           // of the stage prepared last (len = 0: none)
    // prepare(k) before the CTA barrier that frees the slot (thread 0 arms the slot's mbarrier), copy() after it
    auto prepare = [&](int k) {
        len = 0;
        if (k < iters) {
            const int u = u0 + k * ustep;
            while (u >= lu1) {                                     // warp-uniform
                ++lp;
                enter_tensor();
            }
            c0 = (u - lu0) * 128;                                  // < 2^31: checked on the host
            len = int(min((long long)128, (long long)(llen - c0)));
            if (threadIdx.x == 0) {
                if (feed == FEED_BOX) sp_mbar_expect_tx(bars + 8u * fill, uint32_t(N) * 512u);
                else if (len == 128) sp_mbar_expect_tx(bars + 8u * fill, uint32_t(feed == FEED_GATHER ? NPAD : N) * 512u);
                else sp_mbar_arrive(bars + 8u * fill);             // ragged unit: plain stores, ordered by the CTA barriers
            }
        }
    };
    auto copy = [&]() {
        const uint32_t slot = smem0 + uint32_t(fill * UNIT) * 4u, bar = bars + 8u * fill;
        if (len > 0) {
            if (feed == FEED_BOX) {
                if (threadIdx.x == 0) sp_box_copy(slot, static_cast<const char*>(v.tmaps) + size_t(lp) * 128, c0, bar);
            } else if (len == 128) {
                if (feed == FEED_GATHER) {
                    if (mygroup < NPAD / 4) {                      // row indices stay in registers: a load here sits between the barrier and the arithmetic
                        const int urow = c0 >> 7;
                        sp_gather4(slot + uint32_t(mygroup) * 2048u, static_cast<const char*>(v.gmap) + 128,
                                   make_int4(grow.x + urow, grow.y + urow, grow.z + urow, grow.w + urow), bar);
                    }
                } else if (myrow < N) {
                    sp_bulk_copy(slot + uint32_t(myrow) * 512u, rowptr + c0, 512u, bar);
                }
            } else {
#pragma unroll 1
                for (int j = 0; j < ROWS; ++j) {
                    const unsigned long long src = __shfl_sync(0xffffffffu, (unsigned long long)(rowptr), j);
                    const int r = warp + j * WP;
                    if (r < N) {
                        const float* sp = reinterpret_cast<const float*>(src) + c0;
                        float4 x;
                        x.x = 4 * lane + 0 < len ? __ldg(sp + 4 * lane + 0) : 0.f;
                        x.y = 4 * lane + 1 < len ? __ldg(sp + 4 * lane + 1) : 0.f;
                        x.z = 4 * lane + 2 < len ? __ldg(sp + 4 * lane + 2) : 0.f;
                        x.w = 4 * lane + 3 < len ? __ldg(sp + 4 * lane + 3) : 0.f;
                        reinterpret_cast<float4*>(smem + fill * UNIT + r * 128)[lane] = x;
                    }
                }
            }
        }
        fill = fill + 1 == STAGES ? 0 : fill + 1;
    };

    float acc[G];
#pragma unroll
    for (int e = 0; e < G; ++e) acc[e] = 0.f;

#pragma unroll 1
    for (int k = 0; k < STAGES - 1; ++k) {
        prepare(k);
        __syncthreads();
        copy();
    }
    int drain = 0;                                                 // ring slot of stage k
    uint32_t phase = 0;
#pragma unroll 1
    for (int k = 0; k < iters; ++k) {
        prepare(k + STAGES - 1);                                   // goes to the slot stage k - 1 used
        sp_mbar_wait(bars + 8u * drain, phase);                    // stage k has landed
        __syncthreads();                                           // every warp is done with stage k - 1 (and has its ragged stores out)
        copy();
        const float* ub = smem + drain * UNIT + lane;
        // four quarters per warp: a loop, not four copies -- the H groups run different code, and 5 groups x 4 x 6 KB of
        // straight-line arithmetic fell out of the instruction cache (2.3 ms at n = 32 in an early version)
#pragma unroll(QU)
        for (int q = 0; q < QW; ++q) sp_step<N, HI>(acc, ub + 32 * (cs * QW + q));
        if (++drain == STAGES) {
            drain = 0;
            phase ^= 1;
        }
    }
    // Lanes -> one FP64 value per accumulator, through shared memory (the ring is idle now): 32 accumulators at a time,
    // every lane writes its 32 values as a column of a [32][33] tile, lane l then adds row l (accumulator 32 b + l over
    // the 32 lanes, FP64, lane order) and the warp stores 32 sums as one 256-byte run of ITS G partials.  Small code on
    // purpose: a fully unrolled shuffle butterfly per accumulator is ~2400 instructions per group, which pushes the hot
    // loops of the H groups tens of KB apart -- their relative placement then decided the instruction-cache behaviour
    // (n = 28: 0.49 or 0.55 ms depending on unrelated changes to this epilogue).
    __syncthreads();                                               // every warp is done reading the ring
    float* tile = smem + warp * (32 * 33);
    double* dst = part + (size_t(blockIdx.x) * WP + warp) * G;
#pragma unroll
    for (int b = 0; b < (G + 31) / 32; ++b) {
#pragma unroll
        for (int i = 0; i < 32; ++i)
            if (32 * b + i < G) tile[i * 33 + lane] = acc[32 * b + i];
        __syncwarp();
        double s = 0.0;
#pragma unroll 1
        for (int l = 0; l < 32; ++l) s += double(tile[lane * 33 + l]);
        if (32 * b + lane < G) dst[32 * b + lane] = s;
        __syncwarp();
    }
}

// part[((cta * WP) + warp) * G + e], warp = cs * H + group
template <int N>
__global__ void __launch_bounds__(sp_threads(N), sp_min_blocks(sp_threads(N), sp_reg_cap(N)))
sqdist_pairs_kernel(ViewDev v, int total_units, double* __restrict__ part) {
    extern __shared__ __align__(128) float sp_smem[];
    __shared__ __align__(8) unsigned long long sp_bars[SP_MAX_STAGES];
    constexpr int H = sp_groups(N);
    const uint32_t bars = uint32_t(__cvta_generic_to_shared(sp_bars));
    if (threadIdx.x == 0) {
#pragma unroll
        for (int s = 0; s < sp_stages(N); ++s) sp_mbar_init(bars + 8u * s);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const int h = (threadIdx.x >> 5) % H;                          // warp-uniform
    if (h == 0) sp_run<N, 0>(v, total_units, part, sp_smem, bars);
    if constexpr (H > 1) { if (h == 1) sp_run<N, 1>(v, total_units, part, sp_smem, bars); }
    if constexpr (H > 2) { if (h == 2) sp_run<N, 2>(v, total_units, part, sp_smem, bars); }
    if constexpr (H > 3) { if (h == 3) sp_run<N, 3>(v, total_units, part, sp_smem, bars); }
    static_assert(H <= SP_MAX_GROUPS, "sqdist_pairs_kernel: more pair groups than the dispatch handles");
}

// One CTA per pair (r < c): 128 threads stride over the pair's (cta, column slot) partials, then a
// fixed-shape shuffle tree and a fixed-order sum over the four warps -- the same order every run.  (One WARP per pair
// walked 28-55 dependent L2 round trips: 15-25 us of a 100-240 us call.)  The last CTA writes the exact-zero diagonal.
constexpr int SP_REDUCE_THREADS = 128;
__global__ void __launch_bounds__(SP_REDUCE_THREADS)
sqdist_pairs_reduce_kernel(const double* __restrict__ part, int terms, int n, int H, int CS, int G, int accumulate,
                           double* __restrict__ D2) {
    __shared__ double wsum[SP_REDUCE_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, np = sp_pairs(n);
    const int e = blockIdx.x;
    if (e >= np) {
        if (!accumulate)
            for (int i = tid; i < n; i += SP_REDUCE_THREADS) D2[size_t(i) * n + i] = 0.0;
        return;
    }
    int r = 0, rem = e;                                            // row-major pair index -> (r, c)
    while (rem >= n - 1 - r) {
        rem -= n - 1 - r;
        ++r;
    }
    const int c = r + 1 + rem;
    double s = 0.0;
    const int h = e / G, eg = e - h * G, WP = H * CS;
#pragma unroll 4
    for (int k = tid; k < terms; k += SP_REDUCE_THREADS) s += part[(size_t(k / CS) * WP + (k % CS) * H + h) * G + eg];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) wsum[tid >> 5] = s;
    __syncthreads();
    if (tid == 0) {
        double t = wsum[0];
#pragma unroll
        for (int w = 1; w < SP_REDUCE_THREADS / 32; ++w) t += wsum[w];
        if (accumulate) {
            D2[size_t(r) * n + c] += t;
            D2[size_t(c) * n + r] += t;
        } else {
            D2[size_t(r) * n + c] = t;
            D2[size_t(c) * n + r] = t;
        }
    }
}

template <int N>
static int sp_go(flagg_view* v, int accumulate, double* D2, cudaStream_t st) {
  if constexpr (!sp_serves(N)) {
    set_error("flagg_sqdist: the pair-group kernel does not serve n = %d (more than %d pair groups)", N, SP_MAX_GROUPS);
    return FLAGG_EUNSUPPORTED;
  } else {
    constexpr int H = sp_groups(N), CS = sp_cslots(H), WP = H * CS, T = 32 * WP, G = sp_group_size(N), QW = 4 / CS;
    constexpr size_t smem = size_t(sp_stages(N)) * sp_npad(N) * 512;
    auto kern = sqdist_pairs_kernel<N>;
    const int64_t units = v->total_units[GRAN_128];
    const int sms = device_sm_count();
    FLAGG_REQUIRE(sms > 0, "cannot query the SM count of the current device");
    FLAGG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    static int resident = 0;                                       // CTAs per SM (registers and the ring), queried once
    if (resident == 0) {
        int r = 0;
        FLAGG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&r, kern, T, smem));
        FLAGG_REQUIRE(r >= 1, "flagg_sqdist: the n = %d kernel does not fit an SM", N);
        resident = r;
    }
    int per_sm = resident;
    const char* env = getenv("FLAGG_SQDIST_CTAS_PER_SM");          // tuning only
    if (env && atoi(env) >= 1) per_sm = std::min(per_sm, atoi(env));
    // a lane sees QW columns of every unit of its CTA: keep that under SP_MAX_ACC_COLS; whole waves
    const int64_t max_units = SP_MAX_ACC_COLS / QW;
    const int64_t wave = int64_t(per_sm) * sms;
    int64_t ctas = std::max<int64_t>((units + max_units - 1) / max_units, std::min<int64_t>(units, wave));
    ctas = (ctas + wave - 1) / wave * wave;
    ctas = std::min<int64_t>(ctas, units);
    const size_t need = sizeof(double) * size_t(ctas) * WP * G;
    int rc = v->ensure_scratch(need);
    if (rc != FLAGG_OK) return rc;
    double* part = static_cast<double*>(v->d_scratch);
    kern<<<unsigned(ctas), T, smem, st>>>(v->dev, int(units), part);
    FLAGG_CUDA(cudaGetLastError());
    sqdist_pairs_reduce_kernel<<<sp_pairs(N) + 1, SP_REDUCE_THREADS, 0, st>>>(part, int(ctas) * CS, N, H, CS, G, accumulate, D2);
    FLAGG_CUDA(cudaGetLastError());
    return FLAGG_OK;
  }
}

// Cohort sizes the pair-group kernel has an instantiation for (at most SP_MAX_GROUPS pair groups under the register cap).
bool sqdist_pairs_serves(int n) {
    if (n < SP_MIN_N || n > SP_MAX_N) return false;
    int last = SP_MIN_N;
    for (int k = SP_MIN_N; k <= SP_MAX_N && sp_groups(k) <= SP_MAX_GROUPS; ++k) last = k;    // sp_groups is monotone in n
    return n <= last;
}

// Host launch; called from flagg_sqdist (gram.cu) for 16-byte aligned rows.
int sqdist_pairs_launch(flagg_view* v, int accumulate, double* D2, cudaStream_t st) {
    const int n = v->dev.n;
    const int64_t units = v->total_units[GRAN_128];
    FLAGG_REQUIRE(n >= SP_MIN_N && n <= SP_MAX_N, "flagg_sqdist: the pair-group kernel serves %d..%d clients (n=%d)", SP_MIN_N, SP_MAX_N, n);
    FLAGG_REQUIRE(v->align_bytes >= 16, "flagg_sqdist: the pair-group kernel needs 16-byte aligned rows");
    FLAGG_REQUIRE(units >= 1 && units < (int64_t(1) << 24), "flagg_sqdist: column count out of range for the pair-group kernel");
    switch (n) {
#define SP_CASE(N) case N: return sp_go<N>(v, accumulate, D2, st);
        SP_CASE(2) SP_CASE(3) SP_CASE(4) SP_CASE(5) SP_CASE(6) SP_CASE(7) SP_CASE(8) SP_CASE(9) SP_CASE(10) SP_CASE(11)
        SP_CASE(12) SP_CASE(13) SP_CASE(14) SP_CASE(15) SP_CASE(16) SP_CASE(17) SP_CASE(18) SP_CASE(19) SP_CASE(20) SP_CASE(21)
        SP_CASE(22) SP_CASE(23) SP_CASE(24) SP_CASE(25) SP_CASE(26) SP_CASE(27) SP_CASE(28) SP_CASE(29) SP_CASE(30) SP_CASE(31)
        SP_CASE(32)
#undef SP_CASE
    }
    return FLAGG_EINVAL;
}

}  // namespace flagg
