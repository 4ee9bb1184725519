// K4 (tensor-core variant): centred Gram matrix G = (X - c)(X - c)^T with exact-FP32
// products on the 5th-generation tensor cores (tcgen05, accumulators in TMEM).
//
// Replaces the n(n-1)/2 Python iterations of torch.norm(flat_i - flat_j).item() at
// krum.py:92-97 for n >= 64 (gram.cu keeps the CUDA-core path for small n).
//
// 3xTF32 split.  x~ = x - c is split on the CUDA cores into hi = x~ & 0xffffe000 (the 19
// bits a TF32 operand keeps) and lo = (x~ - hi) & 0xffffe000 (x~ - hi is exact in fp32), so
// x~ = hi + lo up to 2^-22 |x~|, and
//     x~_i x~_j  ~=  hi_i hi_j + hi_i lo_j + lo_i hi_j        (lo_i lo_j <= 2^-22 dropped)
// is three tcgen05.mma.kind::tf32 per k-step into ONE fp32 TMEM accumulator.  Both operand
// sets are written explicitly, so the result does not depend on how the tensor core rounds
// a full fp32 container to TF32.
//
// Shape of one CTA (416 threads, one CTA per SM, grid = upper-triangular tiles x K-splits):
//   warps 0-3   epilogue: every KC columns drain the TMEM accumulator into a second fp32
//               accumulator that also lives in TMEM (tcgen05.ld / add / tcgen05.st), every
//               FOLD drains fold that one into the CTA's FP64 partial tile in global memory
//   warps 4-11  producers: coalesced 128-bit global loads of 32 columns x 128 rows of the
//               row tile (and of the column tile when it differs), centre, split, store hi/lo
//               into 128B-swizzled K-major operand tiles in shared memory (3 stages)
//   warp 12     allocates TMEM; one lane issues the MMAs and commits them to mbarriers
// X is read once per tile from L2/HBM and never materialised as hi/lo in HBM.
//
// Precision: the tensor core adds each 8-deep dot product into the fp32 accumulator with one
// rounding; KC = 256 columns bound that to 96 roundings per drain, the second level adds
// drains in fp32 (round to nearest) for at most FOLD*KC = 8192 columns, and everything
// above that is FP64.  D^2 and sqrt are formed in FP64 by K5 (gram.cu).
#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "common.cuh"

namespace flagg {
namespace tc {

constexpr int TILE = 128;              // output tile edge = UMMA M = UMMA N
constexpr int BK = 32;                 // columns per k-block: 128 bytes = one swizzle row
constexpr int UMMA_K = 8;              // tf32
constexpr int STAGES = 3;
constexpr int KC_BLOCKS = 8;           // k-blocks per TMEM drain (KC = 256 columns)
constexpr int FOLD = 32;               // drains per FP64 fold (8192 columns)
constexpr int OPER_BYTES = TILE * BK * 4;            // 16 KB: one operand tile
constexpr int STAGE_BYTES = 4 * OPER_BYTES;          // A_hi, A_lo, B_hi, B_lo
constexpr int EPI_WARPS = 4, PROD_WARPS = 8;
constexpr int THREADS = (EPI_WARPS + PROD_WARPS + 1) * 32;   // 416
constexpr int MMA_WARP = EPI_WARPS + PROD_WARPS;             // 12
constexpr int PROD_ROWS = PROD_WARPS * 4;                    // rows covered by one pass of the producers (32)
constexpr int PROD_PASSES = TILE / PROD_ROWS;                // row slots per producer thread (4)
constexpr int TMEM_COLS = 512;         // D0 [0,128) D1 [128,256) ACC [256,384); power of two
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*alignment slack*/ + 256 /*barriers*/;

// UMMA instruction descriptor, kind::tf32: D fp32, A/B tf32, both K-major, M = N = 128
// (bit layout: cute/arch/mma_sm100_desc.hpp InstrDescriptor).
constexpr uint32_t IDESC = (1u << 4) | (2u << 7) | (2u << 10) | (uint32_t(TILE >> 3) << 17) | (uint32_t(TILE >> 4) << 24);

// ---- PTX wrappers -----------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return uint32_t(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "LAB_WAIT: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra LAB_DONE;\n\tbra LAB_WAIT;\n\tLAB_DONE:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
// Shared-memory matrix descriptor: K-major, SWIZZLE_128B, 8-row groups 1024 B apart
// (cute/arch/mma_sm100_desc.hpp SmemDescriptor; canonical layout ((8,m),(T,2)):((8T,SBO),(1,T))).
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr) {
    return uint64_t((saddr & 0x3ffffu) >> 4) | (uint64_t(1) << 16) | (uint64_t(1024 >> 4) << 32) | (uint64_t(1) << 46) |
           (uint64_t(2) << 61);
}
__device__ __forceinline__ void umma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(IDESC), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
constexpr int EG = 16;   // columns per epilogue TMEM access
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&r)[EG]) {
    uint32_t* u = reinterpret_cast<uint32_t*>(r);
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                 : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]), "=r"(u[4]), "=r"(u[5]), "=r"(u[6]), "=r"(u[7]), "=r"(u[8]), "=r"(u[9]), "=r"(u[10]), "=r"(u[11]), "=r"(u[12]), "=r"(u[13]), "=r"(u[14]), "=r"(u[15])
                 : "r"(taddr)
                 : "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const float (&r)[EG]) {
    const uint32_t* u = reinterpret_cast<const uint32_t*>(r);
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr), "r"(u[0]), "r"(u[1]), "r"(u[2]), "r"(u[3]), "r"(u[4]), "r"(u[5]), "r"(u[6]), "r"(u[7]), "r"(u[8]), "r"(u[9]), "r"(u[10]), "r"(u[11]), "r"(u[12]), "r"(u[13]), "r"(u[14]), "r"(u[15])
                 : "memory");
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// ---- k-block enumeration -----------------------------------------------------------------
// The CTA's K range is a run of 128-column units [u0, u1); a unit never crosses a tensor
// boundary and holds ceil(len/32) k-blocks.  Every role walks the same sequence.
struct KbIter {
    int64_t c0, seglen;        // first column of the unit inside its segment; length of segment p
    int u, u1, us_p, us_p1;    // current unit, end, cached unit range of segment p (host checks units < 2^31)
    int p, len, ks;            // segment, unit length (<= 128), k-block offset inside the unit
    __device__ __forceinline__ bool valid() const { return u < u1; }
    __device__ __forceinline__ void load_seg(const ViewDev& v) {
        us_p = int(v.unit_start[GRAN_128][p]);
        us_p1 = int(v.unit_start[GRAN_128][p + 1]);
        seglen = v.seg_start[p + 1] - v.seg_start[p];
    }
    __device__ __forceinline__ void load_unit(const ViewDev& v) {
        if (u >= u1) return;
        while (us_p1 <= u) { ++p; load_seg(v); }
        c0 = int64_t(u - us_p) * 128;
        len = int(min((long long)(128), (long long)(seglen - c0)));
        ks = 0;
    }
    __device__ __forceinline__ void init(const ViewDev& v, int64_t u0_, int64_t u1_) {
        u = int(u0_); u1 = int(u1_); c0 = 0; len = 0; ks = 0; us_p = us_p1 = 0; seglen = 0;
        p = (u < u1) ? find_segment(v.unit_start[GRAN_128], v.P, u0_) : 0;
        if (u < u1) load_seg(v);
        load_unit(v);
    }
    __device__ __forceinline__ void next(const ViewDev& v) {
        ks += BK;
        if (ks >= len) { ++u; load_unit(v); }
    }
};

__device__ int64_t count_kblocks(const ViewDev& v, int64_t u0, int64_t u1) {
    const int64_t* __restrict__ us = v.unit_start[GRAN_128];
    int64_t nkb = 0;
    for (int p = 0; p < v.P; ++p) {
        const int64_t a = max(u0, us[p]), b = min(u1, us[p + 1]);
        if (b <= a) continue;
        nkb += (b - a) * 4;
        if (b == us[p + 1]) {  // the segment's last (possibly short) unit is in range
            const int64_t seglen = v.seg_start[p + 1] - v.seg_start[p];
            const int lastlen = int(seglen - (us[p + 1] - us[p] - 1) * 128);
            nkb -= 4 - (lastlen + BK - 1) / BK;
        }
    }
    return nkb;
}

struct ProdRegs {
    float4 a[PROD_PASSES], b[PROD_PASSES], c;
};

// Per-segment state of a producer thread: row pointers (rows >= n are clamped to row n-1 and
// their stores predicated off, so loads stay branch-free) and the centre's segment base.
struct ProdRows {
    const float* a[PROD_PASSES];
    const float* b[PROD_PASSES];
    const float* cbase;   // centre + seg_start[p], or nullptr
    bool c_vec;           // cbase is 16-byte aligned
};

// Issue the global loads of one k-block for this producer thread; columns beyond the unit are zero.
__device__ __forceinline__ void prod_load(const KbIter& it, const ProdRows& rows, bool diag, int chunk, ProdRegs& r) {
    const int64_t col = it.c0 + it.ks + chunk * 4;      // first column of this thread's 16 bytes, inside the segment
    const int rem = it.len - it.ks - chunk * 4;         // valid columns from there (may be <= 0)
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    if (rem >= 4) {
#pragma unroll
        for (int j = 0; j < PROD_PASSES; ++j) r.a[j] = ldg_stream4(reinterpret_cast<const float4*>(rows.a[j] + col));
        if (!diag) {
#pragma unroll
            for (int j = 0; j < PROD_PASSES; ++j) r.b[j] = ldg_stream4(reinterpret_cast<const float4*>(rows.b[j] + col));
        }
        r.c = z;
        if (rows.cbase) {
            const float* cp = rows.cbase + col;
            if (rows.c_vec) r.c = __ldg(reinterpret_cast<const float4*>(cp));
            else r.c = make_float4(__ldg(cp), __ldg(cp + 1), __ldg(cp + 2), __ldg(cp + 3));
        }
    } else {
        auto part = [&](const float* row) {
            float4 x = z;
            if (row && rem > 0) {
                x.x = __ldg(row + col);
                if (rem > 1) x.y = __ldg(row + col + 1);
                if (rem > 2) x.z = __ldg(row + col + 2);
            }
            return x;
        };
#pragma unroll
        for (int j = 0; j < PROD_PASSES; ++j) r.a[j] = part(rows.a[j]);
        if (!diag) {
#pragma unroll
            for (int j = 0; j < PROD_PASSES; ++j) r.b[j] = part(rows.b[j]);
        }
        r.c = part(rows.cbase);
    }
}

// (x - c) -> hi = top 19 bits, lo = top 19 bits of the (exact) remainder;
// packed fp32x2 subtracts (FADD2), predicated 128-bit shared stores.
__device__ __forceinline__ void split_store(uint32_t hi_addr, uint32_t lo_addr, float4 x, float4 c, uint32_t ok) {
    uint64_t x01, x23, c01, c23, t01, t23, h01, h23, l01, l23;
    asm("mov.b64 %0, {%1, %2};" : "=l"(x01) : "f"(x.x), "f"(x.y));
    asm("mov.b64 %0, {%1, %2};" : "=l"(x23) : "f"(x.z), "f"(x.w));
    asm("mov.b64 %0, {%1, %2};" : "=l"(c01) : "f"(c.x), "f"(c.y));
    asm("mov.b64 %0, {%1, %2};" : "=l"(c23) : "f"(c.z), "f"(c.w));
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(t01) : "l"(x01), "l"(c01));
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(t23) : "l"(x23), "l"(c23));
    h01 = t01 & 0xffffe000ffffe000ull;
    h23 = t23 & 0xffffe000ffffe000ull;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(l01) : "l"(t01), "l"(h01));
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(l23) : "l"(t23), "l"(h23));
    l01 &= 0xffffe000ffffe000ull;
    l23 &= 0xffffe000ffffe000ull;
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %6, 0;\n\t"
        "@p st.shared.v2.b64 [%0], {%2, %3};\n\t"
        "@p st.shared.v2.b64 [%1], {%4, %5};\n\t}" ::"r"(hi_addr),
        "r"(lo_addr), "l"(h01), "l"(h23), "l"(l01), "l"(l23), "r"(ok)
        : "memory");
}

// Epilogue role (warps 0-3), shared by the SS and TS kernels: every KC columns drain the TMEM
// accumulator D[buf] into the second-level fp32 accumulator ACC (also TMEM), every FOLD drains
// fold ACC into the CTA's FP64 partial tile (column-major [128 col][128 row]) in global memory.
__device__ __forceinline__ void epilogue_role(uint32_t tmem_base, double* __restrict__ part, int ndrain, int warp, int lane,
                                              uint32_t tfull0, uint32_t tempty0) {
    const int row = warp * 32 + lane;
    const uint32_t lane_base = tmem_base + (uint32_t(warp * 32) << 16);
    double* my = part + (size_t(blockIdx.y) * gridDim.x + blockIdx.x) * (TILE * TILE) + row;
    bool folded = false;
    int in_fold = 0;
    for (int dr = 0; dr < ndrain; ++dr) {
        const int buf = dr & 1;
        mbar_wait(tfull0 + 8u * buf, uint32_t((dr >> 1) & 1));
        tc_fence_after();
#pragma unroll 1
        for (int g = 0; g < TILE / EG; ++g) {
            float x[EG];
            tmem_ld16(lane_base + uint32_t(buf * TILE + g * EG), x);
            if (in_fold > 0) {
                float a[EG];
                tmem_ld16(lane_base + uint32_t(2 * TILE + g * EG), a);
                tmem_wait_ld();
#pragma unroll
                for (int e = 0; e < EG; ++e) x[e] = __fadd_rn(x[e], a[e]);
            } else {
                tmem_wait_ld();
            }
            tmem_st16(lane_base + uint32_t(2 * TILE + g * EG), x);
        }
        tmem_wait_st();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(tempty0 + 8u * buf);
        ++in_fold;
        if (in_fold == FOLD || dr == ndrain - 1) {
#pragma unroll 1
            for (int g = 0; g < TILE / EG; ++g) {
                float a[EG];
                tmem_ld16(lane_base + uint32_t(2 * TILE + g * EG), a);
                tmem_wait_ld();
#pragma unroll
                for (int e = 0; e < EG; ++e) {
                    double* q = my + size_t(g * EG + e) * TILE;
                    *q = folded ? *q + double(a[e]) : double(a[e]);
                }
            }
            folded = true;
            in_fold = 0;
        }
    }
    if (ndrain == 0) {
        for (int c = 0; c < TILE; ++c) my[size_t(c) * TILE] = 0.0;
    }
}

// grid = (tiles, splits).  part: per-CTA FP64 partial tile, COLUMN-major [128 col][128 row].
__global__ void __launch_bounds__(THREADS, 1)
gram_tc_kernel(ViewDev v, const float* __restrict__ center, int nb, int64_t total_units, int64_t units_per_split,
               double* __restrict__ part) {
    extern __shared__ unsigned char smem_raw[];
    const uint32_t sbase = (smem_u32(smem_raw) + 1023u) & ~1023u;   // operand tiles need 1024 B alignment (SWIZZLE_128B)
    unsigned char* smem = smem_raw + (sbase - smem_u32(smem_raw));
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
    // bars[0..S) full, [S..2S) empty, [2S..2S+2) tmem_full, [2S+2..2S+4) tmem_empty
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);
    int* nkb_slot = reinterpret_cast<int*>(bars + 2 * STAGES + 6);
    const uint32_t bar0 = sbase + STAGES * STAGE_BYTES;
    auto FULL = [&](int s) { return bar0 + 8u * s; };
    auto EMPTY = [&](int s) { return bar0 + 8u * (STAGES + s); };
    auto TFULL = [&](int b) { return bar0 + 8u * (2 * STAGES + b); };
    auto TEMPTY = [&](int b) { return bar0 + 8u * (2 * STAGES + 2 + b); };

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    int t = blockIdx.x, bi = 0;
    while (t >= nb - bi) { t -= nb - bi; ++bi; }
    const int bj = bi + t;
    const bool diag = (bi == bj);
    const int64_t u0 = int64_t(blockIdx.y) * units_per_split;
    const int64_t u1 = min(total_units, u0 + units_per_split);

    if (tid == 0) {
        for (int s = 0; s < STAGES; ++s) { mbar_init(FULL(s), PROD_WARPS); mbar_init(EMPTY(s), 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(TFULL(b), 1); mbar_init(TEMPTY(b), EPI_WARPS); }
        fence_barrier_init();
        *nkb_slot = int(count_kblocks(v, u0, u1));
    }
    if (warp == MMA_WARP) tmem_alloc(smem_u32(tmem_slot), TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const int nkb = *nkb_slot;
    const int ndrain = (nkb + KC_BLOCKS - 1) / KC_BLOCKS;

    if (warp >= EPI_WARPS && warp < MMA_WARP) {
        // ================================ producers ================================
        const int pt = tid - EPI_WARPS * 32;
        const int rsub = (pt >> 5) * 4 + (lane >> 3), chunk = lane & 7;
        const uint32_t off0 = uint32_t(rsub >> 3) * 1024u + uint32_t(rsub & 7) * 128u + (uint32_t(chunk ^ (rsub & 7)) << 4);
        uint32_t ok = 0;   // bit j: A row slot j is a real client; bit 8+j: B row slot j
#pragma unroll
        for (int j = 0; j < PROD_PASSES; ++j) {
            ok |= uint32_t((bi * TILE + j * PROD_ROWS + rsub) < v.n) << j;
            ok |= uint32_t((bj * TILE + j * PROD_ROWS + rsub) < v.n) << (8 + j);
        }
        // zero this thread's slots of every stage once (rows >= n stay zero for the whole kernel)
        for (int s = 0; s < STAGES; ++s)
            for (int o = 0; o < 4; ++o)
#pragma unroll
                for (int j = 0; j < PROD_PASSES; ++j)
                    *reinterpret_cast<float4*>(smem + s * STAGE_BYTES + o * OPER_BYTES + j * (PROD_ROWS * 128) + off0) =
                        make_float4(0.f, 0.f, 0.f, 0.f);
        ProdRows rows;
        int cur_p = -1;
        auto refresh_rows = [&](int p) {
#pragma unroll
            for (int j = 0; j < PROD_PASSES; ++j) {
                const int ra = min(bi * TILE + j * PROD_ROWS + rsub, v.n - 1), rb = min(bj * TILE + j * PROD_ROWS + rsub, v.n - 1);
                rows.a[j] = v.seg_ptrs[size_t(ra) * v.P + p];
                rows.b[j] = v.seg_ptrs[size_t(rb) * v.P + p];
            }
            rows.cbase = center ? center + v.seg_start[p] : nullptr;
            rows.c_vec = (reinterpret_cast<uintptr_t>(rows.cbase) & 15) == 0;
            cur_p = p;
        };
        KbIter it;
        it.init(v, u0, u1);
        ProdRegs r0, r1;
        auto fetch = [&](ProdRegs& r) {
            if (it.valid()) {
                if (it.p != cur_p) refresh_rows(it.p);
                prod_load(it, rows, diag, chunk, r);
                it.next(v);
            }
        };
        fetch(r0);   // two k-blocks of loads in flight per thread
        fetch(r1);
        int stage = 0;
        uint32_t phase = 1;   // parity to wait for on EMPTY: a fresh barrier passes parity 1
        auto emit = [&](const ProdRegs& r) {
            mbar_wait(EMPTY(stage), phase);
            const uint32_t st = sbase + stage * STAGE_BYTES + off0;
#pragma unroll
            for (int j = 0; j < PROD_PASSES; ++j)
                split_store(st + j * (PROD_ROWS * 128), st + OPER_BYTES + j * (PROD_ROWS * 128), r.a[j], r.c, ok & (1u << j));
            if (!diag) {
#pragma unroll
                for (int j = 0; j < PROD_PASSES; ++j)
                    split_store(st + 2 * OPER_BYTES + j * (PROD_ROWS * 128), st + 3 * OPER_BYTES + j * (PROD_ROWS * 128), r.b[j],
                                r.c, ok & (256u << j));
            }
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) mbar_arrive(FULL(stage));
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
        };
        for (int kb = 0; kb < nkb; kb += 2) {
            emit(r0);
            fetch(r0);
            if (kb + 1 < nkb) {
                emit(r1);
                fetch(r1);
            }
        }
    } else if (warp == MMA_WARP) {
        // ================================ MMA issuer ================================
        int s = 0, inb = 0, buf = 0;
        uint32_t full_phase = 0, tempty_phase = 3u;   // bit b = parity to wait for on TEMPTY(b); fresh barriers pass parity 1
        for (int kb = 0; kb < nkb; ++kb) {
            if (inb == 0) {
                mbar_wait(TEMPTY(buf), (tempty_phase >> buf) & 1u);
                tempty_phase ^= 1u << buf;
                tc_fence_after();
            }
            mbar_wait(FULL(s), full_phase);
            tc_fence_after();
            const bool last_in_drain = (inb == KC_BLOCKS - 1) || (kb == nkb - 1);
            if (lane == 0) {
                const uint32_t a_hi = sbase + s * STAGE_BYTES, a_lo = a_hi + OPER_BYTES;
                const uint32_t b_hi = diag ? a_hi : a_hi + 2 * OPER_BYTES, b_lo = diag ? a_lo : a_hi + 3 * OPER_BYTES;
                const uint32_t d = tmem_base + uint32_t(buf * TILE);
#pragma unroll
                for (int k = 0; k < BK / UMMA_K; ++k) {
                    const uint32_t ko = k * UMMA_K * 4;
                    umma_tf32(d, make_desc(a_hi + ko), make_desc(b_hi + ko), (inb > 0 || k > 0) ? 1u : 0u);
                    umma_tf32(d, make_desc(a_hi + ko), make_desc(b_lo + ko), 1u);
                    umma_tf32(d, make_desc(a_lo + ko), make_desc(b_hi + ko), 1u);
                }
                umma_commit(EMPTY(s));
                if (last_in_drain) umma_commit(TFULL(buf));
            }
            __syncwarp();
            if (++s == STAGES) { s = 0; full_phase ^= 1; }
            if (last_in_drain) { inb = 0; buf ^= 1; } else { ++inb; }
        }
    } else {
        // ================================ epilogue ================================
        epilogue_role(tmem_base, part, ndrain, warp, lane, TFULL(0), TEMPTY(0));
    }
    tc_fence_before();
    __syncthreads();
    if (warp == MMA_WARP) {
        tc_fence_after();
        tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

#include "gram_ts.cuh"
#include "gram_ts2.cuh"

// Sum the split partials in split order and scatter the tile (and its mirror) into G.
__global__ void gram_tc_reduce_kernel(const double* __restrict__ part, int tiles, int splits, int nb, int n,
                                      int accumulate, double* __restrict__ G) {
    int t = blockIdx.x, bi = 0;
    while (t >= nb - bi) { t -= nb - bi; ++bi; }
    const int bj = bi + t;
    for (int e = blockIdx.y * blockDim.x + threadIdx.x; e < TILE * TILE; e += gridDim.y * blockDim.x) {
        const int lc = e / TILE, lr = e % TILE;   // column-major partial: consecutive threads -> consecutive rows
        const int r = bi * TILE + lr, c = bj * TILE + lc;
        if (r >= n || c >= n) continue;
        if (bi == bj && c < r) continue;
        double s = 0.0;
        for (int sp = 0; sp < splits; ++sp) s += part[(size_t(sp) * tiles + blockIdx.x) * (TILE * TILE) + e];
        if (accumulate) {
            G[size_t(r) * n + c] += s;
            if (r != c) G[size_t(c) * n + r] += s;
        } else {
            G[size_t(r) * n + c] = s;
            if (r != c) G[size_t(c) * n + r] = s;
        }
    }
}

}  // namespace tc

// Host launch; called from flagg_gram (gram.cu).  Requires 16-byte aligned segment bases.
// Kernel choice (FLAGG_GRAM_IMPL=ss|ts|ts1|ts2 overrides): n > 128 -> CTA-pair 256x256 tiles, n <= 128 ->
// the same body on single CTAs (gram_ts2.cuh); ss / ts are the earlier 128x128 kernels kept for comparison.
int gram_tc_launch(flagg_view* v, const float* center, int accumulate, double* G, cudaStream_t st) {
    using namespace tc;
    const int n = v->dev.n;
    const int64_t units = v->total_units[GRAN_128];
    FLAGG_REQUIRE(units < (int64_t(1) << 31), "flagg_gram: too many columns for the tensor-core path");
    const int sms = device_sm_count();
    FLAGG_REQUIRE(sms > 0, "cannot query the SM count of the current device");
    const char* impl = getenv("FLAGG_GRAM_IMPL");
    auto is = [&](const char* name) { return impl && !strcmp(impl, name); };
    const bool want_ss = is("ss"), want_ts2 = is("ts2"), want_ts1 = is("ts1");
    const bool forced = want_ss || want_ts2 || want_ts1;
    // measured: one (diagonal) pair tile beats three 128x128 tiles from n = 129 (tools/k4_small_n.py)
    const bool pair = want_ts2 || (!forced && n > TILE);
    const bool single = want_ts1 || (!forced && n <= TILE);
    static bool attr_set = false;
    if (!attr_set) {
        FLAGG_CUDA(cudaFuncSetAttribute(gram_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
        FLAGG_CUDA(cudaFuncSetAttribute(gram_ts2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, T2_SMEM_BYTES));
        FLAGG_CUDA(cudaFuncSetAttribute(gram_ts1_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, T2_SMEM_BYTES));
        attr_set = true;
    }
    if (pair || single) {
        // A-from-TMEM kernels (gram_ts2.cuh): CTA pairs on 256-row tiles, or single CTAs on 128-row tiles
        const int ncta = pair ? 2 : 1, rows = pair ? T2_ROWS : TILE;
        const int nb2 = (n + rows - 1) / rows;
        const int tiles2 = nb2 * (nb2 + 1) / 2;
        const int64_t ctas = int64_t(ncta) * tiles2;
        const int64_t per_cta_doubles = pair ? 2 * TILE * TILE : TILE * TILE;   // two N-halves per CTA
        // Split-K count: CTAs run in waves of `sms`; pick the count whose last wave is fullest
        // (n = 1000: 20 CTAs per split, 37 splits = 740 CTAs = exactly 5 waves of 148, where
        // 7 splits = 140 CTAs leave 8 SMs idle for the whole kernel).  Partials stay under 512 MB.
        int64_t splits = 1;
        {
            const int64_t per_split_bytes = int64_t(sizeof(double)) * ctas * per_cta_doubles;
            const int64_t smax = std::max<int64_t>(1, std::min<int64_t>({units, pair ? int64_t(96) : int64_t(2 * sms), (int64_t(512) << 20) / per_split_bytes}));
            const double per_split = pair ? 2e-4 : 2e-4 / 8;   // a little per-CTA overhead (prologue, FP64 partial tile)
            double best = 1e30;
            for (int64_t sp = 1; sp <= smax; ++sp) {
                const double cost = double((ctas * sp + sms - 1) / sms) / double(sp) + per_split * double(sp);   // waves per unit of K
                if (cost < best - 1e-12) { best = cost; splits = sp; }
            }
        }
        const int64_t upb = (units + splits - 1) / splits;
        splits = (units + upb - 1) / upb;
        const size_t need = sizeof(double) * size_t(splits) * size_t(ctas) * size_t(per_cta_doubles);
        int rc = v->ensure_scratch(need);
        if (rc != FLAGG_OK) return rc;
        double* part = static_cast<double*>(v->d_scratch);
        const dim3 grid{unsigned(ctas), unsigned(splits), 1u};
        // reduce: one thread per output element when there are few tiles (its loop over the splits is latency-bound)
        const unsigned ry = ctas >= 8 ? 16u : unsigned(per_cta_doubles / 256);
        if (pair) {
            gram_ts2_kernel<<<grid, THREADS, T2_SMEM_BYTES, st>>>(v->dev, center, nb2, units, upb, part);
            FLAGG_CUDA(cudaGetLastError());
            gram_t2_reduce_kernel<true><<<dim3(unsigned(ctas), ry), 256, 0, st>>>(part, int(splits), nb2, n, accumulate, G);
        } else {
            gram_ts1_kernel<<<grid, THREADS, T2_SMEM_BYTES, st>>>(v->dev, center, nb2, units, upb, part);
            FLAGG_CUDA(cudaGetLastError());
            gram_t2_reduce_kernel<false><<<dim3(unsigned(ctas), ry), 256, 0, st>>>(part, int(splits), nb2, n, accumulate, G);
        }
        FLAGG_CUDA(cudaGetLastError());
        return FLAGG_OK;
    }
    const int nb = (n + TILE - 1) / TILE;
    const int tiles = nb * (nb + 1) / 2;
    int64_t splits = std::max<int64_t>(1, sms / tiles);
    splits = std::min<int64_t>(splits, units);
    const int64_t upb = (units + splits - 1) / splits;
    splits = (units + upb - 1) / upb;
    const size_t need = sizeof(double) * size_t(splits) * tiles * TILE * TILE;
    int rc = v->ensure_scratch(need);
    if (rc != FLAGG_OK) return rc;
    double* part = static_cast<double*>(v->d_scratch);
    dim3 grid(tiles, unsigned(splits));
    gram_tc_kernel<<<grid, THREADS, SMEM_BYTES, st>>>(v->dev, center, nb, units, upb, part);
    FLAGG_CUDA(cudaGetLastError());
    gram_tc_reduce_kernel<<<dim3(tiles, 16), 256, 0, st>>>(part, tiles, int(splits), nb, n, accumulate, G);
    FLAGG_CUDA(cudaGetLastError());
    return FLAGG_OK;
}

}  // namespace flagg
