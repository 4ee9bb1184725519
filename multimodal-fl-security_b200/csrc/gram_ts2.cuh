// K4, CTA-pair variant: 256 x 256 output tiles on two SMs (tcgen05 cta_group::2), A operand from
// tensor memory, B from shared memory.  Included by gram_tc.cu inside namespace flagg::tc after
// gram_ts.cuh (shares ts_load, the PTX wrappers, KbIter).  Replaces krum.py:92-97 for n > 128.
//
// Why pairs.  The 128 x 128 kernels are bound by the L2 -> SM operand feed: every CTA pulls 32 KB
// per 32-column k-block for 12 MMAs, and the pattern tops out at ~4.9 TB/s over the chip
// (profiles/r01_k4_load_ubench.txt: 23.3 ms for the C4 shard with no other work).  A CTA pair
// computing 256 x 256 pulls the same 32 KB per CTA per k-block for 24 MMAs: half the bytes per flop.
//
// Geometry of pair tile (BI, BJ), CTA rank r in {0, 1}:
//   A rows  BI*256 + 128 r + [0,128)  -> this CTA's TMEM lanes (UMMA M = 256 spans both CTAs)
//   B rows  BJ*256 + 128 r + [0,128)  -> this CTA's shared memory; local row b belongs to N-half
//           hf = b / 64 and is row b % 64 of that half's operand tile.  On diagonal tiles
//           (BI == BJ) the B rows ARE the A rows, so one set of loads feeds both.
//   Each k-step issues, per N-half, three MMAs (M = 256, N = 128: 64 rows of B from each CTA)
//   hi*hi, hi*lo, lo*hi into that half's accumulator region.  Output column c of half hf is
//   B row 128 (c / 64) + 64 hf + (c % 64) of the block (gram_ts2_reduce_kernel undoes this).
//
// TMEM (512 columns): three 128-column accumulator regions + 128 columns of A stages.  The two
// N-halves accumulate into two regions while the third is drained; their KC-column chunks are
// staggered by KC/2 so drain events alternate (event e: half (e+1)&1, region e%3; the chunk that
// starts after event e takes region (e+2)%3).  The second accumulation level (fp32, at most
// FOLD drains = 8192 columns) lives in SHARED memory here (128 KB), the third is FP64 in global.
//
// Synchronisation (a stage = 16 K positions for pairs, a 32-column k-block for single CTAs): FULL[st] (16 producer-warp arrivals, both CTAs) and TEMPTY[region] (8 epilogue-
// warp arrivals) live in the leader CTA (rank 0) whose MMA warp issues every MMA; EMPTY[st] and
// TFULL[region] are per CTA and signalled by multicast tcgen05.commit.
#pragma once
#include <type_traits>

constexpr int T2_ROWS = 256;                          // pair-tile edge
constexpr int T2_BTILE_BYTES = 64 * BK * 4;           // one B operand tile: 64 rows x 32 K = 8 KB
constexpr int T2_SLOT_BYTES = 4 * T2_BTILE_BYTES;     // [half][hi, lo]
constexpr int T2_KB_PAIR = 2, T2_KB_SINGLE = 4;       // k-blocks in flight (one 32 KB B slot + 64 TMEM columns of A each)
constexpr int T2_ACC_BYTES = 2 * TILE * TILE * 4;     // second-level accumulator [half][col][row] fp32
constexpr int T2_BAR_OFF = T2_KB_PAIR * T2_SLOT_BYTES + T2_ACC_BYTES;   // = 4 slots + half the accumulator bytes in single-CTA form
static_assert(T2_KB_SINGLE * T2_SLOT_BYTES + T2_ACC_BYTES / 2 == T2_BAR_OFF, "both forms share one shared-memory size");
constexpr int T2_SMEM_BYTES = T2_BAR_OFF + 1024 + 256;
constexpr int T2_CHUNK = KC_BLOCKS;                   // k-blocks per first-level chunk (KC = 256 columns)
constexpr uint32_t IDESC2 = (1u << 4) | (2u << 7) | (2u << 10) | (uint32_t(128 >> 3) << 17) | (uint32_t(256 >> 4) << 24);
// Register budget: 13 warps x 128 at launch; the epilogue warps hand back 88 each, the producers take 44.
constexpr int T2_EPI_REGS = 40, T2_PROD_REGS = 168;
static_assert((128 - T2_EPI_REGS) * EPI_WARPS >= (T2_PROD_REGS - 128) * PROD_WARPS, "setmaxnreg budget");
static_assert(T2_BTILE_BYTES == 8192, "the predicated B store hard-codes the hi -> lo tile distance");
static_assert(T2_SMEM_BYTES <= 232448, "shared memory budget");

__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `local_addr` in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t map_to_cta(uint32_t local_addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "LAB_WAITC: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra LAB_DONEC;\n\tbra LAB_WAITC;\n\tLAB_DONEC:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void tmem_alloc2(uint32_t dst_smem, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc2(uint32_t taddr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
__device__ __forceinline__ void umma2_tf32_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::tf32 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d_tmem),
        "r"(a_tmem), "l"(b_desc), "r"(IDESC2), "r"(accumulate)
        : "memory");
}
// same, the shared-memory descriptor passed as its two 32-bit words (the high word is a constant:
// SBO = 1024 B, version 1, SWIZZLE_128B; only the start-address field of the low word moves)
constexpr uint32_t DESC_HI = uint32_t((uint64_t(1024 >> 4) << 32 | uint64_t(1) << 46 | uint64_t(2) << 61) >> 32);
__device__ __forceinline__ void umma2_tf32_ts_lohi(uint32_t d_tmem, uint32_t a_tmem, uint32_t b_lo, uint32_t b_hi, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 bd;\n\tsetp.ne.b32 p, %5, 0;\n\tmov.b64 bd, {%2, %3};\n\t"
        "tcgen05.mma.cta_group::2.kind::tf32 [%0], [%1], bd, %4, p;\n\t}" ::"r"(d_tmem),
        "r"(a_tmem), "r"(b_lo), "r"(b_hi), "r"(IDESC2), "r"(accumulate)
        : "memory");
}
// arrive (count 1) on the barrier at this shared::cta offset in BOTH CTAs of the pair once all
// prior MMAs of this thread have completed
__device__ __forceinline__ void umma2_commit_both(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
                 "h"(uint16_t(3))
                 : "memory");
}

// one lane of a converged warp (elect.sync); the compiler keeps warp-uniform operands of the
// guarded tcgen05 instructions in uniform registers instead of re-broadcasting them per MMA
__device__ __forceinline__ bool elect_one() {
    uint32_t p;
    asm volatile("{\n\t.reg .pred q;\n\telect.sync _|q, 0xffffffff;\n\tselp.u32 %0, 1, 0, q;\n\t}" : "=r"(p));
    return p != 0;
}
// ---- The same body runs as a CTA pair (PAIR: cta_group::2, 256-row tiles, M = 256 x N = 128 per N-half,
// ---- 2 stages) or as a single CTA (cta_group::1, 128-row tiles, M = 128 x N = 64 per N-half: half the
// ---- accumulator columns, which pays for 4 stages -- a single CTA's k-block is ~1150 tensor cycles, so
// ---- the producer -> MMA -> producer round trip needs the deeper pipeline to hide).
__device__ __forceinline__ constexpr uint32_t idesc1(int ncols) {
    return (1u << 4) | (2u << 7) | (2u << 10) | (uint32_t(ncols >> 3) << 17) | (uint32_t(128 >> 4) << 24);
}
template <bool PAIR>
__device__ __forceinline__ void t2_mma(uint32_t d_tmem, uint32_t a_tmem, uint32_t b_lo, uint32_t b_hi, uint32_t accumulate, uint32_t idesc) {
    if constexpr (PAIR) {
        umma2_tf32_ts_lohi(d_tmem, a_tmem, b_lo, b_hi, accumulate);
    } else {
        asm volatile(
            "{\n\t.reg .pred p;\n\t.reg .b64 bd;\n\tsetp.ne.b32 p, %5, 0;\n\tmov.b64 bd, {%2, %3};\n\t"
            "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], bd, %4, p;\n\t}" ::"r"(d_tmem),
            "r"(a_tmem), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate)
            : "memory");
    }
}
template <bool PAIR>
__device__ __forceinline__ void t2_commit(uint32_t bar) {
    if constexpr (PAIR) umma2_commit_both(bar); else umma_commit(bar);
}
__device__ __forceinline__ uint64_t pack2(float a, float b) {
    uint64_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
    return r;
}
__device__ __forceinline__ uint64_t half2x(uint64_t a) {   // (0.5 a.x, 0.5 a.y): exact
    uint64_t r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(0x3f0000003f000000ull));
    return r;
}
__device__ __forceinline__ uint64_t sub2(uint64_t a, uint64_t b) {
    uint64_t r;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}


// One k-block of a producer thread as fp32 PAIRS (value pair g = columns 8q + 2g, 8q + 2g + 1):
// 256-bit / 128-bit loads straight into 64-bit registers, so the packed subtracts need no moves.
struct T2Regs {
    uint64_t a[2][4], b[2][4], c[4];
};
__device__ __forceinline__ void ldg_pairs8(const float* p, uint64_t (&v)[4]) {
    asm volatile("ld.global.nc.L1::no_allocate.v4.b64 {%0,%1,%2,%3}, [%4];" : "=l"(v[0]), "=l"(v[1]), "=l"(v[2]), "=l"(v[3]) : "l"(p));
}
__device__ __forceinline__ void ldg_pairs44(const float* p, uint64_t (&v)[4]) {
    asm volatile("ld.global.nc.L1::no_allocate.v2.b64 {%0,%1}, [%2];" : "=l"(v[0]), "=l"(v[1]) : "l"(p));
    asm volatile("ld.global.nc.L1::no_allocate.v2.b64 {%0,%1}, [%2];" : "=l"(v[2]), "=l"(v[3]) : "l"(p + 4));
}
// (An L2 prefetch of the lines 6-12 k-blocks ahead was measured and rejected: it doubles the L2 request
// count, and both forms lost 10-25 %.)
__device__ __forceinline__ void t2_load(const KbIter& it, const TsRows& rows, bool diag, int q, T2Regs& r) {
    const int64_t col = it.c0 + it.ks + q * 8;
    const int rem = it.len - it.ks - q * 8;
    if (rem >= 8) {
        if (rows.v8) {
#pragma unroll
            for (int h = 0; h < 2; ++h) ldg_pairs8(rows.a[h] + col, r.a[h]);
            if (!diag) {
#pragma unroll
                for (int h = 0; h < 2; ++h) ldg_pairs8(rows.b[h] + col, r.b[h]);
            }
        } else {
#pragma unroll
            for (int h = 0; h < 2; ++h) ldg_pairs44(rows.a[h] + col, r.a[h]);
            if (!diag) {
#pragma unroll
                for (int h = 0; h < 2; ++h) ldg_pairs44(rows.b[h] + col, r.b[h]);
            }
        }
        if (rows.cbase) {
            const float* cp = rows.cbase + col;
            if ((reinterpret_cast<uintptr_t>(cp) & 15) == 0) {
                asm volatile("ld.global.nc.v2.b64 {%0,%1}, [%2];" : "=l"(r.c[0]), "=l"(r.c[1]) : "l"(cp));
                asm volatile("ld.global.nc.v2.b64 {%0,%1}, [%2];" : "=l"(r.c[2]), "=l"(r.c[3]) : "l"(cp + 4));
            } else {
#pragma unroll
                for (int g = 0; g < 4; ++g) r.c[g] = pack2(__ldg(cp + 2 * g), __ldg(cp + 2 * g + 1));
            }
        } else {
#pragma unroll
            for (int g = 0; g < 4; ++g) r.c[g] = 0ull;
        }
    } else {
        auto part = [&](const float* row, uint64_t (&v)[4]) {
#pragma unroll
            for (int g = 0; g < 4; ++g)
                v[g] = pack2((row && 2 * g < rem) ? __ldg(row + col + 2 * g) : 0.f, (row && 2 * g + 1 < rem) ? __ldg(row + col + 2 * g + 1) : 0.f);
        };
#pragma unroll
        for (int h = 0; h < 2; ++h) part(rows.a[h], r.a[h]);
        if (!diag) {
#pragma unroll
            for (int h = 0; h < 2; ++h) part(rows.b[h], r.b[h]);
        }
        part(rows.cbase, r.c);
    }
}

template <bool PAIR>
__device__ __forceinline__ void gram_t2_body(const ViewDev& v, const float* __restrict__ center, int nb2, int64_t total_units,
                                             int64_t units_per_split, double* __restrict__ part) {
    constexpr int NCTA = PAIR ? 2 : 1;            // CTAs per tile
    constexpr int ROWS = PAIR ? T2_ROWS : TILE;   // tile edge
    constexpr int NC = PAIR ? TILE : TILE / 2;    // accumulator columns per N-half (= TMEM region width)
    // Pipeline stage: 16 K positions (half a k-block) in the pair form -- its two k-blocks in flight are then four
    // stages and the MMA warp starts on a half-filled k-block -- and a whole k-block in the single-CTA form.
    constexpr bool HALF = PAIR;
    constexpr int NKB = PAIR ? T2_KB_PAIR : T2_KB_SINGLE;
    constexpr int NSTG = HALF ? 2 * NKB : NKB;
    static_assert(NSTG <= 4, "four FULL and four EMPTY barriers");
    constexpr int TMEM_A0 = 3 * NC;               // first TMEM column of the A stages
    static_assert(TMEM_A0 + NKB * 64 <= TMEM_COLS, "TMEM budget");
    constexpr uint32_t HF_OFF = 2u * T2_BTILE_BYTES;                 // B rows 64-127 of this CTA (slot layout [half][hi, lo])
    constexpr uint32_t LO_OFF = uint32_t(T2_BTILE_BYTES);            // hi -> lo operand
    extern __shared__ unsigned char smem_raw[];
    const uint32_t sbase = (smem_u32(smem_raw) + 1023u) & ~1023u;
    unsigned char* smem = smem_raw + (sbase - smem_u32(smem_raw));
    float* acc = reinterpret_cast<float*>(smem + NKB * T2_SLOT_BYTES);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + T2_BAR_OFF);
    // bars: [0,4) FULL (leader's used), [4,8) EMPTY, [8,11) TFULL, [11,14) TEMPTY (leader's used)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 14);
    int* nkb_slot = reinterpret_cast<int*>(bars + 15);
    const uint32_t bar0 = sbase + T2_BAR_OFF;
    auto FULL = [&](int s) { return bar0 + 8u * s; };
    auto EMPTY = [&](int s) { return bar0 + 8u * (4 + s); };
    auto TFULL = [&](int r) { return bar0 + 8u * (8 + r); };
    auto TEMPTY = [&](int r) { return bar0 + 8u * (11 + r); };

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t rank = PAIR ? cluster_ctarank() : 0u;
    int t = int(PAIR ? blockIdx.x >> 1 : blockIdx.x), BI = 0;
    while (t >= nb2 - BI) { t -= nb2 - BI; ++BI; }
    const int BJ = BI + t;
    const bool diag = (BI == BJ);
    const int64_t u0 = int64_t(blockIdx.y) * units_per_split;
    const int64_t u1 = min(total_units, u0 + units_per_split);

    if (tid == 0) {
        for (int s = 0; s < NSTG; ++s) { mbar_init(FULL(s), NCTA * PROD_WARPS); mbar_init(EMPTY(s), 1); }
        for (int r = 0; r < 3; ++r) { mbar_init(TFULL(r), 1); mbar_init(TEMPTY(r), NCTA * EPI_WARPS); }
        fence_barrier_init();
        *nkb_slot = int(count_kblocks(v, u0, u1));
    }
    if (warp == MMA_WARP) {
        if constexpr (PAIR) tmem_alloc2(smem_u32(tmem_slot), TMEM_COLS); else tmem_alloc(smem_u32(tmem_slot), TMEM_COLS);
    }
    tc_fence_before();
    __syncthreads();
    if constexpr (PAIR) cluster_sync_all();   // both CTAs' barriers are initialised before any remote arrive
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const int nkb = *nkb_slot;
    // drain events: one every CHUNK/2 k-blocks strictly inside the range, two at the end
    const int n_ev = nkb > 0 ? (nkb - 1) / (T2_CHUNK / 2) + 2 : 0;

    if (warp >= EPI_WARPS && warp < MMA_WARP) {
        // ================================ producers ================================
        asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(T2_PROD_REGS));
        const int qd = warp & 3, h16 = (warp - EPI_WARPS) >> 2;
        const int lane0 = 32 * qd + 16 * h16;           // first TMEM lane = first local row of this warp
        const int r8 = lane >> 2, q = lane & 3;
        const int rloc = lane0 + r8;                    // local rows rloc, rloc + 8
        const int hf = qd >> 1, brow = rloc & 63;       // N-half and row inside that half's B tile
        // B byte offset of K position 2q of row brow (see gram_ts.cuh); odd rows of an 8-row group
        // take the OTHER k-step of the stage (^ 64 bytes) so a half-warp's 8-byte stores hit 16
        // distinct bank pairs
        const uint32_t boff = uint32_t(hf) * HF_OFF + uint32_t(brow >> 3) * 1024u + uint32_t(brow & 7) * 128u +
                              ((uint32_t((q >> 1) ^ (brow & 7)) << 4) + uint32_t(q & 1) * 8u);
        const bool odd = (r8 & 1) != 0;
        bool okA[2], okB[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            okA[h] = (BI * ROWS + int(rank) * TILE + rloc + 8 * h) < v.n;
            okB[h] = (BJ * ROWS + int(rank) * TILE + rloc + 8 * h) < v.n;
        }
        // padding rows of B stay zero for the whole kernel (their stores are predicated off)
        for (int s = 0; s < NKB; ++s)
            for (int o = 0; o < 2; ++o)
#pragma unroll
                for (int h = 0; h < 2; ++h)
#pragma unroll
                    for (int g = 0; g < 4; ++g)
                        *reinterpret_cast<uint2*>(smem + s * T2_SLOT_BYTES + o * LO_OFF + ((boff + h * 1024u) ^ (uint32_t(g) << 5))) =
                            make_uint2(0u, 0u);
        TsRows rows;
        int cur_p = -1;
        auto refresh_rows = [&](int p) {
            uintptr_t bits = 0;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int ra = min(BI * ROWS + int(rank) * TILE + rloc + 8 * h, v.n - 1);
                const int rb = min(BJ * ROWS + int(rank) * TILE + rloc + 8 * h, v.n - 1);
                rows.a[h] = v.seg_ptrs[size_t(ra) * v.P + p];
                rows.b[h] = v.seg_ptrs[size_t(rb) * v.P + p];
                bits |= reinterpret_cast<uintptr_t>(rows.a[h]) | reinterpret_cast<uintptr_t>(rows.b[h]);
            }
            rows.cbase = center ? center + v.seg_start[p] : nullptr;
            rows.v8 = (bits & 31) == 0;
            cur_p = p;
        };
        KbIter it;
        it.init(v, u0, u1);
        T2Regs r0 = {}, r1 = {}, r2 = {};   // three k-blocks of loads in flight per thread
        auto fetch = [&](T2Regs& r) {
            if (it.valid()) {
                if (it.p != cur_p) refresh_rows(it.p);
#ifndef T2_DBG_NOLOAD
                t2_load(it, rows, diag, q, r);
#endif
                it.next(v);
            }
        };
        fetch(r0);
        fetch(r1);
        fetch(r2);
        int stage = 0;                                  // stage being filled
        uint32_t phase = 1;
        const uint32_t a_lane = tmem_base + (uint32_t(lane0) << 16) + TMEM_A0;
        const uint32_t full_leader = PAIR ? map_to_cta(FULL(0), 0) : FULL(0);
        constexpr uint64_t HI2 = 0xffffe000ffffe000ull;
        const bool pad_any = !(okA[0] && okA[1] && okB[0] && okB[1]);
        // Single-CTA form: a k-block is published (FULL arrive) at the start of the NEXT emit, after the
        // loads of the following k-block have been issued, so the TMEM-store latency overlaps them
        // (n = 128 x 6.25M: 1.35 -> 1.04 ms).
        int pend = -1;
        auto publish = [&]() {
            if (pend < 0) return;
            tmem_wait_st();
            tc_fence_before();
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(full_leader + 8u * uint32_t(pend));
            pend = -1;
        };
        auto emit = [&](T2Regs& r, auto diag_c) {
            constexpr bool DIAG = decltype(diag_c)::value;
            publish();
            if (pad_any) {   // padding rows: x := c, so the centred value is exactly zero (A is stored unconditionally)
#pragma unroll
                for (int h = 0; h < 2; ++h)
                    if (!okA[h]) {
#pragma unroll
                        for (int g = 0; g < 4; ++g) r.a[h][g] = r.c[g];
                    }
            }
#pragma unroll
            for (int sub = 0; sub < 2; ++sub) {
                // this stage's two k-steps are value pairs g = sub and g = sub + 2 of the thread's 8 columns
                uint64_t ah[2][2], al[2][2], bh[2][2], bl[2][2];   // [row h][k-step j]
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    const int g = sub + 2 * j;
                    const uint64_t c2 = r.c[g];
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const uint64_t ta = sub2(r.a[h][g], c2);
                        ah[h][j] = ta & HI2;
                        al[h][j] = sub2(ta, ah[h][j]);
                        if constexpr (DIAG) {
                            // diagonal tile: D' = hi (hi/2 + lo)^T, G = D' + D'^T (two products instead of three,
                            // no lo operand in TMEM); the reduce kernel adds the transpose
                            bh[h][j] = half2x(ah[h][j]);
                            bl[h][j] = al[h][j];
                        } else {
                            const uint64_t tb = sub2(r.b[h][g], c2);
                            bh[h][j] = tb & HI2;
                            bl[h][j] = sub2(tb, bh[h][j]);
                        }
                    }
                }
                if (HALF || sub == 0) {
                    mbar_wait(EMPTY(stage), phase);
                    tc_fence_after();
                }
                // A -> TMEM: fragment register 4 j + 2 h + e = (row h, k-step j, element e)
                uint32_t fh[8], fl[8];
#pragma unroll
                for (int j = 0; j < 2; ++j)
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        fh[4 * j + 2 * h] = uint32_t(ah[h][j]);
                        fh[4 * j + 2 * h + 1] = uint32_t(ah[h][j] >> 32);
                        fl[4 * j + 2 * h] = uint32_t(al[h][j]);
                        fl[4 * j + 2 * h + 1] = uint32_t(al[h][j] >> 32);
                    }
#ifndef T2_DBG_NOSTTM
                tmem_st_16x256b_x2(a_lane + uint32_t(HALF ? stage * 32 : stage * 64 + sub * 32), fh);
                if constexpr (!DIAG) tmem_st_16x256b_x2(a_lane + uint32_t(HALF ? stage * 32 : stage * 64 + sub * 32) + 16u, fl);
#else
                asm volatile("" ::"r"(fh[0] ^ fh[1] ^ fh[2] ^ fh[3] ^ fh[4] ^ fh[5] ^ fh[6] ^ fh[7] ^ fl[0] ^ fl[1] ^ fl[2] ^ fl[3] ^ fl[4] ^ fl[5] ^ fl[6] ^ fl[7]));
#endif
                // B -> smem: store i writes k-step j = i for even rows, j = 1 - i for odd rows
                const uint32_t sb = sbase + uint32_t(HALF ? stage >> 1 : stage) * T2_SLOT_BYTES;
#pragma unroll
                for (int h = 0; h < 2; ++h) {
#pragma unroll
                    for (int i = 0; i < 2; ++i) {
                        const uint64_t vh = odd ? bh[h][1 - i] : bh[h][i], vl = odd ? bl[h][1 - i] : bl[h][i];
                        const int ge = sub + 2 * i, go = sub + 2 * (1 - i);
                        const uint32_t addr = sb + ((boff + h * 1024u) ^ (odd ? uint32_t(go) << 5 : uint32_t(ge) << 5));
                        asm volatile(
                            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %3, 0;\n\t"
                            "@p st.shared.b64 [%0], %1;\n\t"
                            "@p st.shared.b64 [%0+%4], %2;\n\t}" ::"r"(addr), "l"(vh), "l"(vl), "r"(uint32_t(okB[h])), "n"(LO_OFF)
                            : "memory");
                    }
                }
                if (HALF || sub == 1) {   // the stage is written
                    pend = stage;
                    if (++stage == NSTG) { stage = 0; phase ^= 1; }
                    if constexpr (PAIR) publish();   // two k-blocks in flight only: a deferred arrive would idle the MMA warp
                }
            }
        };
        // (Five register sets on single-CTA diagonal tiles -- the B registers are dead there -- were measured
        // and rejected: 0.96 / 0.97 / 1.06 ms at n = 64 / 100 / 128 against 0.90 / 0.93 / 1.03 with three.)
        auto run = [&](auto diag_c) {
            for (int kb = 0; kb < nkb; kb += 3) {
                emit(r0, diag_c);
                fetch(r0);
                if (kb + 1 < nkb) {
                    emit(r1, diag_c);
                    fetch(r1);
                }
                if (kb + 2 < nkb) {
                    emit(r2, diag_c);
                    fetch(r2);
                }
            }
        };
        if (diag) run(std::true_type{}); else run(std::false_type{});
        publish();
    } else if (warp == MMA_WARP) {
        // ================================ MMA issuer (leader CTA only) ================================
        if (rank == 0) {
            int s = 0;
            uint32_t full_phase = 0;
            const uint32_t blo0 = uint32_t(make_desc(sbase));      // low descriptor word of slot 0, half 0, hi tile, k-step 0
            const uint32_t a_tm0 = tmem_base + TMEM_A0;
            uint32_t tempty_uses = 0;          // bit r: parity of the next wait on TEMPTY[r]
            int reg[2] = {1, 0};               // accumulator region of half 0 / half 1
            bool fresh[2] = {true, true};      // next MMA of the half starts a chunk (overwrite)
            int ev = 0;                        // drain events committed so far
            // single CTA, n - 128 BJ <= 64: the second N-half holds only padding rows, so its MMAs are skipped
            // (its drain events still run; the reduce kernel never reads those columns)
            const int nh = (!PAIR && v.n - BJ * ROWS <= TILE / 2) ? 1 : 2;
            constexpr uint32_t idesc = idesc1(64);
            for (int kb = 0; kb < nkb; ++kb) {
                const uint32_t d0 = tmem_base + uint32_t(reg[0] * NC), d1 = tmem_base + uint32_t(reg[1] * NC);
                const uint32_t acc0 = fresh[0] ? 0u : 1u, acc1 = fresh[1] ? 0u : 1u;
                // the 3 x 2 MMAs of sub-stage `sub`, k-step j (K position 32 (sub + 2 j) bytes into the 128-byte row;
                // descriptor start address (>> 4) advances by 2 per 32-byte k-step)
                auto issue = [&](int sub, int j, uint32_t lo0, uint32_t a_hi) {
                    const uint32_t kofs = uint32_t(2 * sub + 4 * j);
#pragma unroll
                    for (int hf = 0; hf < 2; ++hf) {
                        if (hf >= nh) continue;
#ifdef T2_DBG_NOMMA
                        continue;
#endif
                        const uint32_t d = hf ? d1 : d0;
                        const uint32_t b_hi = lo0 + kofs + uint32_t(hf) * (HF_OFF >> 4), b_lo = b_hi + (LO_OFF >> 4);
                        t2_mma<PAIR>(d, a_hi + 8 * j, b_hi, DESC_HI, (sub == 0 && j == 0) ? (hf ? acc1 : acc0) : 1u, idesc);
                        t2_mma<PAIR>(d, a_hi + 8 * j, b_lo, DESC_HI, 1u, idesc);
                        if (!diag) t2_mma<PAIR>(d, a_hi + 16 + 8 * j, b_hi, DESC_HI, 1u, idesc);
                    }
                };
                if constexpr (HALF) {
#pragma unroll
                    for (int sub = 0; sub < 2; ++sub) {
                        mbar_wait_cluster(FULL(s), full_phase);
                        tc_fence_after();
                        const uint32_t lo0 = blo0 + uint32_t(s >> 1) * (T2_SLOT_BYTES >> 4);
                        const uint32_t a_hi = a_tm0 + uint32_t(s * 32);
                        const uint32_t empty_s = EMPTY(s);
                        if (elect_one()) {
#pragma unroll
                            for (int j = 0; j < 2; ++j) issue(sub, j, lo0, a_hi);
                            t2_commit<PAIR>(empty_s);
                        }
                        __syncwarp();
                        if (++s == NSTG) { s = 0; full_phase ^= 1; }
                    }
                } else {
                    mbar_wait_cluster(FULL(s), full_phase);
                    tc_fence_after();
                    const uint32_t lo0 = blo0 + uint32_t(s) * (T2_SLOT_BYTES >> 4);
                    const uint32_t a_hi0 = a_tm0 + uint32_t(s * 64);
                    const uint32_t empty_s = EMPTY(s);
                    if (elect_one()) {
#pragma unroll
                        for (int sub = 0; sub < 2; ++sub)
#pragma unroll
                            for (int j = 0; j < 2; ++j) issue(sub, j, lo0, a_hi0 + uint32_t(32 * sub));
                        t2_commit<PAIR>(empty_s);
                    }
                    __syncwarp();
                    if (++s == NSTG) { s = 0; full_phase ^= 1; }
                }
                fresh[0] = fresh[1] = false;
                // chunk boundaries after k-block kb: half 1 at 4, 12, 20, ...; half 0 at 8, 16, ...; both at the end
                const int done = kb + 1;
                int nclose = 0;
                if (done == nkb) nclose = 2;
                else if (done % (T2_CHUNK / 2) == 0) nclose = 1;
                for (int c = 0; c < nclose; ++c) {
                    const int hf = (ev + 1) & 1, r = ev % 3;
                    if (elect_one()) t2_commit<PAIR>(TFULL(r));
                    __syncwarp();
                    if (done < nkb) {
                        // the half continues in region (ev + 2) % 3, which event ev - 1 drained
                        const int nr = (ev + 2) % 3;
                        if (ev >= 1) {
                            mbar_wait_cluster(TEMPTY(nr), (tempty_uses >> nr) & 1u);
                            tempty_uses ^= 1u << nr;
                            tc_fence_after();
                        }
                        if (hf == 0) { reg[0] = nr; fresh[0] = true; } else { reg[1] = nr; fresh[1] = true; }
                    }
                    ++ev;
                }
            }
        }
    } else {
        // ================================ epilogue ================================
        asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(T2_EPI_REGS));
        const int row = warp * 32 + lane;
        const uint32_t lane_base = tmem_base + (uint32_t(warp * 32) << 16);
        double* my = part + (size_t(blockIdx.y) * gridDim.x + blockIdx.x) * (2 * NC * TILE) + row;   // (split, tile, rank)
        const uint32_t tempty_leader = PAIR ? map_to_cta(TEMPTY(0), 0) : TEMPTY(0);
        int cnt[2] = {0, 0};
        bool folded[2] = {false, false};
        for (int e = 0; e < n_ev; ++e) {
            const int r = e % 3, hf = (e + 1) & 1;
            mbar_wait(TFULL(r), uint32_t((e / 3) & 1));
            tc_fence_after();
            float* a = acc + hf * (NC * TILE) + row;
            const bool add = cnt[hf] > 0;
#pragma unroll 1
            for (int g = 0; g < NC / EG; ++g) {
                float x[EG];
                tmem_ld16(lane_base + uint32_t(r * NC + g * EG), x);
                tmem_wait_ld();
#pragma unroll
                for (int i = 0; i < EG; ++i) {
                    float* p = a + (g * EG + i) * TILE;
                    *p = add ? __fadd_rn(*p, x[i]) : x[i];
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(tempty_leader + 8u * r);
            ++cnt[hf];
            if (cnt[hf] == FOLD || e >= n_ev - 2) {
                double* o = my + size_t(hf) * (NC * TILE);
#pragma unroll 4
                for (int c = 0; c < NC; ++c) {
                    // later folds add in place with a fire-and-forget reduction (same thread, same address,
                    // program order: the FP64 sum is the one a load-add-store would give, without the load latency)
                    const double x = double(a[c * TILE]);
                    if (folded[hf]) asm volatile("red.global.add.f64 [%0], %1;" ::"l"(o + size_t(c) * TILE), "d"(x) : "memory");
                    else o[size_t(c) * TILE] = x;
                }
                folded[hf] = true;
                cnt[hf] = 0;
            }
        }
        if (n_ev == 0) {
            for (int c = 0; c < 2 * NC; ++c) my[size_t(c) * TILE] = 0.0;
        }
    }
    tc_fence_before();
    __syncthreads();
    if constexpr (PAIR) cluster_sync_all();   // no CTA leaves (or frees TMEM) while its peer may still signal it
    if (warp == MMA_WARP) {
        tc_fence_after();
        if constexpr (PAIR) tmem_dealloc2(tmem_base, TMEM_COLS); else tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(THREADS, 1)
gram_ts2_kernel(const __grid_constant__ ViewDev v, const float* __restrict__ center, int nb2, int64_t total_units,
                int64_t units_per_split, double* __restrict__ part) {
    gram_t2_body<true>(v, center, nb2, total_units, units_per_split, part);
}
// single-CTA form (128 x 128 tiles): the kernel for 64 <= n <= 128, where one CTA pair would leave half its rows as padding
__global__ void __launch_bounds__(THREADS, 1)
gram_ts1_kernel(const __grid_constant__ ViewDev v, const float* __restrict__ center, int nb, int64_t total_units,
                int64_t units_per_split, double* __restrict__ part) {
    gram_t2_body<false>(v, center, nb, total_units, units_per_split, part);
}

// Sum the split partials in split order and scatter tile (BI, BJ) (and its mirror) into G.
// Partial layout per (split, tile, rank): [half][col NC][row 128] doubles.
template <bool PAIR>
__global__ void gram_t2_reduce_kernel(const double* __restrict__ part, int splits, int nb2, int n, int accumulate, double* __restrict__ G) {
    constexpr int NCTA = PAIR ? 2 : 1, ROWS = PAIR ? T2_ROWS : TILE, NC = PAIR ? TILE : TILE / 2;
    int t = int(blockIdx.x) / NCTA, BI = 0;
    const int rank = int(blockIdx.x) % NCTA;
    while (t >= nb2 - BI) { t -= nb2 - BI; ++BI; }
    const int BJ = BI + t;
    constexpr int PER = 2 * NC * TILE;
    for (int e = blockIdx.y * blockDim.x + threadIdx.x; e < PER; e += gridDim.y * blockDim.x) {
        const int lr = e % TILE, c = (e / TILE) % NC, hf = e / (NC * TILE);
        const int i = BI * ROWS + rank * TILE + lr;
        const int j = BJ * ROWS + (PAIR ? 128 * (c >> 6) + 64 * hf + (c & 63) : 64 * hf + c);
        if (i >= n || j >= n) continue;
        if (BI == BJ && j < i) continue;
        double s = 0.0;
        if (BI == BJ) {   // diagonal tile: the kernel left D' = hi (hi/2 + lo)^T; G = D' + D'^T
            const int ii = i - BI * ROWS, jj = j - BJ * ROWS;
            const int rank2 = PAIR ? jj / TILE : 0, lr2 = jj % TILE;
            const int e2 = PAIR ? ((ii % TILE) / 64) * (NC * TILE) + (64 * (ii / TILE) + ii % 64) * TILE + lr2
                                : (ii / 64) * (NC * TILE) + (ii % 64) * TILE + lr2;
            const size_t b1 = size_t(blockIdx.x), b2 = size_t(blockIdx.x) - rank + rank2;
            for (int sp = 0; sp < splits; ++sp)
                s += part[(size_t(sp) * gridDim.x + b1) * PER + e] + part[(size_t(sp) * gridDim.x + b2) * PER + e2];
        } else {
            for (int sp = 0; sp < splits; ++sp) s += part[(size_t(sp) * gridDim.x + blockIdx.x) * PER + e];
        }
        if (accumulate) {
            G[size_t(i) * n + j] += s;
            if (i != j) G[size_t(j) * n + i] += s;
        } else {
            G[size_t(i) * n + j] = s;
            if (i != j) G[size_t(j) * n + i] = s;
        }
    }
}
