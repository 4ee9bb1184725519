// K4, "TS" variant: the A operand of every tcgen05.mma comes from TENSOR MEMORY, B from
// shared memory.  Included by gram_tc.cu inside namespace flagg::tc (shares its PTX wrappers,
// KbIter and the reduce kernel).  Replaces krum.py:92-97 like the SS kernel does.
//
// Why: with both operands in shared memory a 128x128x8 tf32 MMA reads 8 KB of smem in its 64
// cycles -- the whole 128 B/clk the SM has -- and the producers' hi/lo stores (85 B/clk at
// full MMA rate) come on top: the SS kernel is shared-memory-bandwidth bound at ~60 % of the
// tensor pipe (measured 42 %, l1tex 65 %: profiles/r01_ncu_k4_gram_tc_summary.txt).  Here the
// producers write the A tile's hi/lo halves straight from registers into TMEM
// (tcgen05.st.16x256b), so per 32-column k-block shared memory sees only B: 32 KB of stores
// + 48 KB of MMA reads = 104 B/clk at full rate.
//
// K permutation.  tcgen05.st.16x256b gives thread t of a warp the elements
// (lane t/4 + 8h, column 8g + 2(t%4) + e) in register 4g + 2h + e.  A thread loads 8
// CONSECUTIVE columns 8q..8q+7 (q = t%4) of its two rows (one 256-bit or two 128-bit loads
// per row, four lanes cover a row's 128-byte line), so inside a k-block actual column
// 8q + 2g + e sits at K position 8g + 2q + e.  B is written to shared memory under the same
// permutation (8-byte stores), and a permutation of K common to both operands leaves the
// product unchanged.
//
// Pipeline: NSTG stages of SK K-positions; stage = A_hi/A_lo in TMEM columns
// [384 + 2*SK*st, +2*SK) and the matching K range of a 128B-swizzled B_hi/B_lo slot in smem.
//   warps 0-3   epilogue (same two-level accumulation as the SS kernel: TMEM D -> TMEM ACC -> FP64)
//   warps 4-11  producers: warp w owns tile rows 32*(w%4) + 16*((w-4)/4) + [0,16) of A and of B
//   warp 12     MMA issuer
#pragma once

constexpr int TS_SK = 16;                        // K positions per stage (2 k-steps of 8)
constexpr int TS_SUBS = BK / TS_SK;              // stages per 32-column k-block
constexpr int TS_NSTG = 128 / (2 * TS_SK);       // TMEM A stages: 128 columns hold hi+lo of NSTG stages
constexpr int TS_SLOTS = TS_NSTG / TS_SUBS;      // smem B slots, one k-block each
constexpr int TS_SLOT_BYTES = 2 * OPER_BYTES;    // B_hi, B_lo
constexpr int TS_TMEM_A0 = 3 * TILE;             // after D0, D1, ACC
constexpr int TS_SMEM_BYTES = TS_SLOTS * TS_SLOT_BYTES + 1024 + 256;
// Register budget (13 warps are allocated 128 registers per thread at launch): the four epilogue
// warps give 64 each back, which is exactly what the eight producer warps take (+32 each).
constexpr int TS_EPI_REGS = 64, TS_PROD_REGS = 160;
static_assert((128 - TS_EPI_REGS) * EPI_WARPS >= (TS_PROD_REGS - 128) * PROD_WARPS, "setmaxnreg budget");
static_assert(TS_SK % UMMA_K == 0 && BK % TS_SK == 0 && TS_NSTG % TS_SUBS == 0, "stage shape");
static_assert(TS_TMEM_A0 + TS_NSTG * 2 * TS_SK <= TMEM_COLS, "TMEM budget");

__device__ __forceinline__ void umma_tf32_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d_tmem),
        "r"(a_tmem), "l"(b_desc), "r"(IDESC), "r"(accumulate)
        : "memory");
}
// 16 lanes x (8*NG) columns; u[4g + 2h + e] = element (lane + t/4 + 8h, col + 8g + 2(t%4) + e)
__device__ __forceinline__ void tmem_st_16x256b_x2(uint32_t taddr, const uint32_t (&u)[8]) {
    asm volatile("tcgen05.st.sync.aligned.16x256b.x2.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr), "r"(u[0]),
                 "r"(u[1]), "r"(u[2]), "r"(u[3]), "r"(u[4]), "r"(u[5]), "r"(u[6]), "r"(u[7])
                 : "memory");
}

struct TsRegs {
    float a[2][8], b[2][8], c[8];
};
struct TsRows {
    const float* a[2];
    const float* b[2];
    const float* cbase;
    bool v8;   // every pointer above is 32-byte aligned -> 256-bit loads
};

__device__ __forceinline__ void ldg_stream8(const float* p, float (&v)[8]) {
    asm volatile("ld.global.nc.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7])
                 : "l"(p));
}
__device__ __forceinline__ void ldg_stream44(const float* p, float (&v)[8]) {
    const float4 x = ldg_stream4(reinterpret_cast<const float4*>(p)), y = ldg_stream4(reinterpret_cast<const float4*>(p) + 1);
    v[0] = x.x; v[1] = x.y; v[2] = x.z; v[3] = x.w; v[4] = y.x; v[5] = y.y; v[6] = y.z; v[7] = y.w;
}

// Global loads of one k-block for this thread: columns 8q..8q+7 of two A rows (and two B rows
// off the diagonal) and of the centre; columns beyond the unit read as zero.
__device__ __forceinline__ void ts_load(const KbIter& it, const TsRows& rows, bool diag, int q, TsRegs& r) {
    const int64_t col = it.c0 + it.ks + q * 8;
    const int rem = it.len - it.ks - q * 8;
    if (rem >= 8) {
        if (rows.v8) {
#pragma unroll
            for (int h = 0; h < 2; ++h) ldg_stream8(rows.a[h] + col, r.a[h]);
            if (!diag) {
#pragma unroll
                for (int h = 0; h < 2; ++h) ldg_stream8(rows.b[h] + col, r.b[h]);
            }
        } else {
#pragma unroll
            for (int h = 0; h < 2; ++h) ldg_stream44(rows.a[h] + col, r.a[h]);
            if (!diag) {
#pragma unroll
                for (int h = 0; h < 2; ++h) ldg_stream44(rows.b[h] + col, r.b[h]);
            }
        }
        if (rows.cbase) {
            const float* cp = rows.cbase + col;
            if ((reinterpret_cast<uintptr_t>(cp) & 15) == 0) {
                const float4 x = __ldg(reinterpret_cast<const float4*>(cp)), y = __ldg(reinterpret_cast<const float4*>(cp) + 1);
                r.c[0] = x.x; r.c[1] = x.y; r.c[2] = x.z; r.c[3] = x.w; r.c[4] = y.x; r.c[5] = y.y; r.c[6] = y.z; r.c[7] = y.w;
            } else {
#pragma unroll
                for (int j = 0; j < 8; ++j) r.c[j] = __ldg(cp + j);
            }
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) r.c[j] = 0.f;
        }
    } else {
        auto part = [&](const float* row, float (&v)[8]) {
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = (row && j < rem) ? __ldg(row + col + j) : 0.f;
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

// (x - c) -> hi, lo bit patterns (mask = 0xffffe000, or 0 for a padding row)
__device__ __forceinline__ void ts_split(const float (&x)[8], const float (&c)[8], uint32_t mask, uint32_t (&hi)[8], uint32_t (&lo)[8]) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const float t = __fsub_rn(x[j], c[j]);
        hi[j] = __float_as_uint(t) & mask;
        lo[j] = __float_as_uint(__fsub_rn(t, __uint_as_float(__float_as_uint(t) & 0xffffe000u))) & mask;
    }
}

__global__ void __launch_bounds__(THREADS, 1)
gram_ts_kernel(ViewDev v, const float* __restrict__ center, int nb, int64_t total_units, int64_t units_per_split,
               double* __restrict__ part) {
    extern __shared__ unsigned char smem_raw[];
    const uint32_t sbase = (smem_u32(smem_raw) + 1023u) & ~1023u;
    unsigned char* smem = smem_raw + (sbase - smem_u32(smem_raw));
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + TS_SLOTS * TS_SLOT_BYTES);
    // bars[0..NSTG) full, [NSTG..2NSTG) empty, then 2 tmem_full, 2 tmem_empty
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * TS_NSTG + 4);
    int* nkb_slot = reinterpret_cast<int*>(bars + 2 * TS_NSTG + 6);
    const uint32_t bar0 = sbase + TS_SLOTS * TS_SLOT_BYTES;
    auto FULL = [&](int s) { return bar0 + 8u * s; };
    auto EMPTY = [&](int s) { return bar0 + 8u * (TS_NSTG + s); };
    auto TFULL = [&](int b) { return bar0 + 8u * (2 * TS_NSTG + b); };
    auto TEMPTY = [&](int b) { return bar0 + 8u * (2 * TS_NSTG + 2 + b); };

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    int t = blockIdx.x, bi = 0;
    while (t >= nb - bi) { t -= nb - bi; ++bi; }
    const int bj = bi + t;
    const bool diag = (bi == bj);
    const int64_t u0 = int64_t(blockIdx.y) * units_per_split;
    const int64_t u1 = min(total_units, u0 + units_per_split);

    if (tid == 0) {
        for (int s = 0; s < TS_NSTG; ++s) { mbar_init(FULL(s), PROD_WARPS); mbar_init(EMPTY(s), 1); }
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
        asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(TS_PROD_REGS));
        const int qd = warp & 3, half = (warp - EPI_WARPS) >> 2;
        const int lane0 = 32 * qd + 16 * half;          // first TMEM lane / tile row of this warp
        const int rloc = lane0 + (lane >> 2), q = lane & 3;
        // byte offset of K position 2q of tile row rloc inside a swizzled operand tile; K position
        // 8g + 2q is the same offset ^ (g << 5) (16-byte chunk index 2g + q/2, XOR-swizzled by row & 7)
        const uint32_t off0 = uint32_t(rloc >> 3) * 1024u + uint32_t(rloc & 7) * 128u +
                              (uint32_t((q >> 1) ^ (rloc & 7)) << 4) + uint32_t(q & 1) * 8u;
        bool okA[2], okB[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            okA[h] = (bi * TILE + rloc + 8 * h) < v.n;
            okB[h] = (bj * TILE + rloc + 8 * h) < v.n;
        }
        // rows >= n of B stay zero for the whole kernel (their stores are predicated off)
        for (int s = 0; s < TS_SLOTS; ++s)
            for (int o = 0; o < 2; ++o)
#pragma unroll
                for (int h = 0; h < 2; ++h)
#pragma unroll
                    for (int g = 0; g < 4; ++g)
                        *reinterpret_cast<uint2*>(smem + s * TS_SLOT_BYTES + o * OPER_BYTES + ((off0 + h * 1024u) ^ (uint32_t(g) << 5))) =
                            make_uint2(0u, 0u);
        TsRows rows;
        int cur_p = -1;
        auto refresh_rows = [&](int p) {
            uintptr_t bits = 0;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int ra = min(bi * TILE + rloc + 8 * h, v.n - 1), rb = min(bj * TILE + rloc + 8 * h, v.n - 1);
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
        TsRegs r0, r1;
        auto fetch = [&](TsRegs& r) {
            if (it.valid()) {
                if (it.p != cur_p) refresh_rows(it.p);
                ts_load(it, rows, diag, q, r);
                it.next(v);
            }
        };
        fetch(r0);
        fetch(r1);
        int stage = 0;
        uint32_t phase = 1;
        const uint32_t a_lane = tmem_base + (uint32_t(lane0) << 16) + TS_TMEM_A0;
        auto emit = [&](const TsRegs& r) {
            uint32_t ahi[2][8], alo[2][8], bhi[2][8], blo[2][8];
#pragma unroll
            for (int h = 0; h < 2; ++h) ts_split(r.a[h], r.c, okA[h] ? 0xffffe000u : 0u, ahi[h], alo[h]);
            if (!diag) {
#pragma unroll
                for (int h = 0; h < 2; ++h) ts_split(r.b[h], r.c, 0xffffe000u, bhi[h], blo[h]);
            } else {   // B rows are the A rows (okB == okA); stores of padding rows are predicated off below
#pragma unroll
                for (int h = 0; h < 2; ++h)
#pragma unroll
                    for (int j = 0; j < 8; ++j) { bhi[h][j] = ahi[h][j]; blo[h][j] = alo[h][j]; }
            }
#pragma unroll
            for (int sub = 0; sub < TS_SUBS; ++sub) {
                mbar_wait(EMPTY(stage), phase);
                tc_fence_after();
                // A -> TMEM: register 4*gl + 2*h + e of the fragment = row h, value j = 2*(g0+gl) + e
                constexpr int NG = TS_SK / 8;
                static_assert(NG == 2, "fragment wrappers below are .x2");
                uint32_t fh[8], fl[8];
#pragma unroll
                for (int gl = 0; gl < NG; ++gl)
#pragma unroll
                    for (int h = 0; h < 2; ++h)
#pragma unroll
                        for (int e = 0; e < 2; ++e) {
                            fh[4 * gl + 2 * h + e] = ahi[h][2 * (sub * NG + gl) + e];
                            fl[4 * gl + 2 * h + e] = alo[h][2 * (sub * NG + gl) + e];
                        }
                tmem_st_16x256b_x2(a_lane + uint32_t(stage * 2 * TS_SK), fh);
                tmem_st_16x256b_x2(a_lane + uint32_t(stage * 2 * TS_SK + TS_SK), fl);
                // B -> smem (same K permutation): value pair (2g, 2g+1) -> K positions 8g + 2q + {0,1}
                const uint32_t sb = sbase + uint32_t(stage / TS_SUBS) * TS_SLOT_BYTES;
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const bool ok = okB[h];
#pragma unroll
                    for (int gl = 0; gl < NG; ++gl) {
                        const int g = sub * NG + gl;
                        const uint32_t addr = sb + ((off0 + h * 1024u) ^ (uint32_t(g) << 5));
                        if (ok) {
                            asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(addr), "r"(bhi[h][2 * g]), "r"(bhi[h][2 * g + 1]) : "memory");
                            asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(addr + OPER_BYTES), "r"(blo[h][2 * g]), "r"(blo[h][2 * g + 1]) : "memory");
                        }
                    }
                }
                tmem_wait_st();
                tc_fence_before();
                fence_proxy_async();
                __syncwarp();
                if (lane == 0) mbar_arrive(FULL(stage));
                if (++stage == TS_NSTG) { stage = 0; phase ^= 1; }
            }
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
        uint32_t full_phase = 0, tempty_phase = 3u;
        for (int kb = 0; kb < nkb; ++kb) {
            if (inb == 0) {
                mbar_wait(TEMPTY(buf), (tempty_phase >> buf) & 1u);
                tempty_phase ^= 1u << buf;
                tc_fence_after();
            }
            const bool last_in_drain = (inb == KC_BLOCKS - 1) || (kb == nkb - 1);
            const uint32_t d = tmem_base + uint32_t(buf * TILE);
#pragma unroll
            for (int sub = 0; sub < TS_SUBS; ++sub) {
                mbar_wait(FULL(s), full_phase);
                tc_fence_after();
                if (lane == 0) {
                    const uint32_t b_hi = sbase + uint32_t(s / TS_SUBS) * TS_SLOT_BYTES, b_lo = b_hi + OPER_BYTES;
                    const uint32_t a_hi = tmem_base + TS_TMEM_A0 + uint32_t(s * 2 * TS_SK), a_lo = a_hi + TS_SK;
#pragma unroll
                    for (int k = 0; k < TS_SK / UMMA_K; ++k) {
                        const uint32_t ko = uint32_t((s % TS_SUBS) * TS_SK + k * UMMA_K) * 4u;   // K byte offset inside the 128-byte row
                        const uint32_t ka = uint32_t(k * UMMA_K);
                        umma_tf32_ts(d, a_hi + ka, make_desc(b_hi + ko), (inb > 0 || sub > 0 || k > 0) ? 1u : 0u);
                        umma_tf32_ts(d, a_hi + ka, make_desc(b_lo + ko), 1u);
                        umma_tf32_ts(d, a_lo + ka, make_desc(b_hi + ko), 1u);
                    }
                    umma_commit(EMPTY(s));
                    if (last_in_drain && sub == TS_SUBS - 1) umma_commit(TFULL(buf));
                }
                __syncwarp();
                if (++s == TS_NSTG) { s = 0; full_phase ^= 1; }
            }
            if (last_in_drain) { inb = 0; buf ^= 1; } else { ++inb; }
        }
    } else {
        // ================================ epilogue ================================
        asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(TS_EPI_REGS));
        epilogue_role(tmem_base, part, ndrain, warp, lane, TFULL(0), TEMPTY(0));
    }
    tc_fence_before();
    __syncthreads();
    if (warp == MMA_WARP) {
        tc_fence_after();
        tmem_dealloc(tmem_base, TMEM_COLS);
    }
}
