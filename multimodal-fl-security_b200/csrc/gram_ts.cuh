// K4, "TS" form: the A operand of every tcgen05.mma comes from TENSOR MEMORY, B from shared memory.
// This header holds what the TS kernels share -- the TMEM store / MMA wrappers, the producers' load and hi/lo
// split, the K permutation -- and is included by gram_tc.cu inside namespace flagg::tc before gram_ts2.cuh.
// (The first TS kernel, 128 x 128 tiles per CTA, 32.9 ms on the C4 shard, was superseded by the CTA-pair and
// single-CTA kernels of gram_ts2.cuh and removed in round 2; its measurements stay in DESIGN.md.)
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
