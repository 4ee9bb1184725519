// K1 for n beyond the register networks (128 < n <= ~7000): column-wise order statistics with the
// column tile resident in SHARED memory, one read of X from HBM (two extra L2 reads for the
// trimmed mean).  Replaces torch.stack + torch.median(dim=0) (trimmed_mean.py:160-163) and
// torch.sort + slice + mean (:79-88) like the network kernels do; the old thread-per-column
// radix select (colselect.cu) re-read every column 8-17 times from L2/HBM and ran at 3-9 % of
// the HBM roofline (tools/k1_large_n.py).
//
// One WARP owns a tile of 8 columns x n rows: lane = (column c = lane & 7, row group rg = lane >> 3)
// owns rows 4j + rg of column c in slots smem[(4j + rg) * 8 + c] -- every lane always hits its own
// bank, whatever j, so no access below ever conflicts.  The tile arrives by cp.async (16-byte
// pieces = 4 columns of one row; 8 lanes of a request cover 4 rows x 32 B, sector-exact).
//
// Selection = MSB-first radix select with sampled 8-bit windows, shared-memory atomics for the
// histogram and in-place compaction (select_rank below).  Columns advance independently (per-lane
// window, rank, counts); the warp loops until all 8 columns are resolved.  Exact: the result is an
// element of the column, bit for bit; columns holding a NaN or resolving to a zero are re-resolved
// by median_slow (torch's tie rules).
//
// Status (B200, tools/k1_large_n.py -> profiles/r01_k1_large_n.txt): 635-845 GB/s = 10-13 % of the
// measured HBM rate for 200 <= n <= 1000 against 5.8-9 % for the radix fallback.  ncu: 1534 warp
// instructions per column at n = 1000 where the HBM time of a column allows ~700, and 5 resident
// warps per SM (one 40 KB tile + histogram each) leave the dependent sweeps latency-bound (issue
// slots 23 % busy).  The select is ~14 thread instructions per element per sweep, two sweeps per
// round: a radix-style select cannot reach the roofline at these n; see DESIGN.md section 8.
#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "colselect_net.cuh"

namespace flagg {
namespace tile {

constexpr int TCOLS = 8;          // columns per warp tile
constexpr int MAX_WARPS = 8;
constexpr unsigned FULL = 0xffffffffu;

__device__ __forceinline__ unsigned f2ukey(float x) { return unsigned(f2key(x)) ^ 0x80000000u; }
__device__ __forceinline__ float ukey2f(unsigned u) { return key2f(int(u ^ 0x80000000u)); }

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async4(uint32_t dst, const void* src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
    asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
}
// sum over the 4 lanes (row groups) that share a column
__device__ __forceinline__ int colsum(int x) {
    x += __shfl_xor_sync(FULL, x, 8);
    x += __shfl_xor_sync(FULL, x, 16);
    return x;
}

// Bring columns [c0, c0 + TCOLS) of tensor p, all n rows, into the warp's tile (columns >= len read 0).
__device__ __forceinline__ void load_tile(const ViewDev& v, int p, int64_t c0, int ncols, bool vec16, float* tile, int lane) {
    const int n = v.n;
    const uint32_t tbase = uint32_t(__cvta_generic_to_shared(tile));
    __syncwarp();                                   // every lane is done with the previous contents
    if (vec16 && ncols == TCOLS) {
        const int half = lane & 1;
        for (int r = lane >> 1; r < n; r += 16) {
            const float* src = v.seg_ptrs[size_t(r) * v.P + p] + c0 + half * 4;
            cp_async16(tbase + uint32_t(r * TCOLS + half * 4) * 4u, src);
        }
    } else {
        const int c = lane & 7, rg = lane >> 3;
        for (int r = rg; r < n; r += 4) {
            if (c < ncols) cp_async4(tbase + uint32_t(r * TCOLS + c) * 4u, v.seg_ptrs[size_t(r) * v.P + p] + c0 + c);
            else tile[r * TCOLS + c] = 0.f;
        }
    }
    cp_async_wait_all();
    __syncwarp();
}

constexpr int HBINS = 260;          // 256 window bins + below (0) + above (257), padded to 4 x 65
constexpr int HSPAN = 65;           // bins scanned per lane of a column (65 = 1 mod 4: the four lanes stay on distinct banks)
constexpr int HIST_WORDS = HBINS * TCOLS;

// rank-th smallest key (0-based) of each of the warp's 8 columns; `mine` = this lane's slot count.
// On entry the tile holds raw floats; it is destroyed.  *has_nan: the column holds a NaN.
//
// Radix select with an 8-bit window placed by a SAMPLE: the first <= 8 candidates of each of the
// column's four lanes give AND / OR; the window is the 8 bits from their highest differing bit
// down.  One sweep histograms every candidate into 256 window bins + "below" / "above" (keys whose
// bits over the window differ from the sample's) with shared-memory atomics, the four lanes scan
// the histogram for the bin holding the rank, a second sweep compacts that bin's members to the
// front of each lane's slots.  Sample members always land in window bins, so every round
// strictly shrinks the candidate set; Gaussian-like columns finish in two rounds.
__device__ __forceinline__ unsigned select_rank(float* tile, int* hist, int lane, int mine, int rank, bool* has_nan) {
    const int c = lane & 7, rg = lane >> 3;
    unsigned* slot = reinterpret_cast<unsigned*>(tile) + (rg * TCOLS + c);   // + j * 4 * TCOLS
    constexpr int STR = 4 * TCOLS;
    int* hcol = hist + c;                                                    // bin b of this column: hcol[b * TCOLS]
    int m = mine, nanf = 0;
    unsigned result = 0u;
    bool done = false, raw = true;
    while (__any_sync(FULL, !done)) {
        // ---- sample -> window
        unsigned a = 0xffffffffu, o = 0u;
        const int ns = done ? 0 : min(m, 8);
        for (int j = 0; j < ns; ++j) {
            const unsigned v = slot[j * STR];
            const unsigned u = raw ? f2ukey(__uint_as_float(v)) : v;
            a &= u;
            o |= u;
        }
        a &= __shfl_xor_sync(FULL, a, 8);  a &= __shfl_xor_sync(FULL, a, 16);
        o |= __shfl_xor_sync(FULL, o, 8);  o |= __shfl_xor_sync(FULL, o, 16);
        const unsigned diff = a ^ o;
        const int tb = diff ? 31 - __clz(diff) : 7;
        const int lo = max(tb - 7, 0), hs = lo + 8;                          // window = bits [lo, lo + 8)
        const unsigned ph = __funnelshift_rc(a, 0u, hs);                     // the sample's bits above the window (0 when hs == 32)
        // ---- clear the histogram, sweep 1: histogram (and keys, first round)
        __syncwarp();
        for (int i = lane; i < HIST_WORDS; i += 32) hist[i] = 0;
        __syncwarp();
        if (!done) {
            for (int j = 0; j < m; ++j) {
                const unsigned v = slot[j * STR];
                const unsigned u = raw ? f2ukey(__uint_as_float(v)) : v;
                if (raw) { slot[j * STR] = u; nanf |= int(u == 0xffffffffu); }
                const unsigned h = __funnelshift_rc(u, 0u, hs);
                const int bin = (h == ph) ? int((u >> lo) & 255u) + 1 : (h < ph ? 0 : 257);
                atomicAdd(hcol + bin * TCOLS, 1);
            }
        }
        __syncwarp();
        // ---- scan: lane rg of a column owns bins [65 rg, 65 rg + 65)
        int mysum = 0;
        for (int i = 0; i < HSPAN; ++i) mysum += hcol[(rg * HSPAN + i) * TCOLS];
        int pre = 0;                                                         // candidates in bins before my span
        {
            const int s1 = __shfl_xor_sync(FULL, mysum, 8);
            const int p1 = (rg & 1) ? s1 : 0;
            const int pair = mysum + s1;
            const int s2 = __shfl_xor_sync(FULL, pair, 16);
            pre = p1 + ((rg & 2) ? s2 : 0);
        }
        int sel = -1, below = 0;                                             // selected bin and the candidates below it
        if (!done && rank >= pre && rank < pre + mysum) {
            int acc = pre;
            for (int i = 0; i < HSPAN; ++i) {
                const int cnt = hcol[(rg * HSPAN + i) * TCOLS];
                if (sel < 0 && rank < acc + cnt) { sel = rg * HSPAN + i; below = acc; }
                acc += cnt;
            }
        }
        sel = max(sel, __shfl_xor_sync(FULL, sel, 8));    sel = max(sel, __shfl_xor_sync(FULL, sel, 16));
        below = max(below, __shfl_xor_sync(FULL, below, 8)); below = max(below, __shfl_xor_sync(FULL, below, 16));
        if (!done) rank -= below;
        // ---- sweep 2: keep the members of the selected bin
        int w = 0;
        if (!done) {
            for (int j = 0; j < m; ++j) {
                const unsigned u = slot[j * STR];
                const unsigned h = __funnelshift_rc(u, 0u, hs);
                const int bin = (h == ph) ? int((u >> lo) & 255u) + 1 : (h < ph ? 0 : 257);
                if (bin == sel) { slot[w * STR] = u; ++w; }
            }
            m = w;
        }
        const int total = colsum(m);
        unsigned cand = (!done && m > 0) ? slot[0] : 0u;
        cand = max(cand, __shfl_xor_sync(FULL, cand, 8));
        cand = max(cand, __shfl_xor_sync(FULL, cand, 16));
        // a window bin at lo == 0 fixes every bit; one survivor is the answer in any case
        if (!done && (total == 1 || (lo == 0 && sel >= 1 && sel <= 256))) { result = cand; done = true; }
        raw = false;
    }
    *has_nan = colsum(nanf) != 0;
    return result;
}

// grid-stride over 8-column tiles; warps are independent.  smem: MAX_WARPS tiles of cap4 * 4 * 8 floats.
__global__ void __launch_bounds__(MAX_WARPS * 32, 1)
colselect_tile_kernel(ViewDev v, int mode, int k, float* __restrict__ out, int64_t total_units, int warps, int tile_floats, int vec16) {
    extern __shared__ __align__(16) float smem_f[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp >= warps) return;
    float* tile = smem_f + size_t(warp) * tile_floats;
    int* hist = reinterpret_cast<int*>(smem_f + size_t(warps) * tile_floats) + warp * HIST_WORDS;
    const int n = v.n, c = lane & 7, rg = lane >> 3;
    const int mine = (n - rg + 3) >> 2;                      // rows rg, rg + 4, ... < n
    const int64_t* __restrict__ ustart = v.unit_start[GRAN_128];
    const int64_t ntiles = total_units * (128 / TCOLS);
    for (int64_t t = int64_t(blockIdx.x) * warps + warp; t < ntiles; t += int64_t(gridDim.x) * warps) {
        const int64_t u = t / (128 / TCOLS);
        const int sub = int(t % (128 / TCOLS));
        const int p = find_segment(ustart, v.P, u);
        const int64_t len = v.seg_start[p + 1] - v.seg_start[p];
        const int64_t c0 = (u - ustart[p]) * 128 + sub * TCOLS;
        if (c0 >= len) continue;
        const int ncols = (len - c0 < TCOLS) ? int(len - c0) : TCOLS;
        const bool ok = c < ncols;
        load_tile(v, p, c0, ncols, vec16 != 0, tile, lane);
        float res;
        if (mode == FLAGG_MEDIAN) {
            bool nan;
            const unsigned uk = select_rank(tile, hist, lane, mine, (n - 1) >> 1, &nan);
            res = ukey2f(uk);
            // torch's tie rules for NaN columns and mixed-sign zeros (keys 0x7fffffff / 0x80000000 are -0 / +0)
            if (ok && rg == 0 && (nan || uk == 0x80000000u || uk == 0x7fffffffu)) res = median_slow(v.seg_ptrs + p, v.P, n, c0 + c, res);
        } else {
            bool nan;
            const unsigned uL = select_rank(tile, hist, lane, mine, k, &nan);
            load_tile(v, p, c0, ncols, vec16 != 0, tile, lane);
            const unsigned uH = select_rank(tile, hist, lane, mine, n - k - 1, &nan);
            load_tile(v, p, c0, ncols, vec16 != 0, tile, lane);
            // same arithmetic as the radix fallback: FP64 sum of the values strictly between, plus the kept copies of the two boundary values
            int ltL = 0, eqL = 0, ltH = 0, eqH = 0;
            double s = 0.0;
            const float* col = tile + rg * TCOLS + c;
            for (int j = 0; j < mine; ++j) {
                const float x = col[j * 4 * TCOLS];
                const unsigned uu = f2ukey(x);
                ltL += (uu < uL); eqL += (uu == uL);
                ltH += (uu < uH); eqH += (uu == uH);
                if (uu > uL && uu < uH) s += double(x);
            }
            ltL = colsum(ltL); eqL = colsum(eqL); ltH = colsum(ltH); eqH = colsum(eqH);
            s += __shfl_xor_sync(FULL, s, 8);
            s += __shfl_xor_sync(FULL, s, 16);
            const int keepL = min(ltL + eqL, n - k) - max(ltL, k);
            s += double(keepL) * double(ukey2f(uL));
            if (uH != uL) {
                const int keepH = min(ltH + eqH, n - k) - max(ltH, k);
                s += double(keepH) * double(ukey2f(uH));
            }
            res = float(s / double(n - 2 * k));
        }
        if (ok && rg == 0) *out_addr(v, out, p, c0 + c) = res;
        __syncwarp();
    }
}

}  // namespace tile

// Host launch; returns FLAGG_EUNSUPPORTED when n does not fit one warp tile in shared memory.
int colselect_tile_launch(const flagg_view* v, int mode, int k, float* out, cudaStream_t st) {
    using namespace tile;
    const int n = v->dev.n;
    const int cap4 = (n + 3) / 4;
    const int tile_floats = cap4 * 4 * TCOLS;
    const size_t tile_bytes = sizeof(float) * size_t(tile_floats) + sizeof(int) * HIST_WORDS;   // + the warp's histogram
    const size_t budget = 220 * 1024;
    if (tile_bytes > budget) return FLAGG_EUNSUPPORTED;
    const int warps = int(std::min<size_t>(MAX_WARPS, budget / tile_bytes));
    const size_t smem = tile_bytes * warps;   // kernel layout: all tiles, then all histograms
    static size_t attr_smem = 0;
    if (smem > attr_smem) {
        FLAGG_CUDA(cudaFuncSetAttribute(colselect_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(budget)));
        attr_smem = budget;
    }
    const int64_t units = v->total_units[GRAN_128];
    const int64_t ntiles = units * (128 / TCOLS);
    const int sms = device_sm_count();
    FLAGG_REQUIRE(sms > 0, "cannot query the SM count of the current device");
    const int ctas_per_sm = int(std::max<size_t>(1, std::min<size_t>(4, budget / smem)));
    const int grid = int(std::min<int64_t>((ntiles + warps - 1) / warps, int64_t(sms) * ctas_per_sm));
    colselect_tile_kernel<<<grid, MAX_WARPS * 32, smem, st>>>(v->dev, mode, k, out, units, warps, tile_floats, v->align_bytes >= 16 ? 1 : 0);
    FLAGG_CUDA(cudaGetLastError());
    return FLAGG_OK;
}

}  // namespace flagg
