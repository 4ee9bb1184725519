// K1 dispatch + the large-n fallback.
//
// n <= FLAGG_NET_MAX (128): register compare-exchange networks (colselect_net.cuh,
// generated sizes in gen/network_sizes.h).
// n  > 128: thread-per-column MSB-first radix select, 4 bits per pass, with a private
// 16-bin histogram per column in shared memory (bank = lane, so no conflicts and no
// atomics).  The column is re-read from L2/HBM on every pass -- 8 passes for a median,
// 17 for a trimmed mean -- so this path is a correctness fallback, not a roofline
// kernel (DESIGN.md says so); every load is still a full coalesced 128-byte line.
#include <cstdlib>
#include <cstring>

#include "colselect_net.cuh"
#include "network_sizes.h"

namespace flagg {

#define X(N) int flagg_colselect_launch_##N(const flagg::ViewDev&, int, int, float*, int64_t, int, cudaStream_t);
}  // namespace flagg
FLAGG_NET_SIZES(X)
#undef X
namespace flagg {

struct NetEntry { int n; colselect_launch_fn fn; };
#define X(N) {N, flagg_colselect_launch_##N},
static const NetEntry g_nets[] = {FLAGG_NET_SIZES(X)};
#undef X

int colselect_tile_launch(const flagg_view* v, int mode, int k, float* out, cudaStream_t st);   // colselect_tile.cu

constexpr int RX_THREADS = 128;

__device__ __forceinline__ unsigned f2ukey(float x) { return unsigned(f2key(x)) ^ 0x80000000u; }
__device__ __forceinline__ float ukey2f(unsigned u) { return key2f(int(u ^ 0x80000000u)); }

// rank-th smallest ukey of the column (0-based).  hist: this thread's 16 bins, stride RX_THREADS.
__device__ unsigned radix_select(const float* const* sptr, int n, int64_t c, int rank, int* hist) {
    unsigned prefix = 0, mask = 0;
    for (int shift = 28; shift >= 0; shift -= 4) {
#pragma unroll
        for (int b = 0; b < 16; ++b) hist[b * RX_THREADS] = 0;
        int i = 0;
        for (; i + 8 <= n; i += 8) {
            float x[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) x[j] = sptr[i + j][c];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const unsigned u = f2ukey(x[j]);
                if ((u & mask) == prefix) hist[((u >> shift) & 15u) * RX_THREADS] += 1;
            }
        }
        for (; i < n; ++i) {
            const unsigned u = f2ukey(sptr[i][c]);
            if ((u & mask) == prefix) hist[((u >> shift) & 15u) * RX_THREADS] += 1;
        }
        unsigned digit = 15;
#pragma unroll
        for (int b = 0; b < 16; ++b) {
            const int cnt = hist[b * RX_THREADS];
            if (digit == 15 && b < 15) {
                if (rank < cnt) digit = b; else rank -= cnt;
            }
        }
        prefix |= digit << shift;
        mask |= 15u << shift;
    }
    return prefix;
}

__global__ void __launch_bounds__(RX_THREADS)
colselect_radix_kernel(ViewDev v, int mode, int k, float* __restrict__ out, int64_t total_units) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const float** sptr = reinterpret_cast<const float**>(smem_raw);                 // [n]
    int* hist_all = reinterpret_cast<int*>(smem_raw + sizeof(float*) * size_t(v.n)); // [16][RX_THREADS]
    const int n = v.n, tid = threadIdx.x;
    int* hist = hist_all + tid;
    const int64_t* __restrict__ ustart = v.unit_start[GRAN_128];
    int cur_p = -1;
    for (int64_t u = blockIdx.x; u < total_units; u += gridDim.x) {
        const int p = find_segment(ustart, v.P, u);
        if (p != cur_p) {
            __syncthreads();
            for (int i = tid; i < n; i += RX_THREADS) sptr[i] = v.seg_ptrs[size_t(i) * v.P + p];
            __syncthreads();
            cur_p = p;
        }
        const int64_t c = (u - ustart[p]) * RX_THREADS + tid;
        if (c >= v.seg_start[p + 1] - v.seg_start[p]) continue;
        float res;
        if (mode == FLAGG_MEDIAN) {
            const unsigned uk = radix_select(sptr, n, c, (n - 1) >> 1, hist);
            res = ukey2f(uk);
            const int kk = int(uk ^ 0x80000000u);
            bool has_nan = false;
            for (int i = 0; i < n; ++i) { const float x = sptr[i][c]; has_nan |= (x != x); }
            if (has_nan || kk == 0 || kk == -1) res = median_slow(sptr, 1, n, c, res);
        } else {
            const unsigned uL = radix_select(sptr, n, c, k, hist);
            const unsigned uH = radix_select(sptr, n, c, n - k - 1, hist);
            int ltL = 0, eqL = 0, ltH = 0, eqH = 0;
            double s = 0.0;
            for (int i = 0; i < n; ++i) {
                const float x = sptr[i][c];
                const unsigned uu = f2ukey(x);
                ltL += (uu < uL); eqL += (uu == uL);
                ltH += (uu < uH); eqH += (uu == uH);
                if (uu > uL && uu < uH) s += double(x);
            }
            const int keepL = min(ltL + eqL, n - k) - max(ltL, k);
            s += double(keepL) * double(ukey2f(uL));
            if (uH != uL) {
                const int keepH = min(ltH + eqH, n - k) - max(ltH, k);
                s += double(keepH) * double(ukey2f(uH));
            }
            res = float(s / double(n - 2 * k));
        }
        *out_addr(v, out, p, c) = res;
    }
}

}  // namespace flagg

using namespace flagg;

extern "C" int flagg_colwise_select(const flagg_view* v, int mode, int k, float* out, void* stream) {
    FLAGG_REQUIRE(v, "flagg_colwise_select: NULL view");
    if (v->dev.d == 0) return FLAGG_OK;   // an empty column shard (d < 1024 x ranks): nothing to write, `out` may be NULL
    FLAGG_REQUIRE(out || v->dev.out_ptrs, "flagg_colwise_select: NULL argument");
    FLAGG_REQUIRE(mode == FLAGG_MEDIAN || mode == FLAGG_TRIMMED_MEAN, "flagg_colwise_select: bad mode %d", mode);
    const_cast<flagg_view*>(v)->last_stream = stream;
    const int n = v->dev.n;
    if (mode == FLAGG_TRIMMED_MEAN)
        FLAGG_REQUIRE(k >= 0 && n - 2 * k >= 1, "flagg_colwise_select: trimmed mean needs n-2k >= 1 (n=%d k=%d)", n, k);
    const int64_t units = v->total_units[GRAN_128];
    if (units == 0) return FLAGG_OK;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int sms = device_sm_count();
    FLAGG_REQUIRE(sms > 0, "cannot query the SM count of the current device");
    if (n == 1) {  // median / trimmed mean of one client is that client (as float)
        for (size_t p = 0; p < v->seg_numel.size(); ++p)
            if (v->seg_numel[p] > 0)
                FLAGG_CUDA(cudaMemcpyAsync(v->out_ptrs.empty() ? static_cast<void*>(out + v->seg_start[p]) : v->out_ptrs[p], v->seg_ptrs[p], sizeof(float) * v->seg_numel[p],
                                           cudaMemcpyDeviceToDevice, st));
        return FLAGG_OK;
    }
    if (n <= FLAGG_NET_MAX) {
        for (int64_t len : v->seg_numel)
            if (len >= (int64_t(1) << 30)) {
                set_error("flagg_colwise_select: a tensor of %lld elements exceeds the 2^30-element limit of the network path", (long long)len);
                return FLAGG_EUNSUPPORTED;
            }
        const NetEntry* e = nullptr;
        for (const NetEntry& cand : g_nets)
            if (cand.n >= n) { e = &cand; break; }
        FLAGG_REQUIRE(e, "flagg_colwise_select: no network for n=%d", n);
        const bool exact = (e->n == n);
        const int variant = (mode == FLAGG_TRIMMED_MEAN) ? (exact ? CS_TRIMMED_EXACT : CS_TRIMMED_ANY)
                                                         : (exact ? CS_MEDIAN_EXACT : CS_MEDIAN_ANY);
        // TMA-staged feed for the exact sizes on 16-byte aligned tensors; FLAGG_K1_IMPL=ldg|bulk overrides (tests, tuning)
        const char* impl = getenv("FLAGG_K1_IMPL");
        const bool want_bulk = impl ? !strcmp(impl, "bulk") : true;
        if (impl && !strcmp(impl, "bulk") && !(exact && (v->dense || v->gather))) {
            set_error("flagg_colwise_select: FLAGG_K1_IMPL=bulk needs n in the generated sizes and a dense or 512-byte-congruent view");
            return FLAGG_EUNSUPPORTED;
        }
        // dense matrix -> one 2-D box per tile; separately allocated 512-byte-congruent rows -> gather4 (four clients per request)
        const int flag = (want_bulk && exact && v->align_bytes >= 16 && (v->dense || v->gather)) ? CS_BULK : 0;
        const bool wide = flag && e->n <= CSB_WIDE_MAX_N && (v->dense ? v->dev.dense_cols == 256 : v->dev.gather_cols == 256);
        return e->fn(v->dev, variant | flag | (wide ? CS_WIDE : 0), k, out, wide ? v->total_units[GRAN_256] : units, sms, st);
    }
    {   // shared-memory tile select (colselect_tile.cu): one HBM read instead of 8-17.  Measured against the radix
        // fallback (tools/k1_large_n.py, profiles/r01_k1_large_n.txt) it wins for the median up to n ~ 1500 and for
        // the trimmed mean (three tile loads) from n ~ 400; FLAGG_K1_IMPL=tile|radix forces one.
        const char* impl = getenv("FLAGG_K1_IMPL");
        const bool force_tile = impl && !strcmp(impl, "tile"), force_radix = impl && !strcmp(impl, "radix");
        const bool pick = (mode == FLAGG_MEDIAN) ? n <= 1500 : (n >= 400 && n <= 1500);
        if (force_tile || (pick && !force_radix)) {
            const int rc = colselect_tile_launch(v, mode, k, out, st);
            if (rc != FLAGG_EUNSUPPORTED) return rc;
        }
    }
    const size_t smem = sizeof(float*) * size_t(n) + sizeof(int) * 16 * RX_THREADS;
    FLAGG_REQUIRE(smem <= 200 * 1024, "flagg_colwise_select: n=%d too large", n);
    FLAGG_CUDA(cudaFuncSetAttribute(colselect_radix_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    const int grid = int(std::min<int64_t>(units, int64_t(sms) * 16));
    colselect_radix_kernel<<<grid, RX_THREADS, smem, st>>>(v->dev, mode, k, out, units);
    FLAGG_CUDA(cudaGetLastError());
    return FLAGG_OK;
}
