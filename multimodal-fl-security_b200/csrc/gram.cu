// K4: centred Gram matrix G = (X - c)(X - c)^T in FP64 (upper triangle computed, mirrored)
// K5: Krum distances and scores from G.
//
// Replaces the n(n-1)/2 Python iterations of torch.norm(flat_i - flat_j).item() at
// krum.py:92-97 (O(n^2 d) bytes of traffic, one host sync per pair) and the per-client
// np.sort / np.sum at krum.py:122-129.
//
// Precision: rows are full weight vectors, so ||x||^2 >> ||x_i - x_j||^2.
// Every element is centred on a common vector c before it is multiplied (distances are
// translation invariant; the caller passes a robust centre, see flagg.engine.gram), products are accumulated in fp32 for at most 1024 columns and
// then folded into FP64; D^2 and sqrt are formed in FP64.
//
// This file is the CUDA-core (FFMA) implementation: 64x64 output tiles, split-K across
// the grid with a deterministic two-stage reduction.  gram_tc.cu (3xTF32 on tcgen05) is
// the tensor-core path for large n.
#include <cstdlib>
#include <cstring>

#include "common.cuh"

namespace flagg {

int gram_tc_launch(flagg_view* v, const float* center, int accumulate, double* G, cudaStream_t st);  // gram_tc.cu
int gram_small_launch(flagg_view* v, const float* center, int accumulate, double* G, cudaStream_t st);  // gram_small.cu (n <= 32)
int sqdist_small_launch(flagg_view* v, int accumulate, double* D2, cudaStream_t st);                          // sqdist_small.cu (n <= 32)
int sqdist_pairs_launch(flagg_view* v, int accumulate, double* D2, cudaStream_t st);                          // sqdist_pairs.cu (16-byte aligned rows)
bool sqdist_pairs_serves(int n);

constexpr int SQDIST_PAIRS_MIN_N = 2;   // flagg_sqdist: pair-group kernel from this cohort size (faster at every n it serves, profiles/r02_krum_small_pairs.txt)
constexpr int GT = 64;          // output tile edge
constexpr int GK = 32;          // columns per k-step (128 B of every row)
constexpr int G_THREADS = 256;  // 16 x 16 threads, 4 x 4 outputs each
constexpr int G_FLUSH_UNITS = 8;  // fold fp32 -> fp64 every 8 x 128 columns

// Stage GK columns of rows [r0, r0+64) into sm[k][row] (centred, zero padded).
__device__ __forceinline__ void gram_stage(const ViewDev& v, int p, int64_t col0, int valid, int r0,
                                           const float* __restrict__ crow, float (*sm)[GT + 1]) {
    // crow: centre values of this segment (already offset to the segment start) or NULL
    const int tid = threadIdx.x, kcol = tid & 31, rsub = tid >> 5;
    const bool ok = kcol < valid;
    const float cval = (ok && crow) ? __ldg(crow + col0 + kcol) : 0.f;
#pragma unroll
    for (int j = 0; j < GT / 8; ++j) {
        const int r = r0 + rsub + 8 * j;
        float x = 0.f;
        if (ok && r < v.n) x = ldg_stream(v.seg_ptrs[size_t(r) * v.P + p] + col0 + kcol) - cval;
        sm[kcol][rsub + 8 * j] = x;
    }
}

// grid = (tiles, splits).  Tile t -> (bi <= bj).  Split s owns 128-column units
// [s*upb, (s+1)*upb).  Writes its 64x64 FP64 partial to part[(s*tiles + t)*4096 ...].
__global__ void __launch_bounds__(G_THREADS)
gram_ffma_kernel(ViewDev v, const float* __restrict__ center, int nb, int64_t total_units, int64_t units_per_split,
                 double* __restrict__ part) {
    __shared__ float As[GK][GT + 1];
    __shared__ float Bs[GK][GT + 1];
    // decode upper-triangular tile index
    int t = blockIdx.x, bi = 0;
    while (t >= nb - bi) { t -= nb - bi; ++bi; }
    const int bj = bi + t;
    const bool diag = (bi == bj);
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const int64_t* __restrict__ ustart = v.unit_start[GRAN_128];

    double acc64[4][4];
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) { acc64[i][j] = 0.0; acc[i][j] = 0.f; }

    const int64_t u0 = int64_t(blockIdx.y) * units_per_split;
    const int64_t u1 = min(total_units, u0 + units_per_split);
    int since_flush = 0;
    for (int64_t u = u0; u < u1; ++u) {
        const int p = find_segment(ustart, v.P, u);
        const int64_t c0 = (u - ustart[p]) * 128;
        const int len = int(min((long long)(128), (long long)v.seg_start[p + 1] - v.seg_start[p] - c0));
        const float* crow = center ? center + v.seg_start[p] : nullptr;
        for (int ks = 0; ks < len; ks += GK) {
            __syncthreads();
            gram_stage(v, p, c0 + ks, len - ks, bi * GT, crow, As);
            if (!diag) gram_stage(v, p, c0 + ks, len - ks, bj * GT, crow, Bs);
            __syncthreads();
            float (*Bp)[GT + 1] = diag ? As : Bs;
#pragma unroll
            for (int kk = 0; kk < GK; ++kk) {
                float a[4], b[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) { a[i] = As[kk][ty * 4 + i]; b[i] = Bp[kk][tx * 4 + i]; }
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
            }
        }
        if (++since_flush == G_FLUSH_UNITS) {
            since_flush = 0;
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) { acc64[i][j] += double(acc[i][j]); acc[i][j] = 0.f; }
        }
    }
    double* dst = part + (size_t(blockIdx.y) * gridDim.x + blockIdx.x) * (GT * GT);
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j)
            dst[(ty * 4 + i) * GT + tx * 4 + j] = acc64[i][j] + double(acc[i][j]);
}

// Sum the split partials in split order and scatter the tile (and its mirror) into G.
__global__ void gram_reduce_kernel(const double* __restrict__ part, int tiles, int splits, int nb, int n,
                                   int accumulate, double* __restrict__ G) {
    int t = blockIdx.x, bi = 0;
    while (t >= nb - bi) { t -= nb - bi; ++bi; }
    const int bj = bi + t;
    for (int e = threadIdx.x; e < GT * GT; e += blockDim.x) {
        const int r = bi * GT + e / GT, c = bj * GT + e % GT;
        if (r >= n || c >= n) continue;
        if (bi == bj && c < r) continue;  // diagonal tile: lower half comes from the mirror
        double s = 0.0;
        for (int sp = 0; sp < splits; ++sp) s += part[(size_t(sp) * tiles + blockIdx.x) * (GT * GT) + e];
        if (accumulate) {
            G[size_t(r) * n + c] += s;
            if (r != c) G[size_t(c) * n + r] += s;
        } else {
            G[size_t(r) * n + c] = s;
            if (r != c) G[size_t(c) * n + r] = s;
        }
    }
}

// ---- K5 ------------------------------------------------------------------------------
// One CTA per client i: D_ij in shared memory, bitonic sort, sum of ranks 1..m
// (rank 0 is the self distance 0, as np.sort(row)[1:m+1] at krum.py:126-129).
// Non-finite values are kept, not clamped: a NaN or infinite d^2 (a client that sent NaN / Inf)
// gives a NaN / inf distance as torch.norm does at krum.py:95, the sort puts NaN last as np.sort
// does (krum.py:126), and the score sum keeps them -- so such a client scores NaN / inf and
// np.argsort (krum.py:174) ranks it last instead of first.
__device__ __forceinline__ bool krum_after(double a, double b) {   // a sorts after b; NaN after everything
    return (a > b) || (a != a && b == b);
}

// MODE 0: G is the centred Gram matrix; 1: G holds distances; 2: G holds SQUARED distances (flagg_sqdist).
template <int MODE>
__global__ void __launch_bounds__(256)
krum_scores_kernel(const double* __restrict__ G, int n, int npad, int m, double* __restrict__ D,
                   double* __restrict__ scores) {
    extern __shared__ double srow[];  // [npad]
    __shared__ double red[256];
    const int i = blockIdx.x, tid = threadIdx.x;
    const double gii = MODE != 0 ? 0.0 : G[size_t(i) * n + i];
    for (int j = tid; j < npad; j += blockDim.x) {
        double dv = __longlong_as_double(0x7ff8000000000000LL);  // NaN padding: sorts last, after real infinities
        if (j < n) {
            if (MODE == 1) {
                dv = G[size_t(i) * n + j];
            } else {
                const double d2 = MODE == 2 ? G[size_t(i) * n + j] : gii + G[size_t(j) * n + j] - 2.0 * G[size_t(i) * n + j];
                dv = (j == i) ? 0.0 : ((d2 != d2) ? d2 : sqrt(fmax(d2, 0.0)));
                if (D) D[size_t(i) * n + j] = dv;
            }
        }
        srow[j] = dv;
    }
    if (!scores) return;   // distances only (CTA-uniform)
    __syncthreads();
    for (int k = 2; k <= npad; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int x = tid; x < npad; x += blockDim.x) {
                const int y = x ^ j;
                if (y > x) {
                    const double a = srow[x], b = srow[y];
                    const bool up = ((x & k) == 0);
                    if (krum_after(a, b) == up) { srow[x] = b; srow[y] = a; }
                }
            }
            __syncthreads();
        }
    }
    double s = 0.0;
    for (int j = 1 + tid; j <= m; j += blockDim.x) s += srow[j];
    red[tid] = s;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (tid < o) red[tid] += red[tid + o];
        __syncthreads();
    }
    if (tid == 0) scores[i] = red[0];
}

}  // namespace flagg

using namespace flagg;

extern "C" {

int flagg_gram(const flagg_view* cv, const float* center, int accumulate, double* G, void* stream) {
    FLAGG_REQUIRE(cv && G, "flagg_gram: NULL argument");
    flagg_view* v = const_cast<flagg_view*>(cv);
    v->last_stream = stream;
    const int n = v->dev.n;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    {   // Kernel choice (FLAGG_GRAM_IMPL=small|ffma|tc|ss|ts1|ts2 overrides, see gram_tc_launch):
        //   n <= 32            register-tile kernel (gram_small.cu), any alignment
        //   n >= 33, aligned   tensor cores (gram_tc.cu: single CTAs up to n = 128, CTA pairs above)
        //   otherwise          64x64 FFMA tiles (below)
        const char* impl = getenv("FLAGG_GRAM_IMPL");
        auto is = [&](const char* name) { return impl && !strcmp(impl, name); };
        const bool force_tc = is("tc") || is("ss") || is("ts1") || is("ts2"), force_ffma = is("ffma"), force_small = is("small");
        if (force_tc && v->align_bytes < 16) {
            set_error("flagg_gram: FLAGG_GRAM_IMPL=tc needs 16-byte aligned tensors");
            return FLAGG_EUNSUPPORTED;
        }
        if (force_small && n > 32) {   // GS_MAX_N
            set_error("flagg_gram: FLAGG_GRAM_IMPL=small needs n <= 32");
            return FLAGG_EUNSUPPORTED;
        }
        if (v->total_units[GRAN_128] > 0 && !force_ffma && !force_tc && (n <= 32 || force_small))
            return gram_small_launch(v, center, accumulate, G, st);
        if (v->total_units[GRAN_128] > 0 && v->align_bytes >= 16 && !force_ffma && (n > 32 || force_tc))
            return gram_tc_launch(v, center, accumulate, G, st);
    }
    const int nb = (n + GT - 1) / GT;
    const int tiles = nb * (nb + 1) / 2;
    const int64_t units = v->total_units[GRAN_128];
    if (units == 0) {
        if (!accumulate) FLAGG_CUDA(cudaMemsetAsync(G, 0, sizeof(double) * size_t(n) * n, st));
        return FLAGG_OK;
    }
    const int sms = device_sm_count();
    FLAGG_REQUIRE(sms > 0, "cannot query the SM count of the current device");
    int64_t splits = std::max<int64_t>(1, (int64_t(sms) * 4 + tiles - 1) / tiles);
    splits = std::min<int64_t>(splits, units);
    splits = std::min<int64_t>(splits, 65535);
    const int64_t upb = (units + splits - 1) / splits;
    splits = (units + upb - 1) / upb;
    const size_t need = sizeof(double) * size_t(splits) * tiles * GT * GT;
    int rc = v->ensure_scratch(need);
    if (rc != FLAGG_OK) return rc;
    double* part = static_cast<double*>(v->d_scratch);
    dim3 grid(tiles, unsigned(splits));
    gram_ffma_kernel<<<grid, G_THREADS, 0, st>>>(v->dev, center, nb, units, upb, part);
    FLAGG_CUDA(cudaGetLastError());
    gram_reduce_kernel<<<tiles, 256, 0, st>>>(part, tiles, int(splits), nb, n, accumulate, G);
    FLAGG_CUDA(cudaGetLastError());
    return FLAGG_OK;
}

static int krum_scores_launch(const double* G, int n, int f, double* D, double* scores, int mode, void* stream) {
    const int m = n - f - 2;
    FLAGG_REQUIRE(n >= 1 && n <= 4096, "flagg_krum_scores: n=%d unsupported (1..4096)", n);
    FLAGG_REQUIRE(!scores || (m >= 1 && m <= n - 1), "flagg_krum_scores: need 1 <= n-f-2 <= n-1 (n=%d f=%d)", n, f);
    int npad = 2;
    while (npad < n) npad <<= 1;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (mode == 1) krum_scores_kernel<1><<<n, 256, sizeof(double) * npad, st>>>(G, n, npad, m, nullptr, scores);
    else if (mode == 2) krum_scores_kernel<2><<<n, 256, sizeof(double) * npad, st>>>(G, n, npad, m, D, scores);
    else krum_scores_kernel<0><<<n, 256, sizeof(double) * npad, st>>>(G, n, npad, m, D, scores);
    FLAGG_CUDA(cudaGetLastError());
    return FLAGG_OK;
}

int flagg_krum_scores(const double* G, int n, int f, double* D, double* scores, void* stream) {
    FLAGG_REQUIRE(G && (scores || D), "flagg_krum_scores: NULL argument");
    return krum_scores_launch(G, n, f, D, scores, 0, stream);
}

int flagg_sqdist(const flagg_view* cv, int accumulate, double* D2, void* stream) {
    FLAGG_REQUIRE(cv && D2, "flagg_sqdist: NULL argument");
    flagg_view* v = const_cast<flagg_view*>(cv);
    v->last_stream = stream;
    if (v->dev.n > 32) {   // GS_MAX_N
        set_error("flagg_sqdist: serves cohorts of up to 32 clients (n=%d); larger cohorts use flagg_gram", v->dev.n);
        return FLAGG_EUNSUPPORTED;
    }
    // Two kernels, the same arithmetic per pair and column: pair groups (sqdist_pairs.cu: every warp the same work, rows
    // by 512-byte bulk copies) for 16-byte aligned rows from SQDIST_PAIRS_MIN_N clients, row-block jobs (sqdist_small.cu)
    // otherwise.  FLAGG_SQDIST_IMPL=pairs|jobs forces one (tests, tuning).
    const int64_t units = v->total_units[GRAN_128];
    const bool pairs_ok = sqdist_pairs_serves(v->dev.n) && v->align_bytes >= 16 && units >= 1 && units < (int64_t(1) << 24);
    const char* impl = getenv("FLAGG_SQDIST_IMPL");
    bool pairs = pairs_ok && v->dev.n >= SQDIST_PAIRS_MIN_N;
    if (impl && *impl) {
        if (!strcmp(impl, "pairs")) {
            if (!pairs_ok) {
                set_error("FLAGG_SQDIST_IMPL=pairs: needs 2 <= n <= 32, 16-byte aligned rows and fewer than 2^31 columns");
                return FLAGG_EUNSUPPORTED;
            }
            pairs = true;
        } else if (!strcmp(impl, "jobs")) {
            pairs = false;
        } else {
            set_error("FLAGG_SQDIST_IMPL=%s: expected pairs or jobs", impl);
            return FLAGG_EINVAL;
        }
    }
    return pairs ? sqdist_pairs_launch(v, accumulate, D2, static_cast<cudaStream_t>(stream))
                 : sqdist_small_launch(v, accumulate, D2, static_cast<cudaStream_t>(stream));
}

int flagg_krum_scores_from_sqdist(const double* D2, int n, int f, double* D, double* scores, void* stream) {
    FLAGG_REQUIRE(D2 && (scores || D), "flagg_krum_scores_from_sqdist: NULL argument");
    return krum_scores_launch(D2, n, f, D, scores, 2, stream);
}

int flagg_krum_scores_from_distances(const double* D, int n, int f, double* scores, void* stream) {
    FLAGG_REQUIRE(D && scores, "flagg_krum_scores_from_distances: NULL argument");
    return krum_scores_launch(D, n, f, nullptr, scores, 1, stream);
}

}  // extern "C"
