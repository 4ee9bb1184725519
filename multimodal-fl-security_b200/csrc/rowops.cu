// K2: streaming row reductions (||x_i||^2, x_i . g0, max|x_i|, ||g0||^2) in ONE pass
//     over the client-update matrix -- replaces the per-client torch.cat + torch.norm
//     + torch.dot calls at fltrust.py:170-176,205-209 and
//     differential_privacy.py:87-88,225-230,311.
// K3: weighted row combination with the reference's exact fp32 operation order --
//     replaces Python sum() over scalar*tensor at base_defense.py:91-95,
//     krum.py:187-190, fltrust.py:262-268, differential_privacy.py:148-155,262-269,325-332.
// Both are HBM-bound streaming kernels: 128-bit coalesced loads, no shared-memory
// staging of X (no reuse), grid sized from the SM count.
#include <algorithm>
#include <mutex>

#include "common.cuh"

namespace flagg {

constexpr int RR_THREADS = 256;
constexpr int RR_WARPS = RR_THREADS / 32;
constexpr int RR_CHUNK = 2048;  // == unit_cols(GRAN_2K): 8 KB of each row per unit

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// Persistent CTAs; CTA b walks units b, b+grid, ...  Within a unit warp w owns rows
// w, w+8, ...: each lane streams 16 x float4 of the row (8 KB per warp-row, all loads
// independent), multiplies in fp32, accumulates float4-partials in FP64, and the warp
// total lands in the CTA's shared per-row accumulator (one owner warp per row: no
// atomics).  Per-CTA partials go to `part`; row_reduce_finalize adds them in CTA order,
// so the result is deterministic.
template <bool VEC4, bool HAS_G0, bool CENTERED>
__global__ void __launch_bounds__(RR_THREADS)
row_reduce_kernel(ViewDev v, const float* __restrict__ g0, double* __restrict__ part_ss,
                  double* __restrict__ part_dot, float* __restrict__ part_linf,
                  double* __restrict__ part_g0, int64_t total_units) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int n = v.n;
    double* acc_ss = reinterpret_cast<double*>(smem_raw);
    double* acc_dot = acc_ss + n;
    float* g0s = reinterpret_cast<float*>(acc_dot + n);   // [RR_CHUNK], 16 B aligned (n*16 bytes before)
    float* acc_inf = g0s + RR_CHUNK;
    __shared__ double red[RR_WARPS];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < n; i += RR_THREADS) { acc_ss[i] = 0.0; acc_dot[i] = 0.0; acc_inf[i] = 0.f; }
    double g0ss = 0.0;
    const int64_t* __restrict__ ustart = v.unit_start[GRAN_2K];

    for (int64_t u = blockIdx.x; u < total_units; u += gridDim.x) {
        const int p = find_segment(ustart, v.P, u);
        const int64_t c0 = (u - ustart[p]) * RR_CHUNK;
        const int64_t segn = v.seg_start[p + 1] - v.seg_start[p];
        const int len = int(min((long long)(RR_CHUNK), (long long)segn - c0));
        __syncthreads();  // previous unit's readers of g0s are done (also orders the init above)
        if (HAS_G0) {
            const float* gsrc = g0 ? g0 + v.seg_start[p] + c0 : nullptr;       // g0 == nullptr: zeros are staged (see the launch)
            for (int j = tid; j < RR_CHUNK; j += RR_THREADS) {
                float g = (gsrc && j < len) ? __ldg(gsrc + j) : 0.f;
                g0s[j] = g;
                g0ss += double(g * g);
            }
            __syncthreads();
        }
        const int nv = VEC4 ? (len >> 2) : 0;
        for (int i = warp; i < n; i += RR_WARPS) {
            const float* row = v.seg_ptrs[size_t(i) * v.P + p] + c0;
            double ss = 0.0, dt = 0.0;
            float mx = 0.f;
            if (VEC4) {
                const float4* row4 = reinterpret_cast<const float4*>(row);
                const float4* g4 = reinterpret_cast<const float4*>(g0s);
                float4 x[16];
#pragma unroll
                for (int jj = 0; jj < 16; ++jj) {
                    const int j = lane + 32 * jj;
                    x[jj] = (j < nv) ? ldg_stream4(row4 + j) : make_float4(0.f, 0.f, 0.f, 0.f);
                }
#pragma unroll
                for (int jj = 0; jj < 16; ++jj) {
                    const int j = lane + 32 * jj;
                    float4 a = x[jj];
                    const float4 g = HAS_G0 ? g4[j & (RR_CHUNK / 4 - 1)] : make_float4(0.f, 0.f, 0.f, 0.f);
                    if (CENTERED && j < nv) { a.x -= g.x; a.y -= g.y; a.z -= g.z; a.w -= g.w; }
                    ss += double((a.x * a.x + a.y * a.y) + (a.z * a.z + a.w * a.w));
                    mx = fmaxf(fmaxf(mx, fmaxf(fabsf(a.x), fabsf(a.y))), fmaxf(fabsf(a.z), fabsf(a.w)));
                    if (HAS_G0) {
                        dt += double((a.x * g.x + a.y * g.y) + (a.z * g.z + a.w * g.w));
                    }
                }
            }
            for (int j = (nv << 2) + lane; j < len; j += 32) {   // tail / unaligned path
                float a = ldg_stream(row + j);
                if (CENTERED) a -= g0s[j];
                ss += double(a * a);
                mx = fmaxf(mx, fabsf(a));
                if (HAS_G0) dt += double(a * g0s[j]);
            }
            ss = warp_sum(ss);
            mx = warp_max(mx);
            if (HAS_G0) dt = warp_sum(dt);
            if (lane == 0) {
                acc_ss[i] += ss;
                acc_inf[i] = fmaxf(acc_inf[i], mx);
                if (HAS_G0) acc_dot[i] += dt;
            }
        }
    }
    __syncthreads();
    for (int i = tid; i < n; i += RR_THREADS) {
        part_ss[size_t(blockIdx.x) * n + i] = acc_ss[i];
        part_dot[size_t(blockIdx.x) * n + i] = acc_dot[i];
        part_linf[size_t(blockIdx.x) * n + i] = acc_inf[i];
    }
    g0ss = warp_sum(g0ss);
    if (lane == 0) red[warp] = g0ss;
    __syncthreads();
    if (tid == 0) {
        double t = 0.0;
        for (int w = 0; w < RR_WARPS; ++w) t += red[w];
        part_g0[blockIdx.x] = t;
    }
}

// One warp per client (plus one for ||g0||^2): lanes stride over the per-CTA partials, then a fixed-shape shuffle tree
// -- the same order every run.  (One THREAD per client walked up to 592 dependent L2 round trips: 50-100 us, a
// quarter of a whole clipping round at the small BASELINE shapes.)
constexpr int RRF_THREADS = 128;
__global__ void __launch_bounds__(RRF_THREADS)
row_reduce_finalize(const double* __restrict__ part_ss, const double* __restrict__ part_dot,
                    const float* __restrict__ part_linf, const double* __restrict__ part_g0,
                    int nblocks, int n, double* sumsq, double* dot, float* linf, double* g0_sumsq) {
    const int lane = threadIdx.x & 31;
    const int i = blockIdx.x * (RRF_THREADS / 32) + (threadIdx.x >> 5);      // warp-uniform
    if (i > n) return;
    if (i == n) {
        if (!g0_sumsq) return;
        double t = 0.0;
        for (int b = lane; b < nblocks; b += 32) t += part_g0[b];
        t = warp_sum(t);
        if (lane == 0) *g0_sumsq = t;
        return;
    }
    double ss = 0.0, dt = 0.0;
    float mx = 0.f;
#pragma unroll 4
    for (int b = lane; b < nblocks; b += 32) {
        ss += part_ss[size_t(b) * n + i];
        dt += part_dot[size_t(b) * n + i];
        mx = fmaxf(mx, part_linf[size_t(b) * n + i]);
    }
    ss = warp_sum(ss);
    dt = warp_sum(dt);
    mx = warp_max(mx);
    if (lane == 0) {
        if (sumsq) sumsq[i] = ss;
        if (dot) dot[i] = dt;
        // torch.max(torch.abs(x)) propagates NaN (differential_privacy.py:230); fmaxf does not.
        if (linf) linf[i] = (ss != ss) ? __int_as_float(0x7fc00000) : mx;
    }
}

// ---- K3 ----------------------------------------------------------------------------
constexpr int RC_THREADS = 256;
constexpr int RC_UNROLL = 8;

// One CTA per 1024-column unit; thread owns 4 columns.  The m selected rows are visited
// strictly in order; loads of RC_UNROLL rows are issued together, then folded in with
// separately rounded multiply / multiply / add (no FMA contraction) so the result is
// bit-identical to the reference's left-to-right Python sum().
template <bool VEC4>
__global__ void __launch_bounds__(RC_THREADS)
row_combine_kernel(ViewDev v, int m, const int* __restrict__ rows, const float* __restrict__ w1,
                   const float* __restrict__ w2, float divisor, const float* __restrict__ divisor_dev,
                   float* __restrict__ out, const float* __restrict__ prev, double* __restrict__ part_delta) {
    if (divisor_dev) divisor = __ldg(divisor_dev);
    const int64_t* __restrict__ ustart = v.unit_start[GRAN_1K];
    const int64_t u = blockIdx.x;
    const int p = find_segment(ustart, v.P, u);
    const int64_t c0 = (u - ustart[p]) * 1024;
    const int64_t segn = v.seg_start[p + 1] - v.seg_start[p];
    const int len = int(min((long long)(1024), (long long)segn - c0));
    const int t = threadIdx.x;
    int col[4];
    bool vec = false;
    if (VEC4 && 4 * t + 3 < len) {
        vec = true;
#pragma unroll
        for (int q = 0; q < 4; ++q) col[q] = 4 * t + q;
    } else if (VEC4) {
        // ragged tail of a vectorised unit: the few leftover columns go to the threads
        // past the last full float4, one column each
        const int full = len >> 2;
#pragma unroll
        for (int q = 0; q < 4; ++q) col[q] = (q == 0 && t >= full && 4 * full + (t - full) < len) ? 4 * full + (t - full) : -1;
    } else {
#pragma unroll
        for (int q = 0; q < 4; ++q) col[q] = (t + RC_THREADS * q < len) ? t + RC_THREADS * q : -1;
    }
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    for (int j0 = 0; j0 < m; j0 += RC_UNROLL) {
        float x[RC_UNROLL][4];
        float a1[RC_UNROLL], a2[RC_UNROLL];
        int ri[RC_UNROLL];
#pragma unroll
        for (int jj = 0; jj < RC_UNROLL; ++jj) {
            const int j = j0 + jj;
            ri[jj] = (j < m) ? (rows ? __ldg(rows + j) : j) : -1;      // a negative entry is skipped (device-side compaction pads with -1)
            if (ri[jj] >= 0) {
                const float* row = v.seg_ptrs[size_t(ri[jj]) * v.P + p] + c0;
                a1[jj] = w1 ? __ldg(w1 + j) : 1.0f;
                a2[jj] = w2 ? __ldg(w2 + j) : 1.0f;
                if (vec) {
                    const float4 q4 = ldg_stream4(reinterpret_cast<const float4*>(row) + t);
                    x[jj][0] = q4.x; x[jj][1] = q4.y; x[jj][2] = q4.z; x[jj][3] = q4.w;
                } else {
#pragma unroll
                    for (int q = 0; q < 4; ++q) x[jj][q] = (col[q] >= 0) ? ldg_stream(row + col[q]) : 0.f;
                }
            }
        }
#pragma unroll
        for (int jj = 0; jj < RC_UNROLL; ++jj) {
            if (ri[jj] >= 0) {
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    acc[q] = __fadd_rn(acc[q], __fmul_rn(a2[jj], __fmul_rn(a1[jj], x[jj][q])));
            }
        }
    }
    float r[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) r[q] = __fdiv_rn(acc[q], divisor);
    // the epilogue form writes straight into the caller's P parameter tensors (ViewDev::out_ptrs)
    float* o = v.out_ptrs ? v.out_ptrs[p] + c0 : out + v.seg_start[p] + c0;
    if (vec && ((reinterpret_cast<uintptr_t>(o) & 15) == 0)) {
        reinterpret_cast<float4*>(o)[t] = make_float4(r[0], r[1], r[2], r[3]);
    } else {
#pragma unroll
        for (int q = 0; q < 4; ++q)
            if (col[q] >= 0) o[col[q]] = r[q];
    }
    if (part_delta) {   // ||out - prev||^2 of this unit (Weiszfeld stopping rule, trimmed_mean.py:255)
        const float* pv = prev + v.seg_start[p] + c0;
        double dl = 0.0;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int cc = vec ? 4 * t + q : col[q];
            if (cc >= 0) {
                const float df = __fsub_rn(r[q], __ldg(pv + cc));
                dl += double(__fmul_rn(df, df));
            }
        }
        dl = warp_sum(dl);
        __shared__ double red[RC_THREADS / 32];
        if ((t & 31) == 0) red[t >> 5] = dl;
        __syncthreads();
        if (t == 0) {
            double tot = 0.0;
#pragma unroll
            for (int w = 0; w < RC_THREADS / 32; ++w) tot += red[w];
            part_delta[blockIdx.x] = tot;
        }
    }
}

// Sum per-CTA partials in CTA order (deterministic).
__global__ void __launch_bounds__(256) sum_partials_kernel(const double* __restrict__ part, int64_t count, double* __restrict__ out) {
    __shared__ double red[256];
    double s = 0.0;
    for (int64_t i = threadIdx.x; i < count; i += 256) s += part[i];
    red[threadIdx.x] = s;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) *out = red[0];
}

// Weiszfeld weights (trimmed_mean.py:245-251): dist_i = clamp(fp32 ||x_i - z||, 1e-10), w_i = 1/dist_i,
// wsum = sum_i w_i.  One warp; n is a client count.
__global__ void __launch_bounds__(32) weiszfeld_weights_kernel(const double* __restrict__ sumsq, int n, float* __restrict__ w,
                                                               float* __restrict__ wsum) {
    double s = 0.0;
    for (int i = threadIdx.x; i < n; i += 32) {
        const float dist = fmaxf(float(sqrt(sumsq[i])), 1e-10f);
        const float wi = __fdiv_rn(1.0f, sumsq[i] != sumsq[i] ? __int_as_float(0x7fc00000) : dist);
        w[i] = wi;
        s += double(wi);
    }
    s = warp_sum(s);
    if (threadIdx.x == 0) *wsum = float(s);
}

// Per-round weights on the device (no host round trip between K2 and K3): the fp32 / double arithmetic of the
// reference's Python, operation for operation, on the K2 statistics.  One CTA: thread i forms client i's terms,
// thread 0 then walks the clients in order for the double-precision running sums Python's sum() forms and for the
// order-preserving compaction of the rows K3 has to read (the rest of `rows` is filled with -1 = skip).
//   info[i]  : FLTRUST trust score (fltrust.py:186), otherwise the client's norm (differential_privacy.py:227-232)
//   counts[0]: rows kept (m)   counts[1]: FLTRUST: 1 if sum(trust) < 1e-10 (:258-260, caller returns g0);
//              CLIP / DPSGD: clients clipped;  NORMBOUND: 1 if every client was rejected (:319-322, all are used)
__global__ void __launch_bounds__(256)
round_weights_kernel(int rule, int n, const double* __restrict__ sumsq, const double* __restrict__ dot,
                     const double* __restrict__ g0_sumsq, const float* __restrict__ linf, const double* __restrict__ num_examples,
                     double p0, double p1, int use_linf, int* __restrict__ rows, float* __restrict__ w1, float* __restrict__ w2,
                     float* __restrict__ divisor, double* __restrict__ info, int* __restrict__ counts, unsigned char* __restrict__ keep) {
    const float nanf_ = __int_as_float(0x7fc00000);
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const double ss = sumsq ? sumsq[i] : 0.0;
        const float norm32 = use_linf ? linf[i] : ((ss != ss) ? nanf_ : float(sqrt(ss)));   // torch.norm / max|x| as an fp32 value
        const double norm = double(norm32);
        float a1 = 1.0f, a2 = num_examples ? float(num_examples[i]) : 1.0f;
        unsigned char k = 1;
        double inf = norm;
        if (rule == FLAGG_W_FLTRUST) {
            const float sn = float(sqrt(*g0_sumsq));
            const bool tiny = norm32 < 1e-10f;
            const float q = __fdiv_rn(float(dot[i]), __fmul_rn(norm32, sn));
            const float trust = (!tiny && sn >= 1e-10f && q > 0.f) ? q : 0.f;          // max(0.0, q): NaN -> 0.0
            a1 = tiny ? 1.0f : __fdiv_rn(sn, norm32);                                  // fltrust.py:208-214
            a2 = trust;
            inf = double(trust);
            // a zero-trust row adds +-0 to a sum that starts at +0: skipped unless it holds Inf / NaN (0 * Inf = NaN)
            k = (trust != 0.f || !(fabs(ss) <= 1.79769313486231570e308)) ? 1 : 0;
        } else if (rule == FLAGG_W_CLIP) {
            if (norm > p0) a1 = float(p0 / norm);                                      // Python float division, then fp32
        } else if (rule == FLAGG_W_DPSGD) {
            const float clip32 = float(p0);
            if (norm32 > clip32) a1 = __fmul_rn(__fdiv_rn(1.0f, norm32), clip32);      // norm.reciprocal() * clip
        } else {
            k = (p1 <= norm && norm <= p0) ? 1 : 0;                                    // min_norm <= norm <= max_norm
        }
        w1[i] = a1;
        w2[i] = a2;
        info[i] = inf;
        keep[i] = k;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int m = 0, flag = 0;
        double total = 0.0;
        if (rule == FLAGG_W_FLTRUST) {
            for (int i = 0; i < n; ++i) total += info[i];                              // Python sum(): left to right, double
            flag = (total < 1e-10) ? 1 : 0;
        } else if (rule == FLAGG_W_NORMBOUND) {
            int kept = 0;
            for (int i = 0; i < n; ++i) kept += keep[i];
            if (kept == 0) { flag = 1; for (int i = 0; i < n; ++i) keep[i] = 1; }
        } else {
            for (int i = 0; i < n; ++i) flag += (rule == FLAGG_W_CLIP) ? (info[i] > p0) : (float(info[i]) > float(p0));
        }
        if (rule != FLAGG_W_FLTRUST)
            for (int i = 0; i < n; ++i) if (keep[i]) total += num_examples ? num_examples[i] : 1.0;   // integer counts in FP64: exact
        for (int i = 0; i < n; ++i) {
            if (keep[i]) {
                rows[m] = i;
                w1[m] = w1[i];      // m <= i: compaction in place never overwrites an unread entry
                w2[m] = w2[i];
                ++m;
            }
        }
        for (int j = m; j < n; ++j) rows[j] = -1;
        *divisor = float(total);
        counts[0] = m;
        counts[1] = flag;
    }
}

}  // namespace flagg

using namespace flagg;

extern "C" {

int flagg_round_weights(int rule, int n, const double* sumsq, const double* dot, const double* g0_sumsq, const float* linf,
                        const double* num_examples, double p0, double p1, int32_t* rows, float* w1, float* w2, float* divisor,
                        double* info, int32_t* counts, void* scratch_n_bytes, void* stream) {
    FLAGG_REQUIRE(rule >= FLAGG_W_FLTRUST && rule <= FLAGG_W_NORMBOUND, "flagg_round_weights: bad rule %d", rule);
    FLAGG_REQUIRE(n >= 1 && rows && w1 && w2 && divisor && info && counts && scratch_n_bytes, "flagg_round_weights: NULL argument");
    FLAGG_REQUIRE(sumsq || linf, "flagg_round_weights: no statistics");
    FLAGG_REQUIRE(rule != FLAGG_W_FLTRUST || (sumsq && dot && g0_sumsq), "flagg_round_weights: FLTrust needs sumsq, dot and g0_sumsq");
    const int use_linf = (rule == FLAGG_W_CLIP && linf && !sumsq) ? 1 : 0;
    FLAGG_REQUIRE(use_linf || sumsq, "flagg_round_weights: this rule needs sumsq");
    round_weights_kernel<<<1, 256, 0, static_cast<cudaStream_t>(stream)>>>(rule, n, sumsq, dot, g0_sumsq, linf, num_examples, p0, p1,
                                                                            use_linf, rows, w1, w2, divisor, info, counts,
                                                                            static_cast<unsigned char*>(scratch_n_bytes));
    FLAGG_CUDA(cudaGetLastError());
    return FLAGG_OK;
}

int flagg_row_reduce(const flagg_view* cv, const float* g0, int centered, double* sumsq, double* dot, float* linf,
                     double* g0_sumsq, void* stream) {
    FLAGG_REQUIRE(cv, "flagg_row_reduce: NULL view");
    if (cv->dev.d == 0) {   // an empty column shard: the rank's partial statistics are zeros (g0 is an empty tensor, i.e. NULL)
        const int n0 = cv->dev.n;
        cudaStream_t s0 = static_cast<cudaStream_t>(stream);
        const_cast<flagg_view*>(cv)->last_stream = stream;
        if (sumsq) FLAGG_CUDA(cudaMemsetAsync(sumsq, 0, sizeof(double) * n0, s0));
        if (dot) FLAGG_CUDA(cudaMemsetAsync(dot, 0, sizeof(double) * n0, s0));
        if (linf) FLAGG_CUDA(cudaMemsetAsync(linf, 0, sizeof(float) * n0, s0));
        if (g0_sumsq) FLAGG_CUDA(cudaMemsetAsync(g0_sumsq, 0, sizeof(double), s0));
        return FLAGG_OK;
    }
    FLAGG_REQUIRE(!(dot && !g0), "flagg_row_reduce: dot requested without g0");
    FLAGG_REQUIRE(!(g0_sumsq && !g0), "flagg_row_reduce: g0_sumsq requested without g0");
    FLAGG_REQUIRE(!(centered && !g0), "flagg_row_reduce: centered statistics need g0");
    flagg_view* v = const_cast<flagg_view*>(cv);
    v->last_stream = stream;
    const int n = v->dev.n;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int64_t units = v->total_units[GRAN_2K];
    const size_t smem = size_t(n) * 16 + RR_CHUNK * 4 + size_t(n) * 4;
    FLAGG_REQUIRE(smem <= 200 * 1024, "flagg_row_reduce: n=%d too large for the per-CTA accumulators", n);
    const int sms = device_sm_count();
    FLAGG_REQUIRE(sms > 0, "cannot query the SM count of the current device");
    int nblocks = int(std::min<int64_t>(std::max<int64_t>(units, 1), int64_t(sms) * 4));
    const size_t b_d = sizeof(double) * size_t(nblocks) * n;
    const size_t b_f = sizeof(float) * size_t(nblocks) * n;
    const size_t need = 2 * b_d + ((b_f + 7) & ~size_t(7)) + sizeof(double) * nblocks;
    int rc = v->ensure_scratch(need);
    if (rc != FLAGG_OK) return rc;
    char* s = static_cast<char*>(v->d_scratch);
    double* part_ss = reinterpret_cast<double*>(s);
    double* part_dot = reinterpret_cast<double*>(s + b_d);
    double* part_g0 = reinterpret_cast<double*>(s + 2 * b_d);
    float* part_linf = reinterpret_cast<float*>(s + 2 * b_d + sizeof(double) * nblocks);
    const bool vec4 = v->align_bytes >= 16;
    auto launch = [&](auto kern) -> int {
        FLAGG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
        kern<<<nblocks, RR_THREADS, smem, st>>>(v->dev, g0, part_ss, part_dot, part_linf, part_g0, units);
        FLAGG_CUDA(cudaGetLastError());
        return FLAGG_OK;
    };
    // Without g0 the vectorised kernel still runs its "with g0" form on staged zeros: the shared-memory reads between
    // the load loop and the arithmetic keep the sixteen 128-bit loads per thread in flight (64 registers); the form
    // compiled without them lets ptxas sink every load to its use (32 registers, ~4 in flight) and ran at 4.3 TB/s
    // against 6.4 (profiles/r02_c5_breakdown.txt).  The wasted dot products are free under the HBM bound.
    if (vec4 && g0 && centered) rc = launch(row_reduce_kernel<true, true, true>);
    else if (g0 && centered) rc = launch(row_reduce_kernel<false, true, true>);
    else if (vec4) rc = launch(row_reduce_kernel<true, true, false>);
    else if (g0) rc = launch(row_reduce_kernel<false, true, false>);
    else rc = launch(row_reduce_kernel<false, false, false>);
    if (rc != FLAGG_OK) return rc;
    row_reduce_finalize<<<(n + 1 + RRF_THREADS / 32 - 1) / (RRF_THREADS / 32), RRF_THREADS, 0, st>>>(part_ss, part_dot, part_linf, part_g0, nblocks, n,
                                                         sumsq, dot, linf, g0_sumsq);
    FLAGG_CUDA(cudaGetLastError());
    return FLAGG_OK;
}

static int row_combine_launch(flagg_view* v, int m, const int* d_rows, const float* d_w1, const float* d_w2, float divisor,
                              const float* d_divisor, float* out, const float* prev, double* delta_sumsq, size_t scratch_off,
                              cudaStream_t st) {
    const int64_t units = v->total_units[GRAN_1K];
    if (units == 0) {
        if (delta_sumsq) FLAGG_CUDA(cudaMemsetAsync(delta_sumsq, 0, sizeof(double), st));
        return FLAGG_OK;
    }
    FLAGG_REQUIRE(units < (int64_t(1) << 31), "flagg_row_combine: too many column units");
    double* part = nullptr;
    if (delta_sumsq) {
        scratch_off = (scratch_off + 63) & ~size_t(63);
        // ensure_scratch may move the block: callers pass offsets, never pointers into it
        FLAGG_REQUIRE(v->scratch_bytes >= scratch_off + sizeof(double) * size_t(units), "flagg_row_combine: internal scratch sizing");
        part = reinterpret_cast<double*>(static_cast<char*>(v->d_scratch) + scratch_off);
    }
    if (v->align_bytes >= 16)
        row_combine_kernel<true><<<unsigned(units), RC_THREADS, 0, st>>>(v->dev, m, d_rows, d_w1, d_w2, divisor, d_divisor, out, prev, part);
    else
        row_combine_kernel<false><<<unsigned(units), RC_THREADS, 0, st>>>(v->dev, m, d_rows, d_w1, d_w2, divisor, d_divisor, out, prev, part);
    FLAGG_CUDA(cudaGetLastError());
    if (delta_sumsq) {
        sum_partials_kernel<<<1, 256, 0, st>>>(part, units, delta_sumsq);
        FLAGG_CUDA(cudaGetLastError());
    }
    return FLAGG_OK;
}

int flagg_row_combine(const flagg_view* cv, int m, const int32_t* rows, const float* w1, const float* w2,
                      float divisor, float* out, void* stream) {
    FLAGG_REQUIRE(cv, "flagg_row_combine: NULL argument");
    flagg_view* v = const_cast<flagg_view*>(cv);
    if (v->dev.d == 0) { v->last_stream = stream; return FLAGG_OK; }   // empty column shard: nothing to write
    FLAGG_REQUIRE(out || v->dev.out_ptrs, "flagg_row_combine: NULL output (and no flagg_view_set_outputs table)");
    v->last_stream = stream;
    const int n = v->dev.n;
    FLAGG_REQUIRE(m >= 1, "flagg_row_combine: m must be >= 1 (got %d)", m);
    FLAGG_REQUIRE(rows || m == n, "flagg_row_combine: rows == NULL requires m == n");
    // one host block, one upload: w1 [m] | w2 [m] | rows [m]
    std::vector<float> hw(3 * size_t(m));
    int32_t* hr = reinterpret_cast<int32_t*>(hw.data() + 2 * size_t(m));
    for (int j = 0; j < m; ++j) {
        hr[j] = rows ? rows[j] : j;
        FLAGG_REQUIRE(hr[j] >= 0 && hr[j] < n, "flagg_row_combine: rows[%d]=%d out of range", j, hr[j]);
        hw[j] = w1 ? w1[j] : 1.0f;
        hw[m + j] = w2 ? w2[j] : 1.0f;
    }
    int rc = v->ensure_scratch(sizeof(float) * 3 * size_t(m) + 64);
    if (rc != FLAGG_OK) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    float* d_w = static_cast<float*>(v->d_scratch);
    int* d_rows = reinterpret_cast<int*>(d_w + 2 * size_t(m));
    FLAGG_CUDA(cudaMemcpyAsync(d_w, hw.data(), sizeof(float) * 3 * size_t(m), cudaMemcpyHostToDevice, st));
    return row_combine_launch(v, m, d_rows, d_w, d_w + m, divisor, nullptr, out, nullptr, nullptr, 0, st);
}

int flagg_row_combine_dev(const flagg_view* cv, int m, const int32_t* d_rows, const float* d_w1, const float* d_w2,
                          const float* d_divisor, float* out, const float* prev, double* delta_sumsq, void* stream) {
    FLAGG_REQUIRE(cv && d_divisor, "flagg_row_combine_dev: NULL argument");
    flagg_view* v = const_cast<flagg_view*>(cv);
    if (v->dev.d == 0) {   // empty column shard: nothing to write; its share of ||out - prev||^2 is 0
        v->last_stream = stream;
        if (delta_sumsq) FLAGG_CUDA(cudaMemsetAsync(delta_sumsq, 0, sizeof(double), static_cast<cudaStream_t>(stream)));
        return FLAGG_OK;
    }
    FLAGG_REQUIRE(out || v->dev.out_ptrs, "flagg_row_combine_dev: NULL output (and no flagg_view_set_outputs table)");
    FLAGG_REQUIRE(m >= 1 && (d_rows || m == v->dev.n), "flagg_row_combine_dev: need m >= 1, and m == n when rows == NULL");
    FLAGG_REQUIRE(!delta_sumsq || prev, "flagg_row_combine_dev: delta_sumsq needs prev");
    v->last_stream = stream;
    if (delta_sumsq) {
        int rc = v->ensure_scratch(sizeof(double) * size_t(std::max<int64_t>(v->total_units[GRAN_1K], 1)) + 64);
        if (rc != FLAGG_OK) return rc;
    }
    return row_combine_launch(v, m, d_rows, d_w1, d_w2, 1.0f, d_divisor, out, prev, delta_sumsq, 0,
                              static_cast<cudaStream_t>(stream));
}

int flagg_weiszfeld_weights(const double* sumsq, int n, float* w, float* wsum, void* stream) {
    FLAGG_REQUIRE(sumsq && w && wsum && n >= 1, "flagg_weiszfeld_weights: bad argument");
    weiszfeld_weights_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(sumsq, n, w, wsum);
    FLAGG_CUDA(cudaGetLastError());
    return FLAGG_OK;
}

}  // extern "C"

// ---- K2 without any allocation (no counterpart in the reference) ----
// flagg_row_reduce keeps its per-CTA partials in the view's scratch block, which ensure_scratch grows with
// cudaMallocAsync.  Callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid it) get the
// same two kernels with the partials in this file-scope array instead: FLAGG_STATIC_SCRATCH_BYTES per GPU that loads
// the library, held until exit.  The name below is a device symbol: host code may not pass it to a kernel; its device
// address is looked up once per device with cudaGetSymbolAddress.
static __device__ __align__(256) unsigned char g_rr_static_scratch[FLAGG_STATIC_SCRATCH_BYTES];

namespace flagg {

constexpr int RRS_MAX_DEVICES = 64;

// Device address of g_rr_static_scratch on the CURRENT device (each device has its own copy of the array).
static int rr_static_scratch(unsigned char** out) {
    static void* addr[RRS_MAX_DEVICES] = {};
    static bool known[RRS_MAX_DEVICES] = {};
    static std::mutex mu;
    int dev = -1;
    FLAGG_CUDA(cudaGetDevice(&dev));
    FLAGG_REQUIRE(dev >= 0 && dev < RRS_MAX_DEVICES, "flagg_row_reduce_static: device ordinal %d outside 0..%d", dev, RRS_MAX_DEVICES - 1);
    std::lock_guard<std::mutex> lock(mu);
    if (!known[dev]) {
        void* p = nullptr;
        FLAGG_CUDA(cudaGetSymbolAddress(&p, g_rr_static_scratch));
        FLAGG_REQUIRE(p != nullptr, "flagg_row_reduce_static: cudaGetSymbolAddress returned a NULL address on device %d", dev);
        addr[dev] = p;
        known[dev] = true;
    }
    *out = static_cast<unsigned char*>(addr[dev]);
    return FLAGG_OK;
}

}  // namespace flagg

extern "C" int flagg_row_reduce_static(const flagg_view* cv, const float* g0, int centered, double* sumsq, double* dot,
                                       float* linf, double* g0_sumsq, void* stream) {
    FLAGG_REQUIRE(cv, "flagg_row_reduce_static: NULL view");
    flagg_view* v = const_cast<flagg_view*>(cv);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int n = v->dev.n;
    if (v->dev.d == 0) {   // empty column shard: zeros, as flagg_row_reduce gives
        if (sumsq) FLAGG_CUDA(cudaMemsetAsync(sumsq, 0, sizeof(double) * n, st));
        if (dot) FLAGG_CUDA(cudaMemsetAsync(dot, 0, sizeof(double) * n, st));
        if (linf) FLAGG_CUDA(cudaMemsetAsync(linf, 0, sizeof(float) * n, st));
        if (g0_sumsq) FLAGG_CUDA(cudaMemsetAsync(g0_sumsq, 0, sizeof(double), st));
        return FLAGG_OK;
    }
    FLAGG_REQUIRE(!(dot && !g0), "flagg_row_reduce_static: dot requested without g0");
    FLAGG_REQUIRE(!(g0_sumsq && !g0), "flagg_row_reduce_static: g0_sumsq requested without g0");
    FLAGG_REQUIRE(!(centered && !g0), "flagg_row_reduce_static: centered statistics need g0");
    int dev = -1;
    FLAGG_CUDA(cudaGetDevice(&dev));
    FLAGG_REQUIRE(dev == v->device, "flagg_row_reduce_static: the view was created on device %d, the current device is %d", v->device, dev);
    v->last_stream = stream;   // the kernels read the view's tables: flagg_view_destroy orders their release after this stream
    const int64_t units = v->total_units[GRAN_2K];
    const size_t smem = size_t(n) * 16 + RR_CHUNK * 4 + size_t(n) * 4;
    FLAGG_REQUIRE(smem <= 200 * 1024, "flagg_row_reduce_static: n=%d too large for the per-CTA accumulators", n);
    const int sms = device_sm_count();
    FLAGG_REQUIRE(sms > 0, "cannot query the SM count of the current device");
    // the grid, the partial layout and both kernels are flagg_row_reduce's: the results are the same bit for bit
    const int nblocks = int(std::min<int64_t>(std::max<int64_t>(units, 1), int64_t(sms) * 4));
    const size_t b_d = sizeof(double) * size_t(nblocks) * n;
    const size_t b_f = sizeof(float) * size_t(nblocks) * n;
    const size_t need = 2 * b_d + ((b_f + 7) & ~size_t(7)) + sizeof(double) * nblocks;
    if (need > size_t(FLAGG_STATIC_SCRATCH_BYTES)) {
        set_error("flagg_row_reduce_static: %d clients x %d CTAs need %zu bytes of partial sums, the static scratch array holds "
                  "%zu (FLAGG_STATIC_SCRATCH_BYTES); use flagg_row_reduce", n, nblocks, need, size_t(FLAGG_STATIC_SCRATCH_BYTES));
        return FLAGG_EUNSUPPORTED;
    }
    unsigned char* s = nullptr;
    int rc = rr_static_scratch(&s);
    if (rc != FLAGG_OK) return rc;
    // Nothing in the array has to start at zero: every CTA writes all of its n x 3 + 1 partials, and the finalize kernel
    // reads exactly the nblocks x n entries this launch wrote -- what an earlier call left behind is never read.
    double* part_ss = reinterpret_cast<double*>(s);
    double* part_dot = reinterpret_cast<double*>(s + b_d);
    double* part_g0 = reinterpret_cast<double*>(s + 2 * b_d);
    float* part_linf = reinterpret_cast<float*>(s + 2 * b_d + sizeof(double) * nblocks);
    const bool vec4 = v->align_bytes >= 16;
    auto launch = [&](auto kern) -> int {
        FLAGG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
        kern<<<nblocks, RR_THREADS, smem, st>>>(v->dev, g0, part_ss, part_dot, part_linf, part_g0, units);
        FLAGG_CUDA(cudaGetLastError());
        return FLAGG_OK;
    };
    if (vec4 && g0 && centered) rc = launch(row_reduce_kernel<true, true, true>);
    else if (g0 && centered) rc = launch(row_reduce_kernel<false, true, true>);
    else if (vec4) rc = launch(row_reduce_kernel<true, true, false>);      // g0 == NULL: staged zeros, see flagg_row_reduce
    else if (g0) rc = launch(row_reduce_kernel<false, true, false>);
    else rc = launch(row_reduce_kernel<false, false, false>);
    if (rc != FLAGG_OK) return rc;
    row_reduce_finalize<<<(n + 1 + RRF_THREADS / 32 - 1) / (RRF_THREADS / 32), RRF_THREADS, 0, st>>>(part_ss, part_dot, part_linf, part_g0, nblocks, n,
                                                         sumsq, dot, linf, g0_sumsq);
    FLAGG_CUDA(cudaGetLastError());
    return FLAGG_OK;
}
