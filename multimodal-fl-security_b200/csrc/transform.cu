// Row transforms: the reference's model-poisoning attacks as one elementwise pass over a client's
// P parameter tensors, in place on the rows of a resident update matrix when out == x
// (SURVEY 8f-3).  Replaces the per-tensor Python loops of
//   ModelReplacementAttack.poison_update   src/attacks/model_poisoning.py:71-78
//   ScalingAttack.poison_update            :236
//   InnerProductManipulationAttack         :271-283
//   AdaptiveKrumAttack.poison_update       :172-184 (deterministic target_direction form)
// with the reference's own fp32 operation order (every multiply / add / subtract rounded
// separately), so results are bit-identical to the CPU path.  HBM-bound: 128-bit loads and stores
// when all three pointers of a tensor are 16-byte aligned, scalar otherwise.
#include "common.cuh"

namespace flagg {

constexpr int TR_SEGS = 32;        // tensors per launch (pointer tables travel as kernel arguments)
constexpr int TR_UNIT = 4096;      // elements per CTA
constexpr int TR_THREADS = 256;

struct TransformArgs {
    const float* x[TR_SEGS];
    const float* y[TR_SEGS];
    float* o[TR_SEGS];
    long long unit_start[TR_SEGS + 1];   // prefix of TR_UNIT-element units per tensor
    long long numel[TR_SEGS];
    int nseg;
};

template <int OP>
__device__ __forceinline__ float tr_apply(float x, float y, float s) {
    if (OP == FLAGG_T_SCALE) return __fmul_rn(x, s);
    if (OP == FLAGG_T_REPLACE) return __fadd_rn(y, __fmul_rn(__fsub_rn(x, y), s));
    if (OP == FLAGG_T_SIGN) return __fmul_rn(s, float(x > 0.f) - float(x < 0.f));   // torch.sign: 0 for +-0 and NaN
    return __fadd_rn(x, __fmul_rn(y, s));                                            // FLAGG_T_AXPY
}

template <int OP>
__global__ void __launch_bounds__(TR_THREADS) transform_kernel(const __grid_constant__ TransformArgs a, float s) {
    const long long u = blockIdx.x;
    int p = 0;
    while (p + 1 < a.nseg && a.unit_start[p + 1] <= u) ++p;
    const long long e0 = (u - a.unit_start[p]) * TR_UNIT;
    const int len = int(min((long long)TR_UNIT, a.numel[p] - e0));
    const float* x = a.x[p] + e0;
    const float* y = (OP == FLAGG_T_REPLACE || OP == FLAGG_T_AXPY) ? a.y[p] + e0 : nullptr;
    float* o = a.o[p] + e0;
    const bool needy = (OP == FLAGG_T_REPLACE || OP == FLAGG_T_AXPY);
    const bool al = ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(o) | (needy ? reinterpret_cast<uintptr_t>(y) : 0)) & 15) == 0;
    const int t = threadIdx.x;
    if (al) {
        const int nv = len >> 2;
#pragma unroll
        for (int j = 0; j < TR_UNIT / 4 / TR_THREADS; ++j) {
            const int q = t + j * TR_THREADS;
            if (q < nv) {
                const float4 xv = reinterpret_cast<const float4*>(x)[q];
                float4 yv = make_float4(0.f, 0.f, 0.f, 0.f);
                if (needy) yv = reinterpret_cast<const float4*>(y)[q];
                float4 r;
                r.x = tr_apply<OP>(xv.x, yv.x, s); r.y = tr_apply<OP>(xv.y, yv.y, s);
                r.z = tr_apply<OP>(xv.z, yv.z, s); r.w = tr_apply<OP>(xv.w, yv.w, s);
                reinterpret_cast<float4*>(o)[q] = r;
            }
        }
        for (int c = (nv << 2) + t; c < len; c += TR_THREADS) o[c] = tr_apply<OP>(x[c], needy ? y[c] : 0.f, s);
    } else {
        for (int c = t; c < len; c += TR_THREADS) o[c] = tr_apply<OP>(x[c], needy ? y[c] : 0.f, s);
    }
}

}  // namespace flagg

using namespace flagg;

extern "C" int flagg_segments_transform(int P, const int64_t* seg_numel, const void* const* x_ptrs,
                                        const void* const* y_ptrs, void* const* out_ptrs, int op, float s,
                                        void* stream) {
    FLAGG_REQUIRE(P >= 1 && seg_numel && x_ptrs && out_ptrs, "flagg_segments_transform: bad argument");
    FLAGG_REQUIRE(op >= FLAGG_T_SCALE && op <= FLAGG_T_AXPY, "flagg_segments_transform: unknown op %d", op);
    const bool needy = (op == FLAGG_T_REPLACE || op == FLAGG_T_AXPY);
    FLAGG_REQUIRE(!needy || y_ptrs, "flagg_segments_transform: this op needs the second operand");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    for (int p0 = 0; p0 < P; p0 += TR_SEGS) {
        TransformArgs a;
        a.nseg = 0;
        a.unit_start[0] = 0;
        for (int p = p0; p < P && a.nseg < TR_SEGS; ++p) {
            FLAGG_REQUIRE(seg_numel[p] >= 0, "flagg_segments_transform: negative tensor size");
            if (seg_numel[p] == 0) continue;
            FLAGG_REQUIRE(x_ptrs[p] && out_ptrs[p] && (!needy || y_ptrs[p]), "flagg_segments_transform: NULL tensor %d", p);
            FLAGG_REQUIRE(((reinterpret_cast<uintptr_t>(x_ptrs[p]) | reinterpret_cast<uintptr_t>(out_ptrs[p]) |
                            (needy ? reinterpret_cast<uintptr_t>(y_ptrs[p]) : 0)) & 3) == 0,
                          "flagg_segments_transform: tensor %d not 4-byte aligned", p);
            const int k = a.nseg++;
            a.x[k] = static_cast<const float*>(x_ptrs[p]);
            a.y[k] = needy ? static_cast<const float*>(y_ptrs[p]) : nullptr;
            a.o[k] = static_cast<float*>(out_ptrs[p]);
            a.numel[k] = seg_numel[p];
            a.unit_start[k + 1] = a.unit_start[k] + (seg_numel[p] + TR_UNIT - 1) / TR_UNIT;
        }
        if (a.nseg == 0) continue;
        const long long units = a.unit_start[a.nseg];
        FLAGG_REQUIRE(units < (1ll << 31), "flagg_segments_transform: too many elements in one launch");
        switch (op) {
            case FLAGG_T_SCALE: transform_kernel<FLAGG_T_SCALE><<<unsigned(units), TR_THREADS, 0, st>>>(a, s); break;
            case FLAGG_T_REPLACE: transform_kernel<FLAGG_T_REPLACE><<<unsigned(units), TR_THREADS, 0, st>>>(a, s); break;
            case FLAGG_T_SIGN: transform_kernel<FLAGG_T_SIGN><<<unsigned(units), TR_THREADS, 0, st>>>(a, s); break;
            default: transform_kernel<FLAGG_T_AXPY><<<unsigned(units), TR_THREADS, 0, st>>>(a, s); break;
        }
        FLAGG_CUDA(cudaGetLastError());
    }
    return FLAGG_OK;
}
