// Micro-benchmark: throughput of the compare-exchange building blocks on sm_100a.
// Prints thread-ops per clock per SM for each variant (cycles from clock64 of one CTA).
#include <cstdio>
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#define ITERS 2048
#define ILP 16
template <int MODE>
__global__ void __launch_bounds__(256) k(int* out, int seed, long long* cyc, int one, int mone) {
    int v[ILP], w[ILP];
    for (int i = 0; i < ILP; ++i) { v[i] = seed * (threadIdx.x + 1) + i * 7919; w[i] = seed ^ (i * 104729 + threadIdx.x); }
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) {
            if (MODE == 0) {            // integer CE: min + max
                int lo = min(v[i], w[i]), hi = max(v[i], w[i]); v[i] = lo + it; w[i] = hi;
            } else if (MODE == 1) {     // float CE: fmin + fmax
                float a = __int_as_float(v[i]), b = __int_as_float(w[i]);
                float lo = fminf(a, b), hi = fmaxf(a, b); v[i] = __float_as_int(lo) + it; w[i] = __float_as_int(hi);
            } else if (MODE == 2) {     // integer CE: min + (a+b-min)
                int lo = min(v[i], w[i]); int hi = v[i] + w[i] - lo; v[i] = lo + it; w[i] = hi;
            } else if (MODE == 3) {     // integer CE: min + xor trick
                int lo = min(v[i], w[i]); int hi = v[i] ^ w[i] ^ lo; v[i] = lo + it; w[i] = hi;
            } else if (MODE == 4) {     // 3-input min and max
                int lo = min(min(v[i], w[i]), v[(i + 1) % ILP]); int hi = max(max(v[i], w[i]), v[(i + 1) % ILP]); v[i] = lo + it; w[i] = hi;
            } else if (MODE == 5) {     // compare + two selects
                bool p = v[i] > w[i]; int lo = p ? w[i] : v[i]; int hi = p ? v[i] : w[i]; v[i] = lo + it; w[i] = hi;
            } else if (MODE == 6) {     // baseline: the "+ it" add alone (to subtract)
                v[i] = v[i] + it; w[i] = w[i] ^ v[i];
            } else if (MODE == 8) {     // f32 min + hi via two runtime-multiplier IMADs on the bit patterns
                float lo = fminf(__int_as_float(v[i]), __int_as_float(w[i])); int t, hi;
                asm("mad.lo.s32 %0, %1, %2, %3;" : "=r"(t) : "r"(v[i]), "r"(one), "r"(w[i]));
                asm("mad.lo.s32 %0, %1, %2, %3;" : "=r"(hi) : "r"(__float_as_int(lo)), "r"(mone), "r"(t));
                v[i] = __float_as_int(lo) + it; w[i] = hi;
            } else if (MODE == 9) {     // IMAD alone (two per item)
                int t, hi;
                asm("mad.lo.s32 %0, %1, %2, %3;" : "=r"(t) : "r"(v[i]), "r"(one), "r"(w[i]));
                asm("mad.lo.s32 %0, %1, %2, %3;" : "=r"(hi) : "r"(w[i]), "r"(mone), "r"(t));
                v[i] = t; w[i] = hi;
            } else if (MODE == 10) {    // 1/3 min+max, 2/3 min+2xIMAD
                float a = __int_as_float(v[i]), b = __int_as_float(w[i]);
                float lo = fminf(a, b); int hi;
                if (i % 3 == 0) { hi = __float_as_int(fmaxf(a, b)); }
                else { int t; asm("mad.lo.s32 %0, %1, %2, %3;" : "=r"(t) : "r"(v[i]), "r"(one), "r"(w[i]));
                       asm("mad.lo.s32 %0, %1, %2, %3;" : "=r"(hi) : "r"(__float_as_int(lo)), "r"(mone), "r"(t)); }
                v[i] = __float_as_int(lo) + it; w[i] = hi;
            } else if (MODE == 11) {    // 1/2 min+max, 1/2 min+2xIMAD
                float a = __int_as_float(v[i]), b = __int_as_float(w[i]);
                float lo = fminf(a, b); int hi;
                if (i % 2 == 0) { hi = __float_as_int(fmaxf(a, b)); }
                else { int t; asm("mad.lo.s32 %0, %1, %2, %3;" : "=r"(t) : "r"(v[i]), "r"(one), "r"(w[i]));
                       asm("mad.lo.s32 %0, %1, %2, %3;" : "=r"(hi) : "r"(__float_as_int(lo)), "r"(mone), "r"(t)); }
                v[i] = __float_as_int(lo) + it; w[i] = hi;
            } else if (MODE == 12) {    // pure min+max, no extra add (network-like): values recirculate
                int lo = min(v[i], w[i]), hi = max(v[i], w[i]); v[i] = hi; w[(i + 1) % ILP] = lo;
            } else if (MODE == 13) {    // f32 min + predicated-select hi on fma pipe: hi = a + b - lo in FLOAT for compare only
                float a = __int_as_float(v[i]), b = __int_as_float(w[i]);
                float lo = fminf(a, b); float hi = fmaf(a, 1.0f, b) - lo;   // (inexact; throughput probe only)
                v[i] = __float_as_int(lo) + it; w[i] = __float_as_int(hi);
            } else if (MODE == 14) {    // packed s16x2 CE (DPX): 2 columns per register
                unsigned a = v[i], b = w[i];
                unsigned lo = __vmins2(a, b), hi = __vmaxs2(a, b); v[i] = lo + it; w[i] = hi;
            } else if (MODE == 15) {    // half2 CE
                __half2 a = *reinterpret_cast<__half2*>(&v[i]), b = *reinterpret_cast<__half2*>(&w[i]);
                __half2 lo = __hmin2(a, b), hi = __hmax2(a, b);
                v[i] = *reinterpret_cast<int*>(&lo) + it; w[i] = *reinterpret_cast<int*>(&hi);
            } else if (MODE == 16) {    // f32 min + hi via one IMAD + one IADD (t on fma pipe, sub on alu)
                float lo = fminf(__int_as_float(v[i]), __int_as_float(w[i])); int t;
                asm("mad.lo.s32 %0, %1, %2, %3;" : "=r"(t) : "r"(v[i]), "r"(one), "r"(w[i]));
                int hi = t - __float_as_int(lo);
                v[i] = __float_as_int(lo) + it; w[i] = hi;
            } else if (MODE == 17) {    // f32 3-input min/max
                float a = __int_as_float(v[i]), b = __int_as_float(w[i]), c = __int_as_float(v[(i + 1) % ILP]);
                float lo, hi;
                asm("min.f32 %0, %1, %2, %3;" : "=f"(lo) : "f"(a), "f"(b), "f"(c));
                asm("max.f32 %0, %1, %2, %3;" : "=f"(hi) : "f"(a), "f"(b), "f"(c));
                v[i] = __float_as_int(lo) + it; w[i] = __float_as_int(hi);
            } else if (MODE == 18) {    // pure f32 min+max recirculating (no add)
                float a = __int_as_float(v[i]), b = __int_as_float(w[i]);
                v[i] = __float_as_int(fmaxf(a, b)); w[(i + 1) % ILP] = __float_as_int(fminf(a, b));
            } else if (MODE == 19) {    // f32 min + 2 IMAD recirculating (no add)
                float lo = fminf(__int_as_float(v[i]), __int_as_float(w[i])); int t, hi;
                asm("mad.lo.s32 %0, %1, %2, %3;" : "=r"(t) : "r"(v[i]), "r"(one), "r"(w[i]));
                asm("mad.lo.s32 %0, %1, %2, %3;" : "=r"(hi) : "r"(__float_as_int(lo)), "r"(mone), "r"(t));
                v[i] = hi; w[(i + 1) % ILP] = __float_as_int(lo);
            } else if (MODE == 20) {    // recirculating mix: 2 of 3 CEs via IMAD
                float a = __int_as_float(v[i]), b = __int_as_float(w[i]);
                float lo = fminf(a, b); int hi;
                if (i % 3 == 0) { hi = __float_as_int(fmaxf(a, b)); }
                else { int t; asm("mad.lo.s32 %0, %1, %2, %3;" : "=r"(t) : "r"(v[i]), "r"(one), "r"(w[i]));
                       asm("mad.lo.s32 %0, %1, %2, %3;" : "=r"(hi) : "r"(__float_as_int(lo)), "r"(mone), "r"(t)); }
                v[i] = hi; w[(i + 1) % ILP] = __float_as_int(lo);
            } else if (MODE == 21) {    // recirculating mix: 1 of 2 CEs via IMAD
                float a = __int_as_float(v[i]), b = __int_as_float(w[i]);
                float lo = fminf(a, b); int hi;
                if (i % 2 == 0) { hi = __float_as_int(fmaxf(a, b)); }
                else { int t; asm("mad.lo.s32 %0, %1, %2, %3;" : "=r"(t) : "r"(v[i]), "r"(one), "r"(w[i]));
                       asm("mad.lo.s32 %0, %1, %2, %3;" : "=r"(hi) : "r"(__float_as_int(lo)), "r"(mone), "r"(t)); }
                v[i] = hi; w[(i + 1) % ILP] = __float_as_int(lo);
            } else if (MODE == 7) {     // pure FFMA (fma pipe reference)
                float a = __int_as_float(v[i]), b = __int_as_float(w[i]);
                a = fmaf(a, 1.0001f, b); b = fmaf(b, 0.9999f, a); v[i] = __float_as_int(a); w[i] = __float_as_int(b);
            }
        }
    }
    long long t1 = clock64();
    int s = 0;
    for (int i = 0; i < ILP; ++i) s ^= v[i] ^ w[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int MODE> void run(const char* name, int ctas_per_sm) {
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    int grid = sms * ctas_per_sm; int* out; long long* cyc; long long h;
    cudaMalloc(&out, grid * 256 * 4); cudaMalloc(&cyc, 8);
    k<MODE><<<grid, 256>>>(out, 3, cyc, 1, -1); cudaDeviceSynchronize();
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a); k<MODE><<<grid, 256>>>(out, 5, cyc, 1, -1); cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    double items = double(ITERS) * ILP * 256 * ctas_per_sm;   // per SM
    printf("%-34s ctas/SM=%d  %8.3f ms  cycles=%lld  items/clk/SM=%6.2f  (chip %7.1f G items/s)\n", name, ctas_per_sm, ms, h,
           items / double(h), items * sms / (ms * 1e6));
    cudaFree(out); cudaFree(cyc);
}
int main() {
    for (int c : {2, 4}) {
        run<0>("int CE  min+max", c); run<1>("f32 CE  fmin+fmax", c); run<2>("int CE  min + (a+b-min)", c);
        run<3>("int CE  min + xor", c); run<4>("int min3+max3", c); run<5>("int cmp + 2 sel", c);
        run<6>("baseline add+xor", c); run<7>("2 x FFMA", c);
        run<8>("f32 min + 2 IMAD", c); run<9>("2 IMAD", c); run<10>("mix 1/3 minmax 2/3 min+2IMAD", c);
        run<11>("mix 1/2 minmax 1/2 min+2IMAD", c); run<12>("int min+max only", c); run<13>("f32 min + FFMA + FADD", c);
        run<14>("s16x2 vimin+vimax (2 cols)", c); run<15>("half2 hmin2+hmax2", c); run<16>("f32 min + IMAD + IADD", c); run<17>("f32 min3+max3", c);
        run<18>("recirc f32 min+max", c); run<19>("recirc f32 min + 2 IMAD", c); run<20>("recirc mix 2/3 IMAD", c); run<21>("recirc mix 1/2 IMAD", c);
    }
    return 0;
}
