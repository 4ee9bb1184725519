// Hardware-layout probe for the K4 TS kernel (GPU box only):
//   (1) which (lane, column) each register of tcgen05.st.16x256b.x2 lands on, read back with
//       tcgen05.ld.32x32b;
//   (2) tcgen05.mma kind::tf32 with A in tensor memory (lane = row, column = k) against a
//       128B-swizzled K-major B tile in shared memory, checked against the host product.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -o /tmp/tmem_probe tools/tmem_probe.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <math.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return uint32_t(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr) {
    return uint64_t((saddr & 0x3ffffu) >> 4) | (uint64_t(1) << 16) | (uint64_t(1024 >> 4) << 32) | (uint64_t(1) << 46) | (uint64_t(2) << 61);
}
constexpr uint32_t IDESC = (1u << 4) | (2u << 7) | (2u << 10) | (uint32_t(128 >> 3) << 17) | (uint32_t(128 >> 4) << 24);

__global__ void __launch_bounds__(128, 1) probe(float* map_out /*[128][16]*/, const float* A /*[128][8]*/, const float* B /*[128][8]*/, float* D /*[128][128]*/) {
    extern __shared__ unsigned char raw[];
    const uint32_t sbase = (smem_u32(raw) + 1023u) & ~1023u;
    unsigned char* smem = raw + (sbase - smem_u32(raw));
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 16384);
    uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 1);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" ::"r"(smem_u32(slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tm = *slot;
    // ---- (1) fragment map: each warp stores two 16-lane halves; value = 1000*reg + lane_in_warp
    for (int half = 0; half < 2; ++half) {
        uint32_t u[8];
        for (int i = 0; i < 8; ++i) u[i] = __float_as_uint(float(1000 * (i + 8 * half) + lane));
        const uint32_t addr = tm + (uint32_t(32 * warp + 16 * half) << 16) + 128;
        asm volatile("tcgen05.st.sync.aligned.16x256b.x2.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(addr), "r"(u[0]), "r"(u[1]), "r"(u[2]), "r"(u[3]), "r"(u[4]), "r"(u[5]), "r"(u[6]), "r"(u[7]) : "memory");
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    {
        uint32_t r[16];
        const uint32_t addr = tm + (uint32_t(32 * warp) << 16) + 128;
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                     : "r"(addr) : "memory");
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        for (int i = 0; i < 16; ++i) map_out[tid * 16 + i] = __uint_as_float(r[i]);
    }
    // ---- (2) A (128 x 8) into TMEM columns [192, 200) with 32x32b (lane = row), B into swizzled smem
    {
        uint32_t a[8];
        for (int k = 0; k < 8; ++k) a[k] = __float_as_uint(A[tid * 8 + k]);
        const uint32_t addr = tm + (uint32_t(32 * warp) << 16) + 192;
        asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(addr), "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(a[4]), "r"(a[5]), "r"(a[6]), "r"(a[7]) : "memory");
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        for (int i = tid; i < 16384 / 4; i += 128) reinterpret_cast<float*>(smem)[i] = 0.f;
        __syncthreads();
        const int r = tid;   // B row (N index)
        for (int k = 0; k < 8; ++k) {
            const uint32_t off = uint32_t(r >> 3) * 1024u + uint32_t(r & 7) * 128u + ((uint32_t(k >> 2) ^ uint32_t(r & 7)) << 4) + uint32_t(k & 3) * 4u;
            *reinterpret_cast<float*>(smem + off) = B[r * 8 + k];
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        if (tid == 0) {
            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tm), "r"(tm + 192), "l"(make_desc(sbase)), "r"(IDESC), "r"(0u) : "memory");
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
        }
        asm volatile("{\n\t.reg .pred p;\n\tW: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n\t@p bra DN;\n\tbra W;\n\tDN:\n\t}" ::"r"(smem_u32(bar)) : "memory");
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        for (int g = 0; g < 8; ++g) {
            uint32_t r16[16];
            const uint32_t addr = tm + (uint32_t(32 * warp) << 16) + g * 16;
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                         : "=r"(r16[0]), "=r"(r16[1]), "=r"(r16[2]), "=r"(r16[3]), "=r"(r16[4]), "=r"(r16[5]), "=r"(r16[6]), "=r"(r16[7]), "=r"(r16[8]), "=r"(r16[9]), "=r"(r16[10]), "=r"(r16[11]), "=r"(r16[12]), "=r"(r16[13]), "=r"(r16[14]), "=r"(r16[15])
                         : "r"(addr) : "memory");
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            for (int i = 0; i < 16; ++i) D[tid * 128 + g * 16 + i] = __uint_as_float(r16[i]);
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" ::"r"(tm) : "memory");
}

int main() {
    float *map, *A, *B, *D;
    cudaMallocManaged(&map, 128 * 16 * 4); cudaMallocManaged(&A, 128 * 8 * 4); cudaMallocManaged(&B, 128 * 8 * 4); cudaMallocManaged(&D, 128 * 128 * 4);
    for (int i = 0; i < 128 * 8; ++i) { A[i] = float((i * 37) % 64 - 32) / 8.f; B[i] = float((i * 53) % 64 - 32) / 16.f; }
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 20000);
    probe<<<1, 128, 20000>>>(map, A, B, D);
    cudaError_t e = cudaDeviceSynchronize();
    printf("probe: %s\n", cudaGetErrorString(e));
    if (e != cudaSuccess) return 1;
    // (1) expected: lane L (within warp) column C holds register 4*(C/8) + 2*((L%16)/8) + (C%2) [+8 for the upper half] of thread 4*(L%8) + (C%8)/2
    int bad = 0;
    for (int L = 0; L < 128; ++L)
        for (int C = 0; C < 16; ++C) {
            const int l = L % 32, half = l / 16, reg = 4 * (C / 8) + 2 * ((l % 16) / 8) + (C % 2) + 8 * half, thr = 4 * (l % 8) + (C % 8) / 2;
            if (map[L * 16 + C] != float(1000 * reg + thr)) {
                if (bad < 8) printf("  map mismatch lane %d col %d: got %.0f expected %d\n", L, C, map[L * 16 + C], 1000 * reg + thr);
                ++bad;
            }
        }
    printf("fragment map 16x256b.x2: %s (%d mismatches)\n", bad ? "DIFFERENT" : "as assumed", bad);
    if (bad) { printf("lanes 0-15, cols 0-15 (1000*reg + thread):\n"); for (int L = 0; L < 16; ++L) { for (int C = 0; C < 16; ++C) printf("%6.0f", map[L * 16 + C]); printf("\n"); } }
    double worst = 0;
    for (int m = 0; m < 128; ++m)
        for (int n = 0; n < 128; ++n) {
            double s = 0;
            for (int k = 0; k < 8; ++k) s += double(A[m * 8 + k]) * B[n * 8 + k];
            worst = fmax(worst, fabs(s - D[m * 128 + n]));
        }
    printf("TS mma (A in TMEM lane=row col=k) max |err| = %.3g -> %s\n", worst, worst < 1e-4 ? "OK" : "WRONG");
    return (bad || worst >= 1e-4) ? 2 : 0;
}
