// flagg_view: the n x P segmented description of the client-update matrix, plus
// library-level plumbing (error string, device info).
#include <cuda.h>
#include <cudaTypedefs.h>
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"

namespace flagg {

static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link-time libcuda dependency).
static PFN_cuTensorMapEncodeTiled_v12000 tensor_map_encoder() {
    static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
    }
    return fn;
}

int device_sm_count() {
    static thread_local int cached_dev = -1, cached = 0;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;   // callers treat 0 as an error: no silent fallback
    if (dev != cached_dev) {
        if (cudaDeviceGetAttribute(&cached, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
        cached_dev = dev;
    }
    return cached;
}

}  // namespace flagg

using namespace flagg;

// Device memory of a view (tables, scratch) comes from CUDA's stream-ordered allocator with the
// pool told to keep what is freed: a federated round creates and destroys a view per aggregate()
// call, and cudaMalloc + cudaFree (a device-wide synchronisation) cost ~1 ms of a call whose
// kernels take microseconds at the small shapes (tools/api_latency.py).  Frees are ordered on the
// stream the view was last used on.
static int pool_ready(int device) {
    static bool done[64] = {};
    if (device < 0 || device >= 64 || done[device]) return FLAGG_OK;
    cudaMemPool_t pool;
    FLAGG_CUDA(cudaDeviceGetDefaultMemPool(&pool, device));
    uint64_t keep = ~uint64_t(0);
    FLAGG_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));
    done[device] = true;
    return FLAGG_OK;
}

int flagg_view::ensure_scratch(size_t bytes) {
    if (bytes <= scratch_bytes) return FLAGG_OK;
    cudaStream_t st = static_cast<cudaStream_t>(last_stream);
    if (d_scratch) FLAGG_CUDA(cudaFreeAsync(d_scratch, st));
    d_scratch = nullptr;
    scratch_bytes = 0;
    size_t want = (bytes + 0xffff) & ~size_t(0xffff);
    FLAGG_CUDA(cudaMallocAsync(&d_scratch, want, st));
    scratch_bytes = want;
    return FLAGG_OK;
}

extern "C" {

int flagg_version(void) { return FLAGG_VERSION; }

const char* flagg_last_error(void) { return g_err; }

int flagg_device_info(int* sm_count, int* cc_major, int* cc_minor) {
    int dev = 0;
    FLAGG_CUDA(cudaGetDevice(&dev));
    int v = 0;
    if (sm_count) { FLAGG_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev)); *sm_count = v; }
    if (cc_major) { FLAGG_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrComputeCapabilityMajor, dev)); *cc_major = v; }
    if (cc_minor) { FLAGG_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrComputeCapabilityMinor, dev)); *cc_minor = v; }
    return FLAGG_OK;
}

int flagg_view_create(flagg_view** out, int n, int P, const int64_t* seg_numel,
                      const void* const* seg_ptrs, void* stream) {
    FLAGG_REQUIRE(out && seg_numel && seg_ptrs, "flagg_view_create: NULL argument");
    FLAGG_REQUIRE(n >= 1 && P >= 1, "flagg_view_create: need n >= 1 and P >= 1 (got n=%d P=%d)", n, P);
    flagg_view* v = new (std::nothrow) flagg_view();
    if (!v) { set_error("flagg_view_create: out of host memory"); return FLAGG_ENOMEM; }
    v->seg_numel.assign(seg_numel, seg_numel + P);
    v->seg_start.resize(P + 1);
    v->seg_start[0] = 0;
    for (int g = 0; g < NUM_GRAN; ++g) v->unit_start[g].assign(P + 1, 0);
    for (int p = 0; p < P; ++p) {
        if (seg_numel[p] < 0) { delete v; set_error("flagg_view_create: negative segment size"); return FLAGG_EINVAL; }
        v->seg_start[p + 1] = v->seg_start[p] + seg_numel[p];
        for (int g = 0; g < NUM_GRAN; ++g)
            v->unit_start[g][p + 1] = v->unit_start[g][p] + (seg_numel[p] + unit_cols(g) - 1) / unit_cols(g);
    }
    for (int g = 0; g < NUM_GRAN; ++g) v->total_units[g] = v->unit_start[g][P];
    v->seg_ptrs.assign(seg_ptrs, seg_ptrs + size_t(n) * P);
    int align = 16;
    for (size_t i = 0; i < v->seg_ptrs.size(); ++i) {
        uintptr_t a = reinterpret_cast<uintptr_t>(v->seg_ptrs[i]);
        if (a == 0 && v->seg_numel[i % P] > 0) { delete v; set_error("flagg_view_create: NULL segment pointer"); return FLAGG_EINVAL; }
        if (a & 3) { delete v; set_error("flagg_view_create: segment pointer not 4-byte aligned"); return FLAGG_EINVAL; }
        if (v->seg_numel[i % P] == 0) continue;
        if (a & 15) align = (a & 7) ? 4 : (align < 8 ? align : 8);
    }
    v->align_bytes = align;
    cudaError_t e = cudaGetDevice(&v->device);
    if (e != cudaSuccess) { delete v; set_error("cudaGetDevice: %s", cudaGetErrorString(e)); return FLAGG_ECUDA; }

    // Dense detection: rows of every tensor equally spaced (one [n x len] matrix per tensor, e.g. the
    // rows of a stacked update matrix or flagg_upload_rows' staging matrix) -> 2-D TMA tensor maps.
    std::vector<CUtensorMap> tmaps;
    if (n >= 2 && n <= 256 && align == 16) {
        PFN_cuTensorMapEncodeTiled_v12000 encode = tensor_map_encoder();
        bool ok = encode != nullptr;
        tmaps.resize(P);
        memset(tmaps.data(), 0, sizeof(CUtensorMap) * P);
        for (int p = 0; p < P && ok; ++p) {
            if (v->seg_numel[p] == 0) continue;
            const char* r0 = static_cast<const char*>(v->seg_ptrs[p]);
            const int64_t pitch = static_cast<const char*>(v->seg_ptrs[size_t(P) + p]) - r0;
            ok = pitch >= 4 * v->seg_numel[p] && (pitch & 15) == 0 && pitch < (int64_t(1) << 40) &&
                 v->seg_numel[p] < (int64_t(1) << 31);
            for (int i = 2; i < n && ok; ++i) ok = static_cast<const char*>(v->seg_ptrs[size_t(i) * P + p]) == r0 + int64_t(i) * pitch;
            if (!ok) break;
            const cuuint64_t gdim[2] = {cuuint64_t(v->seg_numel[p]), cuuint64_t(n)};
            const cuuint64_t gstride[1] = {cuuint64_t(pitch)};
            // 256-column boxes (two 256-thread CTAs per SM: eight warps in one barrier domain share their instruction
            // fetches, no_instruction stalls 0.72 -> 0.17 per issue) where the networks are issue-bound, 80 <= n <= 110:
            // C3 2.03 -> 1.995 ms; 128-column boxes (3-4 CTAs of 128) elsewhere -- the HBM-bound sizes measured 2-8 %
            // slower with the wide form (profiles/r02_k1_layouts.txt).  FLAGG_K1_DENSE_COLS=128|256 forces one (tests).
            const char* dc = getenv("FLAGG_K1_DENSE_COLS");
            cuuint32_t bcols = (n >= 80 && n <= 110) ? 256u : 128u;
            if (dc && atoi(dc) == 128) bcols = 128u;
            if (dc && atoi(dc) == 256 && n <= 110) bcols = 256u;
            v->dev.dense_cols = int(bcols);
            const cuuint32_t box[2] = {bcols, cuuint32_t(n)};
            const cuuint32_t estride[2] = {1u, 1u};
            const CUresult r = encode(&tmaps[p], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<char*>(r0), gdim, gstride, box,
                                      estride, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                      CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            ok = (r == CUDA_SUCCESS);
        }
        v->dense = ok;
    }

    // Gather map for rows that are NOT one matrix (see ViewDev::gmap).
    CUtensorMap gmap[2];      // [0] 256-column overlapped rows (when the encoder takes them), [1] 128-column rows
    std::vector<int32_t> seg_rows;
    memset(gmap, 0, sizeof(gmap));
    if (!v->dense && n >= 2 && n <= 256 && align == 16) {
        PFN_cuTensorMapEncodeTiled_v12000 encode = tensor_map_encoder();
        uintptr_t lo = ~uintptr_t(0), hi = 0;
        for (size_t i = 0; i < v->seg_ptrs.size(); ++i) {
            if (v->seg_numel[i % P] == 0) continue;
            const uintptr_t a = reinterpret_cast<uintptr_t>(v->seg_ptrs[i]);
            lo = a < lo ? a : lo;
            const uintptr_t end = a + ((uintptr_t(v->seg_numel[i % P]) * 4 + 511) & ~uintptr_t(511));
            hi = end > hi ? end : hi;
        }
        bool ok = encode != nullptr && hi > lo && ((hi - lo) >> 9) < (uintptr_t(1) << 31);
        const int npad = (n + 3) & ~3;                      // transposed [P][npad]: one 16-byte load per gather4 request
        seg_rows.assign(size_t(P) * npad, 0);
        for (size_t i = 0; i < v->seg_ptrs.size() && ok; ++i) {
            if (v->seg_numel[i % P] == 0) continue;
            // truly 512-byte aligned (not merely congruent): a tensor's ragged last unit is then read to the end of a
            // 512-byte granule that holds valid data, which no allocation boundary or page boundary can cut
            const uintptr_t a = reinterpret_cast<uintptr_t>(v->seg_ptrs[i]);
            ok = (a & 511) == 0;
            seg_rows[(i % P) * npad + i / P] = int32_t((a - lo) >> 9);
        }
        for (int p = 0; p < P; ++p)
            for (int i = n; i < npad; ++i) seg_rows[size_t(p) * npad + i] = seg_rows[size_t(p) * npad + n - 1];
        if (ok) {
            // Rows of 256 columns at a 512-byte pitch (consecutive rows overlap by half): a 256-column unit of a tensor
            // is row seg_rows + 2 * unit, and a gather4 request carries 4 KB.  The last row is kept inside the range.
            // Used where the networks are issue-bound (80 <= n <= 110, like the dense feed); below that 128-column
            // rows (2 KB per request, four CTAs of 128 threads) measured 3-14 % faster, n = 32: 6.8 against 5.9 TB/s
            // (profiles/r02_k1_layouts.txt).  FLAGG_K1_GATHER_COLS=128|256 forces one (tests, tuning).
            const char* gc = getenv("FLAGG_K1_GATHER_COLS");
            const cuuint64_t rows = cuuint64_t((hi - lo) >> 9);
            const cuuint64_t gstride[1] = {512u};
            const cuuint32_t estride[2] = {1u, 1u};
            v->dev.gather_cols = 128;
            const cuuint64_t gdim1[2] = {128u, rows};
            const cuuint32_t box1[2] = {128u, 1u};         // tile::gather4: four such rows per request
            CUresult r = encode(&gmap[1], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, reinterpret_cast<void*>(lo), gdim1, gstride, box1, estride,
                                CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            const bool want_wide = gc ? atoi(gc) == 256 : (n >= 80 && n <= 110);
            if (r == CUDA_SUCCESS && want_wide && rows >= 2) {
                const cuuint64_t gdim2[2] = {256u, rows - 1};
                const cuuint32_t box2[2] = {256u, 1u};
                if (encode(&gmap[0], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, reinterpret_cast<void*>(lo), gdim2, gstride, box2, estride,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS)
                    v->dev.gather_cols = 256;
            }
            ok = (r == CUDA_SUCCESS);
        }
        v->gather = ok;
    }

    const size_t b_ptrs = sizeof(void*) * size_t(n) * P;
    const size_t b_pref = sizeof(int64_t) * (P + 1);
    const size_t b_tabs = (b_ptrs + (1 + NUM_GRAN) * b_pref + 127) & ~size_t(127);   // tensor maps need 64-byte alignment
    const size_t b_tmaps = v->dense ? sizeof(CUtensorMap) * P : (v->gather ? 2 * sizeof(CUtensorMap) : 0);
    const size_t b_rows = v->gather ? sizeof(int32_t) * seg_rows.size() : 0;   // a multiple of 16 bytes
    const size_t total = b_tabs + b_tmaps + b_rows + sizeof(void*) * P;   // tail: output pointer table (flagg_view_set_outputs)
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    v->last_stream = stream;
    if (pool_ready(v->device) != FLAGG_OK) { delete v; return FLAGG_ECUDA; }
    e = cudaMallocAsync(&v->d_tables, total, st);
    if (e != cudaSuccess) { delete v; set_error("cudaMallocAsync(view tables, %zu B): %s", total, cudaGetErrorString(e)); return FLAGG_ECUDA; }
    char* base = static_cast<char*>(v->d_tables);
    // ONE upload: the tables are assembled in a host staging block laid out like the device block (eight separate
    // pageable copies cost ~30 us of host time per view, a tenth to a third of a whole aggregate() call at the small
    // BASELINE shapes).  Ordering w.r.t. later kernels is by stream.
    const size_t b_up = b_tabs + b_tmaps + b_rows;
    std::vector<char>& stage = v->upload;      // kept with the view: no reliance on when the driver has read it
    stage.assign(b_up, 0);
    memcpy(stage.data(), v->seg_ptrs.data(), b_ptrs);
    memcpy(stage.data() + b_ptrs, v->seg_start.data(), b_pref);
    for (int g = 0; g < NUM_GRAN; ++g) memcpy(stage.data() + b_ptrs + (1 + g) * b_pref, v->unit_start[g].data(), b_pref);
    if (v->dense) memcpy(stage.data() + b_tabs, tmaps.data(), sizeof(CUtensorMap) * P);
    if (v->gather) {
        memcpy(stage.data() + b_tabs, gmap, 2 * sizeof(CUtensorMap));
        memcpy(stage.data() + b_tabs + b_tmaps, seg_rows.data(), b_rows);
    }
    e = cudaMemcpyAsync(base, stage.data(), b_up, cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) {
        set_error("flagg_view_create: table upload failed: %s", cudaGetErrorString(e));
        cudaFreeAsync(v->d_tables, st);
        delete v;
        return FLAGG_ECUDA;
    }
    v->dev.seg_ptrs = reinterpret_cast<const float* const*>(base);
    v->dev.seg_start = reinterpret_cast<const int64_t*>(base + b_ptrs);
    for (int g = 0; g < NUM_GRAN; ++g)
        v->dev.unit_start[g] = reinterpret_cast<const int64_t*>(base + b_ptrs + (1 + g) * b_pref);
    v->dev.tmaps = v->dense ? base + b_tabs : nullptr;
    v->dev.gmap = v->gather ? base + b_tabs : nullptr;
    v->dev.seg_rows = v->gather ? reinterpret_cast<const int32_t*>(base + b_tabs + b_tmaps) : nullptr;
    v->d_out_ptrs = base + b_tabs + b_tmaps + b_rows;
    v->dev.out_ptrs = nullptr;
    v->dev.n = n;
    v->dev.P = P;
    v->dev.d = v->seg_start[P];
    *out = v;
    return FLAGG_OK;
}

int flagg_view_destroy(flagg_view* v) {
    if (!v) return FLAGG_OK;
    cudaStream_t st = static_cast<cudaStream_t>(v->last_stream);
    cudaError_t e1 = v->d_tables ? cudaFreeAsync(v->d_tables, st) : cudaSuccess;
    cudaError_t e2 = v->d_scratch ? cudaFreeAsync(v->d_scratch, st) : cudaSuccess;
    delete v;
    if (e1 != cudaSuccess || e2 != cudaSuccess) {
        set_error("flagg_view_destroy: cudaFreeAsync failed: %s", cudaGetErrorString(e1 != cudaSuccess ? e1 : e2));
        return FLAGG_ECUDA;
    }
    return FLAGG_OK;
}

int flagg_view_set_outputs(flagg_view* v, void* const* out_seg_ptrs, void* stream) {
    FLAGG_REQUIRE(v, "flagg_view_set_outputs: NULL view");
    v->last_stream = stream;
    if (!out_seg_ptrs) {
        v->dev.out_ptrs = nullptr;
        v->out_ptrs.clear();
        return FLAGG_OK;
    }
    const int P = v->dev.P;
    for (int p = 0; p < P; ++p) {
        const uintptr_t a = reinterpret_cast<uintptr_t>(out_seg_ptrs[p]);
        FLAGG_REQUIRE(a != 0 || v->seg_numel[p] == 0, "flagg_view_set_outputs: NULL output tensor %d", p);
        FLAGG_REQUIRE((a & 3) == 0, "flagg_view_set_outputs: output tensor %d not 4-byte aligned", p);
    }
    v->out_ptrs.assign(out_seg_ptrs, out_seg_ptrs + P);
    FLAGG_CUDA(cudaMemcpyAsync(v->d_out_ptrs, v->out_ptrs.data(), sizeof(void*) * P, cudaMemcpyHostToDevice,
                               static_cast<cudaStream_t>(stream)));
    v->dev.out_ptrs = static_cast<float* const*>(v->d_out_ptrs);
    return FLAGG_OK;
}

int64_t flagg_view_cols(const flagg_view* v) { return v ? v->dev.d : -1; }
int flagg_view_rows(const flagg_view* v) { return v ? v->dev.n : -1; }

}  // extern "C"
