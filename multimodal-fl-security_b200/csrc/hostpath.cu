// Host-resident inputs: the Flower server path hands the defense CPU tensors
// (robust_server.py:99-108) and wants CPU tensors back (:125).  Columns are streamed
// through the GPU in chunks, H2D copies of chunk c+1 overlapping the select kernel and
// the D2H copy of chunk c (two internal streams, three staging buffers).  PCIe-bound by
// construction; reported separately from the HBM roofline (bench.py "e2e").
// The staging buffers, streams, events and views are kept between calls (one cached pipe per
// process, rebuilt when n / chunk width / device change, freed by flagg_host_release()):
// a federated round calls this once per round with the same shapes.  Rows that sit at a uniform
// pitch in host memory (one [n x d] matrix) move with ONE 2-D copy per chunk and segment.
#include <algorithm>
#include <cstring>
#include <mutex>
#include <thread>

#include "common.cuh"

using namespace flagg;

namespace {

struct HostPipe {
    static constexpr int NBUF = 3;
    float* stage[NBUF] = {nullptr, nullptr, nullptr};
    float* outb[NBUF] = {nullptr, nullptr, nullptr};
    float* hout[NBUF] = {nullptr, nullptr, nullptr};   // pinned bounce buffers for pageable outputs
    int64_t pend_c0[NBUF] = {0, 0, 0}, pend_w[NBUF] = {0, 0, 0};   // result chunk waiting in hout[b]
    flagg_view* view[NBUF] = {nullptr, nullptr, nullptr};
    cudaEvent_t filled[NBUF] = {}, drained[NBUF] = {};
    cudaStream_t s_copy = nullptr, s_comp = nullptr;
    bool ev_ok = false;
    int n = 0, device = -1, nbuf = 0;
    int64_t chunk_cols = 0;
    ~HostPipe() {
        for (int b = 0; b < NBUF; ++b) {
            if (view[b]) flagg_view_destroy(view[b]);
            if (stage[b]) cudaFree(stage[b]);
            if (outb[b]) cudaFree(outb[b]);
            if (hout[b]) cudaFreeHost(hout[b]);
            if (ev_ok) { cudaEventDestroy(filled[b]); cudaEventDestroy(drained[b]); }
        }
        if (s_copy) cudaStreamDestroy(s_copy);
        if (s_comp) cudaStreamDestroy(s_comp);
    }
};

HostPipe* g_pipe = nullptr;   // cached between calls; intentionally not destroyed at process exit
std::mutex g_pipe_mu;         // guards g_pipe: held for a whole host-path call and by flagg_host_release

int make_dense_view(flagg_view** out, float* base, int n, int64_t ld, int64_t cols, cudaStream_t st) {
    std::vector<const void*> ptrs(n);
    for (int i = 0; i < n; ++i) ptrs[i] = base + size_t(i) * ld;
    return flagg_view_create(out, n, 1, &cols, ptrs.data(), st);
}


// ---- pageable sources (the Flower path: torch.tensor(ndarray) per client per tensor, robust_server.py:99-104) ----
// cudaMemcpyAsync from pageable memory is staged by the driver one call at a time on the calling thread; a
// ResNet-50-sized cohort is n x P = 32,000 such calls.  Instead T worker threads each own two pinned bounce
// buffers and a stream: a worker gathers the next batch (up to UP_BATCH bytes of the row-major virtual [n x d]
// matrix, crossing tensor and client boundaries) into a bounce buffer with plain memcpy while the copy engine
// drains its other buffer.  Host-side gather bandwidth scales with T; the H2D copies are few and large.
// The bounce buffers, streams and events are cached between calls (same lifetime rule as HostPipe).
struct UploadWorker {
    static constexpr int NB = 2;
    char* bounce[NB] = {nullptr, nullptr};
    cudaEvent_t done[NB] = {};
    cudaStream_t s = nullptr;
    bool ok = false;
};
constexpr size_t UP_BATCH = size_t(8) << 20;
constexpr int UP_MAX_THREADS = 8;
struct UploadPool {
    UploadWorker w[UP_MAX_THREADS];
    int device = -1, threads = 0;
    ~UploadPool() {
        for (int t = 0; t < UP_MAX_THREADS; ++t) {
            if (!w[t].ok) continue;
            for (int b = 0; b < UploadWorker::NB; ++b) { cudaFreeHost(w[t].bounce[b]); cudaEventDestroy(w[t].done[b]); }
            cudaStreamDestroy(w[t].s);
        }
    }
};
UploadPool* g_upool = nullptr;   // guarded by g_pipe_mu

int upload_pageable(int n, int P, const std::vector<int64_t>& seg_start, const void* const* host_seg_ptrs, float* dst,
                    int64_t ld, cudaStream_t st) {
    std::lock_guard<std::mutex> lock(g_pipe_mu);
    int device = 0;
    FLAGG_CUDA(cudaGetDevice(&device));
    const int64_t d = seg_start[P];
    const size_t total = sizeof(float) * size_t(n) * size_t(d);
    const int64_t nbatch = int64_t((total + UP_BATCH - 1) / UP_BATCH);
    int hw = int(std::thread::hardware_concurrency());
    const int T = int(std::max<int64_t>(1, std::min<int64_t>(std::min(UP_MAX_THREADS, hw > 2 ? hw - 1 : 1), nbatch)));
    if (g_upool && g_upool->device != device) { delete g_upool; g_upool = nullptr; }
    if (!g_upool) { g_upool = new UploadPool; g_upool->device = device; }
    for (int t = 0; t < T; ++t) {
        UploadWorker& w = g_upool->w[t];
        if (w.ok) continue;
        FLAGG_CUDA(cudaStreamCreateWithFlags(&w.s, cudaStreamNonBlocking));
        for (int b = 0; b < UploadWorker::NB; ++b) {
            FLAGG_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&w.bounce[b]), UP_BATCH, cudaHostAllocDefault));
            FLAGG_CUDA(cudaEventCreateWithFlags(&w.done[b], cudaEventDisableTiming));
        }
        w.ok = true;
    }
    // the destination may still be in use by earlier work on the caller's stream
    cudaEvent_t ready;
    FLAGG_CUDA(cudaEventCreateWithFlags(&ready, cudaEventDisableTiming));
    FLAGG_CUDA(cudaEventRecord(ready, st));
    std::vector<cudaError_t> err(T, cudaSuccess);
    auto work = [&](int t) {
        UploadWorker& w = g_upool->w[t];
        cudaError_t e = cudaSetDevice(device);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(w.s, ready, 0);
        int use = 0;
        for (int64_t bt = t; bt < nbatch && e == cudaSuccess; bt += T, use ^= 1) {
            e = cudaEventSynchronize(w.done[use]);          // the copy that last read this bounce buffer is done
            if (e != cudaSuccess) break;
            // virtual byte range [v0, v1) of the row-major [n x d] fp32 matrix
            const size_t v0 = size_t(bt) * UP_BATCH, v1 = std::min(total, v0 + UP_BATCH);
            size_t v = v0;
            while (v < v1 && e == cudaSuccess) {
                const int64_t row = int64_t(v / (sizeof(float) * size_t(d)));
                const size_t in_row = v - size_t(row) * sizeof(float) * size_t(d);
                const size_t row_end = std::min(v1, (size_t(row) + 1) * sizeof(float) * size_t(d));
                // gather this row's byte range [in_row, in_row + len) from its tensors
                const size_t len = row_end - v;
                size_t o = in_row;
                int p = int(std::upper_bound(seg_start.begin(), seg_start.end(), int64_t(o / sizeof(float))) - seg_start.begin()) - 1;
                char* bp = w.bounce[use] + (v - v0);
                while (o < in_row + len) {
                    const size_t s0 = sizeof(float) * size_t(seg_start[p]), s1 = sizeof(float) * size_t(seg_start[p + 1]);
                    const size_t take = std::min(in_row + len, s1) - o;
                    if (take) memcpy(bp, static_cast<const char*>(host_seg_ptrs[size_t(row) * P + p]) + (o - s0), take);
                    bp += take;
                    o += take;
                    ++p;
                }
                e = cudaMemcpyAsync(reinterpret_cast<char*>(dst + size_t(row) * ld) + in_row, w.bounce[use] + (v - v0), len,
                                    cudaMemcpyHostToDevice, w.s);
                v = row_end;
            }
            if (e == cudaSuccess) e = cudaEventRecord(w.done[use], w.s);
        }
        err[t] = e;
    };
    std::vector<std::thread> th;
    for (int t = 1; t < T; ++t) th.emplace_back(work, t);
    work(0);
    for (auto& x : th) x.join();
    int rc = FLAGG_OK;
    for (int t = 0; t < T; ++t) {
        // the caller's stream continues after every worker stream has drained (stream-ordered, no host wait)
        cudaError_t e = err[t];
        if (e == cudaSuccess) e = cudaEventRecord(ready, g_upool->w[t].s);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(st, ready, 0);
        if (e != cudaSuccess && rc == FLAGG_OK) {
            set_error("flagg_upload_rows: %s", cudaGetErrorString(e));
            rc = FLAGG_ECUDA;
        }
    }
    if (rc != FLAGG_OK) for (int t = 0; t < T; ++t) cudaStreamSynchronize(g_upool->w[t].s);
    cudaEventDestroy(ready);
    return rc;
}

}  // namespace

extern "C" {

int flagg_colwise_select_host(int n, int P, const int64_t* seg_numel, const void* const* host_seg_ptrs,
                              int mode, int k, float* host_out, int64_t chunk_cols) {
    FLAGG_REQUIRE(seg_numel && host_seg_ptrs && host_out, "flagg_colwise_select_host: NULL argument");
    FLAGG_REQUIRE(n >= 1 && P >= 1, "flagg_colwise_select_host: need n >= 1 and P >= 1");
    if (mode == FLAGG_TRIMMED_MEAN)
        FLAGG_REQUIRE(k >= 0 && n - 2 * k >= 1, "flagg_colwise_select_host: trimmed mean needs n-2k >= 1 (n=%d k=%d)", n, k);
    std::vector<int64_t> seg_start(P + 1, 0);
    for (int p = 0; p < P; ++p) {
        FLAGG_REQUIRE(seg_numel[p] >= 0, "flagg_colwise_select_host: negative segment size");
        seg_start[p + 1] = seg_start[p] + seg_numel[p];
    }
    const int64_t d = seg_start[P];
    if (d == 0) return FLAGG_OK;
    if (chunk_cols <= 0) chunk_cols = std::max<int64_t>(128 * 1024, (int64_t(256) << 20) / (4 * int64_t(n)));
    chunk_cols = std::min<int64_t>((chunk_cols + 127) & ~int64_t(127), (d + 127) & ~int64_t(127));
    const int64_t nchunks = (d + chunk_cols - 1) / chunk_cols;

    std::lock_guard<std::mutex> lock(g_pipe_mu);   // one host-path call at a time per process (it owns the copy engines anyway)
    int device = 0;
    FLAGG_CUDA(cudaGetDevice(&device));
    const int nbuf = int(std::min<int64_t>(HostPipe::NBUF, nchunks));
    if (g_pipe && (g_pipe->n != n || g_pipe->chunk_cols != chunk_cols || g_pipe->device != device || g_pipe->nbuf < nbuf)) {
        delete g_pipe;
        g_pipe = nullptr;
    }
    if (!g_pipe) {
        HostPipe* np = new HostPipe;
        np->n = n; np->chunk_cols = chunk_cols; np->device = device; np->nbuf = nbuf;
        auto build = [&]() -> int {
            FLAGG_CUDA(cudaStreamCreateWithFlags(&np->s_copy, cudaStreamNonBlocking));
            FLAGG_CUDA(cudaStreamCreateWithFlags(&np->s_comp, cudaStreamNonBlocking));
            for (int b = 0; b < HostPipe::NBUF; ++b) {
                FLAGG_CUDA(cudaEventCreateWithFlags(&np->filled[b], cudaEventDisableTiming));
                FLAGG_CUDA(cudaEventCreateWithFlags(&np->drained[b], cudaEventDisableTiming));
            }
            np->ev_ok = true;
            for (int b = 0; b < nbuf; ++b) {
                FLAGG_CUDA(cudaMalloc(&np->stage[b], sizeof(float) * size_t(n) * chunk_cols));
                FLAGG_CUDA(cudaMalloc(&np->outb[b], sizeof(float) * size_t(chunk_cols)));
                FLAGG_CUDA(cudaHostAlloc(&np->hout[b], sizeof(float) * size_t(chunk_cols), cudaHostAllocDefault));
                int rc = make_dense_view(&np->view[b], np->stage[b], n, chunk_cols, chunk_cols, np->s_comp);
                if (rc != FLAGG_OK) return rc;
            }
            return FLAGG_OK;
        };
        int rc = build();
        if (rc != FLAGG_OK) { delete np; return rc; }
        g_pipe = np;
    }
    HostPipe& hp = *g_pipe;
    bool out_pinned = false;
    {
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, host_out) == cudaSuccess) out_pinned = (at.type == cudaMemoryTypeHost);
        else (void)cudaGetLastError();
    }
    // per segment: do the n rows sit at one pitch in host memory?
    std::vector<int64_t> pitch(P, 0);
    for (int p = 0; p < P; ++p) {
        if (n < 2) continue;
        const char* b0 = static_cast<const char*>(host_seg_ptrs[p]);
        const int64_t st = static_cast<const char*>(host_seg_ptrs[size_t(P) + p]) - b0;
        bool ok = st >= int64_t(sizeof(float)) * seg_numel[p] && st > 0;
        for (int i = 2; i < n && ok; ++i) ok = (static_cast<const char*>(host_seg_ptrs[size_t(i) * P + p]) - b0) == st * i;
        pitch[p] = ok ? st : 0;
    }
    flagg_view* tail_view = nullptr;  // the last chunk is usually narrower
    // The chunk loop runs inside a lambda so that a failing CUDA call (FLAGG_CUDA returns) still
    // reaches the common exit below: both streams are synchronised (no copy into the cached staging
    // buffers or out of the caller's memory stays in flight) and the tail view is destroyed.
    auto run_chunks = [&]() -> int {
    int rc = FLAGG_OK;
    for (int64_t ch = 0; ch < nchunks && rc == FLAGG_OK; ++ch) {
        const int b = int(ch % nbuf);
        const int64_t c0 = ch * chunk_cols, c1 = std::min(d, c0 + chunk_cols), w = c1 - c0;
        if (ch >= nbuf) {
            FLAGG_CUDA(cudaStreamWaitEvent(hp.s_copy, hp.drained[b], 0));
            if (!out_pinned) {   // the chunk that used this buffer last has landed in hout[b]: hand it to the caller
                FLAGG_CUDA(cudaEventSynchronize(hp.drained[b]));
                memcpy(host_out + hp.pend_c0[b], hp.hout[b], sizeof(float) * size_t(hp.pend_w[b]));
            }
        }
        // first segment overlapping [c0, c1)
        int p0 = int(std::upper_bound(seg_start.begin(), seg_start.end(), c0) - seg_start.begin()) - 1;
        for (int p = p0; p < P && seg_start[p] < c1; ++p) {
            const int64_t a = std::max(c0, seg_start[p]), e = std::min(c1, seg_start[p + 1]);
            if (e <= a) continue;
            if (pitch[p] > 0) {
                const float* src = static_cast<const float*>(host_seg_ptrs[p]) + (a - seg_start[p]);
                FLAGG_CUDA(cudaMemcpy2DAsync(hp.stage[b] + (a - c0), sizeof(float) * size_t(chunk_cols), src, size_t(pitch[p]),
                                             sizeof(float) * size_t(e - a), size_t(n), cudaMemcpyHostToDevice, hp.s_copy));
                continue;
            }
            for (int i = 0; i < n; ++i) {
                const float* src = static_cast<const float*>(host_seg_ptrs[size_t(i) * P + p]) + (a - seg_start[p]);
                FLAGG_CUDA(cudaMemcpyAsync(hp.stage[b] + size_t(i) * chunk_cols + (a - c0), src,
                                           sizeof(float) * size_t(e - a), cudaMemcpyHostToDevice, hp.s_copy));
            }
        }
        FLAGG_CUDA(cudaEventRecord(hp.filled[b], hp.s_copy));
        FLAGG_CUDA(cudaStreamWaitEvent(hp.s_comp, hp.filled[b], 0));
        const flagg_view* use = hp.view[b];
        if (w != chunk_cols) {
            rc = make_dense_view(&tail_view, hp.stage[b], n, chunk_cols, w, hp.s_comp);
            if (rc != FLAGG_OK) break;
            use = tail_view;
        }
        rc = flagg_colwise_select(use, mode, k, hp.outb[b], hp.s_comp);
        if (rc != FLAGG_OK) break;
        // a D2H copy into pageable memory blocks this thread per chunk and stalls the H2D queue: go through pinned bounce buffers
        FLAGG_CUDA(cudaMemcpyAsync(out_pinned ? host_out + c0 : hp.hout[b], hp.outb[b], sizeof(float) * size_t(w), cudaMemcpyDeviceToHost, hp.s_comp));
        hp.pend_c0[b] = c0;
        hp.pend_w[b] = w;
        FLAGG_CUDA(cudaEventRecord(hp.drained[b], hp.s_comp));
    }
    return rc;
    };
    const int rc = run_chunks();
    cudaError_t e1 = cudaStreamSynchronize(hp.s_copy), e2 = cudaStreamSynchronize(hp.s_comp);
    if (rc == FLAGG_OK && e1 == cudaSuccess && e2 == cudaSuccess && !out_pinned)
        for (int64_t ch = std::max<int64_t>(0, nchunks - nbuf); ch < nchunks; ++ch) {
            const int b = int(ch % nbuf);
            memcpy(host_out + hp.pend_c0[b], hp.hout[b], sizeof(float) * size_t(hp.pend_w[b]));
        }
    if (tail_view) flagg_view_destroy(tail_view);
    if (rc != FLAGG_OK) return rc;
    if (e1 != cudaSuccess || e2 != cudaSuccess) {
        set_error("flagg_colwise_select_host: %s", cudaGetErrorString(e1 != cudaSuccess ? e1 : e2));
        return FLAGG_ECUDA;
    }
    return FLAGG_OK;
}

int flagg_host_release(void) {
    std::lock_guard<std::mutex> lock(g_pipe_mu);
    delete g_pipe;
    g_pipe = nullptr;
    delete g_upool;
    g_upool = nullptr;
    return FLAGG_OK;
}

int flagg_upload_rows(int n, int P, const int64_t* seg_numel, const void* const* host_seg_ptrs, float* dst,
                      int64_t ld, void* stream) {
    FLAGG_REQUIRE(seg_numel && host_seg_ptrs && dst, "flagg_upload_rows: NULL argument");
    FLAGG_REQUIRE(n >= 1 && P >= 1, "flagg_upload_rows: need n >= 1 and P >= 1");
    std::vector<int64_t> seg_start(P + 1, 0);
    for (int p = 0; p < P; ++p) {
        FLAGG_REQUIRE(seg_numel[p] >= 0, "flagg_upload_rows: negative segment size");
        seg_start[p + 1] = seg_start[p] + seg_numel[p];
    }
    const int64_t d = seg_start[P];
    FLAGG_REQUIRE(ld >= d, "flagg_upload_rows: ld (%lld) < d (%lld)", (long long)ld, (long long)d);
    if (d == 0) return FLAGG_OK;
    bool all_pinned = true;
    for (size_t i = 0; i < size_t(n) * P; ++i) {
        if (seg_numel[i % P] == 0) continue;
        FLAGG_REQUIRE(host_seg_ptrs[i], "flagg_upload_rows: NULL segment pointer (client %zu, tensor %zu)", i / P, i % P);
        if (all_pinned) {
            cudaPointerAttributes at;
            if (cudaPointerGetAttributes(&at, host_seg_ptrs[i]) != cudaSuccess) { (void)cudaGetLastError(); all_pinned = false; }
            else all_pinned = (at.type == cudaMemoryTypeHost);
        }
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (all_pinned) {
        // Pinned sources: the copy engine reads them directly.  One 2-D copy per tensor when the n rows
        // sit at one pitch (a stacked host matrix), else one copy per (client, tensor).
        for (int p = 0; p < P; ++p) {
            if (seg_numel[p] == 0) continue;
            const char* b0 = static_cast<const char*>(host_seg_ptrs[p]);
            int64_t pitch = n >= 2 ? static_cast<const char*>(host_seg_ptrs[size_t(P) + p]) - b0 : 0;
            bool ok = n >= 2 && pitch >= int64_t(sizeof(float)) * seg_numel[p];
            for (int i = 2; i < n && ok; ++i) ok = (static_cast<const char*>(host_seg_ptrs[size_t(i) * P + p]) - b0) == pitch * i;
            if (ok) {
                FLAGG_CUDA(cudaMemcpy2DAsync(dst + seg_start[p], sizeof(float) * size_t(ld), b0, size_t(pitch),
                                             sizeof(float) * size_t(seg_numel[p]), size_t(n), cudaMemcpyHostToDevice, st));
                continue;
            }
            for (int i = 0; i < n; ++i)
                FLAGG_CUDA(cudaMemcpyAsync(dst + size_t(i) * ld + seg_start[p], host_seg_ptrs[size_t(i) * P + p],
                                           sizeof(float) * size_t(seg_numel[p]), cudaMemcpyHostToDevice, st));
        }
        return FLAGG_OK;
    }
    return upload_pageable(n, P, seg_start, host_seg_ptrs, dst, ld, st);
}

}  // extern "C"
