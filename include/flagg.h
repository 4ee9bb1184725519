/*
 * flagg.h -- C ABI of libflagg.so, the B200 (sm_100a) robust-aggregation engine.
 *
 * This is the drop-in boundary for the server-side defense hot path of
 * Shashank8834/multimodal-fl-security (reference paths below are relative to that
 * repository).  The reference has no native code: its defenses are Python classes
 * with the plugin interface
 *
 *     BaseDefense.aggregate(client_updates: List[List[Tensor]], num_examples)
 *                                              src/defenses/base_defense.py:31-47
 *
 * The Python mirror of those classes (multimodal-fl-security_b200/src/defenses) binds
 * the entry points below with ctypes; INTEGRATION.md shows the stub a maintainer of
 * the reference would add.  No torch types cross this boundary: plain pointers,
 * sizes and a cudaStream_t passed as void*.
 *
 * Conventions
 *   - All "device" pointers are CUDA device addresses owned by the caller (the
 *     PyTorch caching allocator in the Python mirror).  All work is enqueued on
 *     `stream` (NULL = legacy default stream) and is asynchronous unless noted.
 *   - Every function returns FLAGG_OK (0) or a negative flagg_status;
 *     flagg_last_error() returns a thread-local description of the last failure.
 *   - There is no CPU fallback anywhere in this library.
 *
 * The client-update "matrix" X[n, d] (n clients, d parameters) is never packed:
 * a flagg_view describes it as n rows x P segments, segment p of client i being a
 * contiguous fp32 device array of seg_numel[p] elements -- exactly the
 * List[List[Tensor]] the reference passes around (krum.py:55-57 flattens it,
 * trimmed_mean.py:79 stacks it; both copies are avoided here).
 */
#ifndef FLAGG_H
#define FLAGG_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FLAGG_VERSION 100 /* 0.1.0 */

typedef enum flagg_status {
    FLAGG_OK = 0,
    FLAGG_EINVAL = -1,       /* bad argument */
    FLAGG_ECUDA = -2,        /* CUDA runtime error (text in flagg_last_error) */
    FLAGG_ENOMEM = -3,
    FLAGG_EUNSUPPORTED = -4  /* valid request this build cannot serve */
} flagg_status;

typedef enum flagg_select_mode {
    FLAGG_MEDIAN = 0,        /* rank (n-1)/2 under (value, client index) order   */
    FLAGG_TRIMMED_MEAN = 1   /* mean of sorted ranks [k, n-k)                     */
} flagg_select_mode;

typedef enum flagg_reduce_op { FLAGG_SUM = 0, FLAGG_MAX = 1 } flagg_reduce_op;

/* Elementwise row transforms (the reference's model-poisoning attacks, src/attacks/model_poisoning.py). */
typedef enum flagg_transform_op {
    FLAGG_T_SCALE = 0,    /* out = x * s                 ScalingAttack :236; IPM negation (s = -1) :273 */
    FLAGG_T_REPLACE = 1,  /* out = y + (x - y) * s       ModelReplacementAttack :74-77 (x = local, y = global) */
    FLAGG_T_SIGN = 2,     /* out = s * sign(x)           InnerProductManipulationAttack :279-281 (s = -epsilon) */
    FLAGG_T_AXPY = 3      /* out = x + y * s             AdaptiveKrumAttack :172-184 (x = centre, y = direction) */
} flagg_transform_op;

typedef struct flagg_view flagg_view;
typedef struct flagg_comm flagg_comm;
#define FLAGG_UNIQUE_ID_BYTES 128

/* ---- library ---------------------------------------------------------------- */
int flagg_version(void);
const char* flagg_last_error(void);
/* SM count and compute capability of the current device. */
int flagg_device_info(int* sm_count, int* cc_major, int* cc_minor);

/* ---- views ------------------------------------------------------------------ */
/* Describe n client updates of P tensors each.  seg_numel: host array [P];
 * seg_ptrs: HOST array [n*P] of DEVICE addresses, seg_ptrs[i*P+p] = client i,
 * tensor p (fp32, at least 4-byte aligned; 16-byte alignment enables 128-bit
 * loads).  The small pointer tables are uploaded on `stream`; the data is not
 * touched.  Replaces: torch.cat/torch.stack packing at krum.py:57,
 * trimmed_mean.py:79,160, fltrust.py:91, differential_privacy.py:58,225,311. */
int flagg_view_create(flagg_view** out, int n, int P, const int64_t* seg_numel,
                      const void* const* seg_ptrs, void* stream);
int flagg_view_destroy(flagg_view* v);
/* Round epilogue (experiments/run_experiments.py:255-259: `param.copy_(new_val)` per tensor
 * after aggregate): out_seg_ptrs = HOST array [P] of DEVICE addresses, tensor p of every later
 * [d] result of this view (flagg_colwise_select, flagg_row_combine*) is written to
 * out_seg_ptrs[p] -- e.g. the global model's own parameter tensors -- instead of the flat
 * `out` argument, which may then be NULL.  NULL clears the table. */
int flagg_view_set_outputs(flagg_view* v, void* const* out_seg_ptrs, void* stream);
int64_t flagg_view_cols(const flagg_view* v); /* d = sum(seg_numel) */
int flagg_view_rows(const flagg_view* v);     /* n */

/* ---- K1: column-wise order statistics -------------------------------------- */
/* out[d] (device, fp32, segments concatenated) = per-coordinate median or trimmed
 * mean over the n clients.
 *   FLAGG_MEDIAN        replaces torch.median(stacked, dim=0)[0]
 *                       (trimmed_mean.py:101,163): bit-exact, including the
 *                       (value, index) tie rule for +-0 and "any NaN -> first NaN".
 *   FLAGG_TRIMMED_MEAN  replaces torch.sort + slice [k:n-k] + mean(dim=0)
 *                       (trimmed_mean.py:82-88); requires 1 <= n-2k; NaN sorts last.
 *                       The kept values are summed in fp32 from the middle ranks outwards
 *                       and multiplied by fp32(1/(n-2k)) (error <= n/2 ulp of mean|kept|,
 *                       inside the 1e-5 tolerance; Tensor.mean's own order is not
 *                       reproducible -- DESIGN.md 3.1). */
int flagg_colwise_select(const flagg_view* v, int mode, int k, float* out, void* stream);

/* ---- K2: streaming row reductions ------------------------------------------ */
/* One pass over X producing, per client i (device outputs, any may be NULL):
 *   sumsq[i] = sum_c x_ic^2 (FP64)   -- torch.norm at krum.py:95 (after centring),
 *                                       fltrust.py:175,208, differential_privacy.py:88,228,311
 *   dot[i]   = sum_c x_ic*g0_c (FP64)-- torch.dot at fltrust.py:174 (needs g0)
 *   linf[i]  = max_c |x_ic|          -- differential_privacy.py:230
 *   g0_sumsq = sum_c g0_c^2          -- fltrust.py:176,209
 * g0: device fp32 [d] (segments concatenated) or NULL.  centered != 0: the per-client
 * statistics are those of (x_i - g0) instead of x_i -- ||x_i - g0||^2 is the Weiszfeld
 * distance of trimmed_mean.py:245.  Deterministic (fixed reduction order). */
int flagg_row_reduce(const flagg_view* v, const float* g0, int centered, double* sumsq,
                     double* dot, float* linf, double* g0_sumsq, void* stream);

/* flagg_row_reduce for callers that may not allocate (CUDA graph capture, serving loops, harnesses that forbid
 * cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same arguments, same
 * kernels, same grid and therefore the same results bit for bit; the per-CTA partial sums live in a file-scope
 * `static __device__` array of FLAGG_STATIC_SCRATCH_BYTES (every GPU that loads libflagg.so holds it until exit)
 * instead of the view's scratch block, and the call makes no cudaMalloc / cudaMallocAsync / cudaFree.  The view is
 * an input: create it beforehand, on the device that is current at this call.
 * Limit: the partials take 20 n + 8 bytes per CTA, min(ceil-units of 2048 columns, 4 x SM count) CTAs -- n <= 708
 * clients on a 148-SM B200 at any d; beyond the array the call returns FLAGG_EUNSUPPORTED and launches nothing.
 * Every call on a device shares the one array: calls that may overlap on the GPU (different streams, no ordering
 * between them) race on it -- serialise them.  Nothing in it needs clearing between calls. */
#define FLAGG_STATIC_SCRATCH_BYTES (8u << 20)
int flagg_row_reduce_static(const flagg_view* v, const float* g0, int centered, double* sumsq,
                            double* dot, float* linf, double* g0_sumsq, void* stream);

/* ---- K3: weighted row combination ------------------------------------------ */
/* out[c] = ( ...((0 + t_0) + t_1) ... + t_{m-1}) / divisor,
 *   t_j = w2[j] * (w1[j] * x[rows[j]][c])
 * with every multiply, add and the final divide rounded to fp32 separately and in
 * this order -- the arithmetic of Python's sum() over scalar*tensor products at
 * base_defense.py:91-95, krum.py:187-190, fltrust.py:262-268,
 * differential_privacy.py:148-155,262-269,325-332 -- so results are bit-identical to
 * the reference's CPU path.  rows/w1/w2: HOST arrays [m] (NULL rows = 0..m-1 with
 * m == n; NULL weight = 1.0f).  out: device fp32 [d]. */
int flagg_row_combine(const flagg_view* v, int m, const int32_t* rows, const float* w1,
                      const float* w2, float divisor, float* out, void* stream);

/* Same combination with every operand resident on the DEVICE (no host round trip between the
 * reductions that produce the weights and this pass): d_rows [m] int32 or NULL (negative entries are skipped), d_w1 / d_w2 [m]
 * fp32 or NULL (= 1.0f), d_divisor one fp32.  With `prev` (device fp32 [d]) and `delta_sumsq`
 * (device, one double) it also returns ||out - prev||^2 (fp32 differences and squares, FP64 sum
 * in a fixed order): the Weiszfeld stopping rule torch.norm(new - current) at trimmed_mean.py:255. */
int flagg_row_combine_dev(const flagg_view* v, int m, const int32_t* d_rows, const float* d_w1,
                          const float* d_w2, const float* d_divisor, float* out, const float* prev,
                          double* delta_sumsq, void* stream);

/* Per-round weights on the device: the reference's Python arithmetic on the flagg_row_reduce statistics, so that
 * a clipping / trust round is K2 -> (all-reduce) -> this -> flagg_row_combine_dev with no host round trip.
 *   FLAGG_W_FLTRUST    trust_i = max(0, fp32 dot_i / (||x_i|| ||g0||)), scale_i = ||g0|| / ||x_i||  (fltrust.py:170-186,
 *                      205-216); divisor = fp32(sum trust, Python order); counts[1] = 1 when sum trust < 1e-10 (:258-260)
 *   FLAGG_W_CLIP       scale_i = fp32(p0 / norm_i) if norm_i > p0 (differential_privacy.py:242-245; norms: L2 from
 *                      sumsq, or Linf when sumsq == NULL and linf is given); counts[1] = clients clipped
 *   FLAGG_W_DPSGD      scale_i = (1 / norm_i) * fp32(p0) in fp32 if norm_i > fp32(p0) (:90-92)
 *   FLAGG_W_NORMBOUND  keep client i iff p1 <= norm_i <= p0 (:313); all rejected -> all kept, counts[1] = 1 (:319-322)
 * num_examples: device FP64 [n] holding the integer counts (w2 = fp32(count) for the non-FLTrust rules; divisor =
 * fp32 of their exact sum over the kept clients, as Python's int sum gives it).
 * Outputs (device): rows [n] int32 = kept clients in order, then -1 (flagg_row_combine_dev skips negative entries:
 * pass m = n); w1, w2 [n] fp32 aligned with rows; divisor [1] fp32; info [n] FP64 (trust score / norm per client);
 * counts [2] int32 (kept, flag); scratch: n bytes. */
typedef enum flagg_weight_rule { FLAGG_W_FLTRUST = 0, FLAGG_W_CLIP = 1, FLAGG_W_DPSGD = 2, FLAGG_W_NORMBOUND = 3 } flagg_weight_rule;
int flagg_round_weights(int rule, int n, const double* sumsq, const double* dot, const double* g0_sumsq, const float* linf,
                        const double* num_examples, double p0, double p1, int32_t* rows, float* w1, float* w2, float* divisor,
                        double* info, int32_t* counts, void* scratch_n_bytes, void* stream);

/* Weiszfeld weights on the device (trimmed_mean.py:245-251): dist_i = max(fp32 sqrt(sumsq_i), 1e-10),
 * w_i = 1 / dist_i, *wsum = sum_i w_i (FP64 sum rounded once).  sumsq: device FP64 [n] (the centred
 * flagg_row_reduce output, all-reduced when sharded); w: device fp32 [n]; wsum: device fp32 [1]. */
int flagg_weiszfeld_weights(const double* sumsq, int n, float* w, float* wsum, void* stream);

/* ---- K4/K5: Krum ------------------------------------------------------------ */
/* G (device, FP64, n*n row-major, symmetric) = (X - c)(X - c)^T where c = `center`
 * (device fp32 [d], segments concatenated; NULL = no centring).  Pairwise distances
 * are translation invariant; centring on a vector inside the benign cluster removes
 * the common component of full-weight rows so fp32 products keep 1e-5 on the
 * distances (DESIGN.md "Gram precision").  Replaces the O(n^2) loop of
 * torch.norm(fi - fj) at krum.py:92-97.  With the parameter dimension sharded,
 * partial G's add: all-reduce them before flagg_krum_scores.
 * `accumulate` != 0 adds into G instead of overwriting it. */
int flagg_gram(const flagg_view* v, const float* center, int accumulate, double* G, void* stream);

/* Small cohorts (n <= 32: the reference's default 10-20 clients): D2 (device, FP64, n*n row-major,
 * symmetric, zero diagonal) = squared pairwise distances ||x_i - x_j||^2 by DIRECT differences, as
 * krum.py:95 forms them (torch.norm(f_i - f_j)): one pass over X, no centre vector and therefore no
 * centre pass, no cancellation; one fp32 subtract + fma per pair and column, fp32 partial sums
 * of at most 1024 columns folded into FP64.  Non-finite rows give NaN / inf entries by IEEE
 * arithmetic, exactly where the reference's distances are NaN / inf.  With the parameter dimension
 * sharded, partial D2's add: all-reduce them before flagg_krum_scores_from_sqdist.
 * `accumulate` != 0 adds into D2.  Returns FLAGG_EUNSUPPORTED for n > 32 (use flagg_gram). */
int flagg_sqdist(const flagg_view* v, int accumulate, double* D2, void* stream);
/* flagg_krum_scores with the squared distances of flagg_sqdist as input (D = sqrt(D2), may be NULL). */
int flagg_krum_scores_from_sqdist(const double* D2, int n, int f, double* D, double* scores, void* stream);

/* D[i,j] = sqrt(max(0, G_ii + G_jj - 2 G_ij)) (un-squared, as krum.py:95), D_ii = 0;
 * scores[i] = sum of the (n-f-2) smallest D[i,j], j != i (krum.py:122-129).
 * D may be NULL.  All device FP64.  Requires 1 <= n-f-2 <= n-1 and n <= 4096.
 * Non-finite entries are propagated, not clamped: a client that sent NaN / Inf gets NaN / inf
 * distances (as torch.norm gives at krum.py:95), NaN sorts last (np.sort, krum.py:126) and stays in
 * the sum, so that client's score is NaN / inf and np.argsort (krum.py:174) rejects it. */
int flagg_krum_scores(const double* G, int n, int f, double* D, double* scores, void* stream);
/* Scores from a distance matrix the caller assembled (device FP64 [n*n], D_ii = 0): the exact path
 * for cohorts with non-finite rows, whose distances come from flagg_row_reduce(centered) instead of
 * the Gram matrix. */
int flagg_krum_scores_from_distances(const double* D, int n, int f, double* scores, void* stream);

/* ---- row transforms: attacks on resident rows ------------------ */
/* One client's P tensors (seg_numel: host [P]; x_ptrs / y_ptrs / out_ptrs: HOST arrays [P] of DEVICE
 * fp32 addresses; y_ptrs may be NULL for SCALE / SIGN; out_ptrs[p] may equal x_ptrs[p]: in place on a
 * row of a resident update matrix).  Every multiply / add / subtract is rounded to fp32 separately in
 * the reference's order, so results are bit-identical to its CPU path. */
int flagg_segments_transform(int P, const int64_t* seg_numel, const void* const* x_ptrs,
                             const void* const* y_ptrs, void* const* out_ptrs, int op, float s,
                             void* stream);

/* ---- host-resident inputs (Flower path, robust_server.py:99-108) ------------ */
/* Same contract as flagg_colwise_select, but seg_ptrs are HOST addresses (pinned
 * memory gives full PCIe rate) and out is a HOST buffer [d].  Columns are streamed
 * through the GPU in chunks of `chunk_cols` (0 = default) with copies and kernels
 * overlapped on internal streams; returns after the result is in `out`. */
int flagg_colwise_select_host(int n, int P, const int64_t* seg_numel,
                              const void* const* host_seg_ptrs, int mode, int k,
                              float* host_out, int64_t chunk_cols);

/* The host path keeps its device staging buffers, streams and events between calls (a
 * federated round repeats the same shapes); this frees them.  Always returns FLAGG_OK. */
int flagg_host_release(void);

/* Copy n host rows (P segments each) into one dense device matrix dst[n, ld]
 * (ld >= d, fp32); work after the call on `stream` sees the data.  Used to make host inputs
 * resident for the two-pass defenses (Krum, FLTrust, clipping; robust_server.py:99-108).
 * Pinned sources are read by the copy engine directly (one 2-D copy per tensor when the rows sit at
 * one pitch); pageable sources (torch.tensor(ndarray), the Flower layout) are gathered by worker
 * threads into cached pinned bounce buffers, 8 MB per H2D copy, gather and DMA overlapped.  The
 * call returns once every source byte has been read (sources may be freed), not once it has landed. */
int flagg_upload_rows(int n, int P, const int64_t* seg_numel,
                      const void* const* host_seg_ptrs, float* dst, int64_t ld, void* stream);

/* ---- parameter-sharded aggregation over several GPUs (one process per GPU) ---- */
/* With the parameter dimension sharded, median / trimmed mean / FedAvg / weighted combine need no
 * communication.  Krum all-reduces its partial n x n FP64 Gram (flagg_gram on the rank's columns,
 * krum.py:73-99), FLTrust / clipping / norm bounding all-reduce the 2n+1 FP64 partial sums of
 * flagg_row_reduce (SUM) and its n fp32 maxima (MAX) (fltrust.py:170-176,
 * differential_privacy.py:225-230, 311); every rank then derives identical scores, selections and
 * weights.  NCCL over NVLink; rank 0 creates the 128-byte id, the host program distributes it.
 * The communicator is bound to the device current at flagg_comm_create; calls are stream-ordered. */
int flagg_comm_unique_id(void* id128);
int flagg_comm_create(flagg_comm** out, int nranks, int rank, const void* id128);
int flagg_comm_destroy(flagg_comm* c);
int flagg_comm_size(const flagg_comm* c);
int flagg_comm_rank(const flagg_comm* c);
int flagg_allreduce_f64(flagg_comm* c, double* buf, size_t count, int op, void* stream);
int flagg_allreduce_f32(flagg_comm* c, float* buf, size_t count, int op, void* stream);
/* Ingest (robust_server.py:99-108 hands the server ONE process's list of whole client updates): rank
 * `root` holds n rows x P tensors on its device (root_seg_ptrs: HOST array [n*P] of DEVICE addresses,
 * NULL elsewhere); rank r receives columns [lo[r], hi[r]) of every row (flattened-vector coordinates,
 * host arrays [nranks], identical on all ranks) as one dense device matrix dst[n, ld].  ncclSend /
 * ncclRecv per (row, peer, tensor piece) over NVLink, stream ordered; the root's own shard is copied
 * device-to-device when dst != NULL (NULL: the root keeps using views of its rows). */
int flagg_comm_scatter_columns(flagg_comm* c, int root, int n, int P, const int64_t* seg_numel,
                               const void* const* root_seg_ptrs, const int64_t* lo, const int64_t* hi,
                               float* dst, int64_t ld, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* FLAGG_H */
