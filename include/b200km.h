/*
 * b200km.h -- C ABI of libb200km.so: the k-means hot path of dask-ml (Lloyd iteration +
 * k-means|| initialisation) as hand-written sm_100a CUDA kernels.
 *
 * The reference (dask/dask-ml) is pure Python and has NO FFI for this path; its boundary is
 * the sklearn estimator protocol plus a few module-level functions.  Each entry point below
 * names the reference interface it replaces (R/ = the dask-ml source tree) -- that is the
 * call a maintainer would re-route through a ctypes stub.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer owned by the caller unless the name says `host`;
 *   - nothing here allocates or frees device memory; scratch is passed in by the caller and
 *     sized with the *_bytes() queries;
 *   - every launch is asynchronous on `stream` (a cudaStream_t passed as void*);
 *   - return value: 0 = ok, <0 = B200KM_E* (invalid argument), >0 = a cudaError_t;
 *   - X is row-major, `ldx` = elements between consecutive rows (>= d);
 *   - never throws, never prints;
 *   - thread-safe for distinct streams + distinct scratch buffers.
 */
#ifndef B200KM_H_
#define B200KM_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200KM_VERSION 100

/* error codes (negative = argument errors) */
#define B200KM_OK 0
#define B200KM_EINVAL (-1)     /* null pointer / non-positive size / bad dtype            */
#define B200KM_ETOOBIG (-2)    /* k or d beyond what the kernels support                  */
#define B200KM_EALIGN (-3)     /* pointer / ldx not aligned as documented                 */
#define B200KM_ECAPACITY (-4)  /* an output list overflowed its capacity (count is valid) */
#define B200KM_ESTATICROWS (-5) /* b200km_assign_static_f32: n > B200KM_STATIC_MAX_ROWS             */
#define B200KM_ESTATICPACK (-6) /* b200km_assign_static_f32: k, d need more than the built-in pack  */

/* dtype tags */
#define B200KM_F32 0
#define B200KM_F64 1

/* limits */
#define B200KM_MAX_D 2048
#define B200KM_MAX_K (1 << 20)
#define B200KM_MINDIST_MAX_M 8 /* candidates per b200km_mindist_update_* launch */

/* options (process-wide; meant for benchmarks and tests) */
#define B200KM_OPT_K1_VARIANT 1 /* 0 = auto (tcgen05 3xTF32 for d <= 2048: small-k kernel when k <= 256 and d <= 32, panel kernel for d > 128),
                                  * 1 = FP32 FFMA kernel, 2 = the general tcgen05 kernel on one CTA per SM,
                                  * 3 = the general tcgen05 kernel on CTA pairs (cta_group::2).
                                  * Set it BEFORE b200km_pack_centers: the tensor-core image differs for 3. */
#define B200KM_OPT_K2_VARIANT 2 /* 0 = auto, 1 = shared-memory owner table, 2 = chunk sort + run sums, 3 = global atomics */
#define B200KM_OPT_FUSED_VARIANT 3 /* fused K1+K2 (k <= 16, d <= 32): 0 = auto, 1 = the sorted-tile kernel always,
                                     * 2 = same as auto (the cp.async kernel for 16-byte aligned rows of 64 / 128 bytes) */
#define B200KM_OPT_K4_VARIANT 4 /* K4 (b200km_mindist_update_f32): 0 = auto (1-2 candidates: rows streamed through a TMA ring when
                                  * they are contiguous, 128 .. 2048 bytes long and there are enough of them; 3-8 candidates: one
                                  * thread per row), 1 = the register-path kernel always, 2 = the ring wherever it fits (the row
                                  * kernel still takes 3-8 candidates).  Ring and register path: same results bit for bit. */
int b200km_set_option(int key, int value);
int b200km_get_option(int key);

int b200km_version(void);
/* static string for a code returned by any entry point (cudaGetErrorString for >0). */
const char *b200km_error_string(int code);

/* ------------------------------------------------------------------------------------------
 * Centre pack.  The fp64 master copy of the centres (k x d, row-major) is expanded once per
 * Lloyd iteration into the layouts the assign kernels read:
 *   mu = mean of the centres (fp32; the float32 kernels work on x - mu and c - mu) and a zero vector that
 *   stands in for it when the rows are already centred; fp32 row-major copy of c - mu; fp32 column-tiled
 *   copy ([tile of 128][d+1][128], zero padded, last row = squared norms, +inf padded); fp32 and fp64
 *   squared norms, max squared norm; fp64 copy of the master; and, for d <= 2048, the tcgen05 operand
 *   image: per tile of 256 centres one augmented K-step block (||c||^2 split in three TF32 terms) and one
 *   block per 8 features holding the TF32 hi and lo parts of -2(c - mu), in the UMMA no-swizzle K-major
 *   layout.  (With B200KM_OPT_K1_VARIANT = 3 the image is laid out for the CTA-pair kernel: set the option
 *   BEFORE packing.)
 * Replaces: the `Y` ndarray argument that R/dask_ml/metrics/pairwise.py:31-33 embeds in every
 * per-block task, and the per-call `row_norms(Y)` inside sklearn's ArgKmin.
 * ---------------------------------------------------------------------------------------- */
size_t b200km_pack_bytes(int k, int d);
int b200km_pack_centers(const double *centers_f64, int k, int d, void *pack, void *stream);

/* ------------------------------------------------------------------------------------------
 * K1  fused distance + argmin (+ min distance).  The n x k distance matrix never exists.
 * Replaces: dask_ml.metrics.pairwise_distances_argmin_min  (R/dask_ml/metrics/pairwise.py:19-47)
 *           -> sklearn ArgKmin per row block, called from the Lloyd loop
 *           (R/dask_ml/cluster/k_means.py:576-579) and from predict (:247-250).
 *
 *   labels[i]   = argmin_j ||x_i - c_j||^2, lowest j on exact ties (int32)
 *   mind2[i]    = that squared distance (X dtype; nullable)
 *   f32 only: rows whose best and second-best fp32 scores are closer than the fp32 rounding
 *   bound are appended to flag_rows[0..*flag_count) (row indices, int32, unordered) so that
 *   b200km_refine_f64 can re-decide them in fp64 -- the reference does all distance arithmetic
 *   in float64 even for float32 input.  flag_rows needs room for n entries; *flag_count must
 *   be zeroed by the caller (or by b200km_assign_f32 when reset_flags != 0).
 *   Passing flag_rows == NULL disables flagging.
 * ---------------------------------------------------------------------------------------- */
int b200km_assign_f32(const float *X, int64_t n, int d, int64_t ldx, const void *pack, int k,
                      int32_t *labels, float *mind2, int32_t *flag_rows, int32_t *flag_count,
                      int reset_flags, void *stream);
int b200km_assign_f64(const double *X, int64_t n, int d, int64_t ldx, const void *pack, int k,
                      int32_t *labels, double *mind2, void *stream);
/* K1 for the Lloyd ITERATION (float32 X): labels and near-tie flags like b200km_assign_f32, but the per-row winner
 * distance is not required of it.  `score` (n floats, caller-owned) is written AT LEAST for the flagged rows, with
 * a value good enough for b200km_refine_pruned_f64's threshold (||x - mu||^2 + best score, expanded form); every
 * other entry is unspecified (kernels that rebuild the exact distance anyway write it for all rows).  The inertia of
 * the iteration then comes from b200km_update_inertia_* below, which sums ||x - c||^2 in float64 from the rows K2
 * reads anyway.  For k <= 256, d <= 32 this removes the winner's centre-row gather and the re-read of the staged
 * row from K1 -- 2/5 of the shared-memory traffic the small-k kernel was bound by (12.3 -> 8.4 ms at C2); the general
 * tensor-core kernel (d <= 128) drops the same work (a 256-byte gather through L2 per row: ~1.5 % at C3 / C4).
 * Replaces: the same call as b200km_assign_f32 inside the loop, R/dask_ml/cluster/k_means.py:576-579. */
int b200km_assign_lloyd_f32(const float *X, int64_t n, int d, int64_t ldx, const void *pack, int k,
                            int32_t *labels, float *score, int32_t *flag_rows, int32_t *flag_count,
                            int reset_flags, void *stream);
/* float64 X on the tensor cores (k > 16).  The reference's arithmetic is float64 for float64 input too, but the labels
 * only need it where the float32 decision is inside its own rounding bound:
 *   b200km_center_convert_f64   out[i] = fl32(x_i - mu)  (mu = the pack's centring vector; n x d floats, ld = d)
 *   b200km_assign_centred_f32   b200km_assign_f32 on rows that are ALREADY centred (the kernels use mu = 0);
 *                               k > 16 only; fills labels / flag_rows exactly like b200km_assign_f32
 *   b200km_refine_x64_f64       b200km_refine_f64 reading the float64 rows
 *   b200km_rowdist_f64          mind2[i] = sum_f (x_if - c_{labels[i], f})^2 in float64 (the float32 kernels' distances
 *                               are good to ~1e-7 only)
 * Replaces: the same call sites as b200km_assign_f64 (R/dask_ml/metrics/pairwise.py:19-47), 8-10x faster than its
 * DFMA kernel for k > 16. */
int b200km_center_convert_f64(const double *X, int64_t n, int d, int64_t ldx, const void *pack, int k, float *out,
                              void *stream);
int b200km_assign_centred_f32(const float *Xc, int64_t n, int d, int64_t ldx, const void *pack, int k,
                              int32_t *labels, float *mind2, int32_t *flag_rows, int32_t *flag_count,
                              int reset_flags, void *stream);
int b200km_refine_x64_f64(const double *X, int64_t n, int d, int64_t ldx, const void *pack, int k,
                          const int32_t *flag_rows, const int32_t *flag_count, int32_t *labels, double *mind2,
                          void *stream);
int b200km_rowdist_f64(const double *X, int64_t n, int d, int64_t ldx, const void *pack, int k,
                       const int32_t *labels, double *mind2, void *stream);
/* b200km_refine_pruned_f64 for float64 rows: the candidate pass reads Xc (the centred float32 block, ld = d) and
 * dist32 (the distances b200km_assign_centred_f32 wrote for it), the float64 scoring reads X; mind2 (nullable) is
 * updated for the re-decided rows.  Falls back to b200km_refine_x64_f64 under the same conditions. */
int b200km_refine_pruned_x64_f64(const double *X, int64_t n, int d, int64_t ldx, const float *Xc, float *dist32,
                                 const void *pack, int k, int32_t *flag_rows, const int32_t *flag_count,
                                 int32_t *labels, double *mind2, void *stream);
int b200km_refine_pruned_x64_len_f64(const double *X, int64_t n, int d, int64_t ldx, const float *Xc, float *dist32,
                                     const void *pack, int k, int32_t *flag_rows, int64_t flag_len,
                                     const int32_t *flag_count, int32_t *labels, double *mind2, void *stream);
/* fp64 re-decision of the flagged rows of a float32 X (labels / mind2 updated in place).
 * `flag_count` is read on the device: no host synchronisation is needed between the two calls.
 * Replaces: nothing separate -- it is what makes K1's float32 labels equal the reference's
 * float64 ArgKmin labels (R/ SURVEY section 0.1). */
int b200km_refine_f64(const float *X, int64_t n, int d, int64_t ldx, const void *pack, int k,
                      const int32_t *flag_rows, const int32_t *flag_count, int32_t *labels,
                      float *mind2, void *stream);
/* The same re-decision, pruned: a flagged row's float64 winner has a float32 score within the rounding bound of
 * the row's best, so a second, small tensor-core pass over the flagged rows (gathered through flag_rows) lists
 * the columns with score <= (mind2 - ||x||^2) + 2 * bound -- typically 2-3 of k -- and only those are scored in
 * float64 (same arithmetic and tie rule as b200km_refine_f64: identical labels).  A row's list holds as many
 * columns as the scratch allows for the number of rows flagged (decided on the device): at least 6, up to 32;
 * rows with more listed columns (duplicate centres, heavily overlapping clusters) get the full scan.
 *   flag_rows: the buffer b200km_assign_f32 filled; entries from *flag_count on are used as scratch (hence not
 *   const).  It has n entries -- or, in the _len form, flag_len >= n entries: a longer buffer (the Python engine
 *   hands over 4 n) keeps the lists long when many rows are flagged (overlapping clusters: 11 % of the rows of the
 *   C4 shape, where 6-column lists sent most of them to the full scan).  mind2 must hold b200km_assign_f32's
 *   output for these rows (or b200km_assign_lloyd_f32's score).
 *   Falls back to b200km_refine_f64 when mind2 == NULL, n < 4096 or K1 variant 3 is selected, and -- decided
 *   on the device, no host synchronisation -- when more than (flag_len - 1) / 9 rows are flagged.
 * Replaces: nothing in the reference (see b200km_refine_f64). */
int b200km_refine_pruned_f64(const float *X, int64_t n, int d, int64_t ldx, const void *pack, int k,
                             int32_t *flag_rows, const int32_t *flag_count, int32_t *labels,
                             float *mind2, void *stream);
int b200km_refine_pruned_len_f64(const float *X, int64_t n, int d, int64_t ldx, const void *pack, int k,
                                 int32_t *flag_rows, int64_t flag_len, const int32_t *flag_count, int32_t *labels,
                                 float *mind2, void *stream);

/* K1 without caller-provided scratch.  No counterpart in the reference (it has no device memory to
 * manage); it exists for callers that cannot allocate on the device: plugins confined to the buffers
 * their host framework hands them, latency-critical threads that must stay out of cudaMalloc (which
 * synchronises the device), processes whose device-memory budget is already spent.  Same results as b200km_pack_centers + b200km_assign_f32 +
 * b200km_refine_f64 (labels int32, mind2 nullable), but the centre pack and the re-decision list
 * live in a fixed array inside the library -- B200KM_STATIC_MAX_PACK_BYTES + 4 * B200KM_STATIC_MAX_ROWS
 * bytes (6 MiB) of device memory that every GPU loading the library holds until the process exits.
 *   Limits: n <= B200KM_STATIC_MAX_ROWS (else B200KM_ESTATICROWS; label a larger X in row slabs) and
 *   b200km_pack_bytes(k, d) <= B200KM_STATIC_MAX_PACK_BYTES (else B200KM_ESTATICPACK; e.g. k = 1024,
 *   d = 64 fits, k = 8192, d = 128 does not).
 *   It never calls cudaMalloc / cudaFree or any allocating routine.  The FIRST call on each device looks
 *   the array's address up (cudaGetSymbolAddress, that device current): make that call before any
 *   stream capture begins.  All calls on a device share the one array: two calls whose kernels can
 *   overlap on the GPU (different streams, no ordering) race on it -- serialise them. */
#define B200KM_STATIC_MAX_ROWS (1 << 20)
#define B200KM_STATIC_MAX_PACK_BYTES (2 << 20)
int b200km_assign_static_f32(const float *X, int64_t n, int d, int64_t ldx, const double *centers_f64,
                             int k, int32_t *labels, float *mind2, void *stream);

/* Candidate weighting of k-means|| (Bahmani et al., steps 7-8: w_c = number of rows whose closest candidate is c):
 * counts[label] += 1 for every row, float64, accumulated INTO counts (the caller zeroes it; one allreduce over the
 * ranks gives the global weights).  The labels come from b200km_assign_* against the candidate set.  Labels outside
 * [0, k) are ignored.
 * The reference leaves this step out ("scikit-learn doesn't have weighted k-means", R/dask_ml/cluster/k_means.py:
 * 481-487) and re-clusters the candidates unweighted (:503-513); here it is opt-in
 * (dask_ml_b200.cluster.k_means.init_scalable(..., weighted=True)), default off, so reference parity is untouched. */
int b200km_label_counts(const int32_t *labels, int64_t n, int k, double *counts, void *stream);

/* ------------------------------------------------------------------------------------------
 * K2  per-cluster sum and count (+ inertia = sum of mind2).
 * Replaces: _centers_dense (R/dask_ml/cluster/k_means.py:622-632), the chain-sum of the block
 *           partials (:595-597), da.bincount(labels, minlength=k) (:598) and
 *           distances.sum() (:616).
 * Accumulates INTO acc = [k*d sums | k counts | 1 inertia] (all float64, packed so that one
 * NCCL allreduce covers it); the caller zeroes acc.  `mind2` nullable (inertia untouched).
 * `scale_sqrt` != 0 adds sqrt(mind2) instead (the reference's non-squared inertia, :612-616).
 * ---------------------------------------------------------------------------------------- */
size_t b200km_acc_len(int k, int d); /* number of doubles in acc */
/* b200km_update_inertia_*: the same sums and counts, and acc's inertia slot += sum_i ||x_i - centers[labels[i]]||^2
 * computed HERE in float64 from the rows and `centers_f64` (k x d, the centres the labels were decided against) --
 * the squared-distance inertia of R/dask_ml/cluster/k_means.py:616 without a per-row distance vector. */
int b200km_update_f32(const float *X, int64_t n, int d, int64_t ldx, const int32_t *labels,
                      const float *mind2, int k, double *acc, int scale_sqrt, void *stream);
int b200km_update_f64(const double *X, int64_t n, int d, int64_t ldx, const int32_t *labels,
                      const double *mind2, int k, double *acc, int scale_sqrt, void *stream);
int b200km_update_inertia_f32(const float *X, int64_t n, int d, int64_t ldx, const int32_t *labels,
                              const double *centers_f64, int k, double *acc, void *stream);
int b200km_update_inertia_f64(const double *X, int64_t n, int d, int64_t ldx, const int32_t *labels,
                              const double *centers_f64, int k, double *acc, void *stream);

/* K1+K2 fused into ONE pass over X for the HBM-bound regime: k <= 16 and d <= 32 (else B200KM_ETOOBIG).
 * labels / mind2 as K1 (float64 arithmetic, so no re-decision is needed), acc += as K2 (squared inertia).
 * 16-byte aligned rows of exactly 64 or 128 bytes (d = 16 / 32 float32, d = 8 / 16 float64; ldx * sizeof a multiple
 * of 16) take the cp.async / FP64-tensor-core kernel, every other shape the sorted-tile kernel
 * (B200KM_OPT_FUSED_VARIANT = 1 keeps the latter for all); same labels up to exact float64 near-ties, same
 * distances to the last bits.  A row whose distances are all NaN gets label -1 and adds to no sum or count.
 * Replaces: the two passes R/dask_ml/cluster/k_means.py:576-598 makes over X per iteration
 *           (pairwise_distances_argmin_min, then _centers_dense + bincount). */
#define B200KM_FUSED_MAX_K 16
#define B200KM_FUSED_MAX_D 32
int b200km_assign_update_f32(const float *X, int64_t n, int d, int64_t ldx, const double *centers_f64, int k,
                             int32_t *labels, float *mind2, double *acc, void *stream);
int b200km_assign_update_f64(const double *X, int64_t n, int d, int64_t ldx, const double *centers_f64, int k,
                             int32_t *labels, double *mind2, double *acc, void *stream);
/* The whole Lloyd iteration of that regime in ONE launch, for a single GPU (no allreduce between the sums and the
 * division): b200km_assign_update_* and, by the CTA that finishes last, K3 -- centers_new = sums / max(count, 1),
 * shift_inertia[0] = ||centers_f64 - centers_new||_F^2, shift_inertia[1] = the iteration's inertia, and then, after
 * a system-scope fence, shift_inertia[2] = tag.  shift_inertia (3 doubles) may be page-locked host memory (UVA: the
 * GPU writes it directly); a host that passes a fresh tag per call can poll shift_inertia[2] for it and read the
 * shift without synchronising the stream.  `acc` (b200km_acc_len doubles) and `ticket` (one uint32) are scratch
 * that must be ZERO on entry and are zero again on return, so consecutive iterations need no memset between them --
 * and iteration j + 1 (centers_f64 := iteration j's centers_new, its own labels / mind2 / centers_new / shift_inertia
 * buffers) may be queued on the same stream BEFORE the host has read iteration j's result: the host's round trip and
 * the launch latency then sit under a running kernel (what dask_ml_b200/cluster/k_means.py does).
 * Replaces: one trip round the loop body R/dask_ml/cluster/k_means.py:576-605. */
int b200km_lloyd_small_f32(const float *X, int64_t n, int d, int64_t ldx, const double *centers_f64, int k,
                           int32_t *labels, float *mind2, double *acc, unsigned int *ticket, double *centers_new,
                           double *shift_inertia, double tag, void *stream);
int b200km_lloyd_small_f64(const double *X, int64_t n, int d, int64_t ldx, const double *centers_f64, int k,
                           int32_t *labels, double *mind2, double *acc, unsigned int *ticket, double *centers_new,
                           double *shift_inertia, double tag, void *stream);

/* out[0] += sum_i v[i]  (or sum_i sqrt(v[i]) when take_sqrt != 0), float64; caller zeroes out.
 * Replaces: distances.sum() (R/dask_ml/cluster/k_means.py:616) on the non-squared final pass. */
int b200km_reduce_sum_f32(const float *v, int64_t n, int take_sqrt, double *out, void *stream);
int b200km_reduce_sum_f64(const double *v, int64_t n, int take_sqrt, double *out, void *stream);

/* ------------------------------------------------------------------------------------------
 * K3  finalize: new = sums / max(count, 1); shift = ||old - new||_F^2.
 * Replaces: R/dask_ml/cluster/k_means.py:600-605 (maximum(counts,1), divide, squared_norm).
 * centers_new (k x d) and shift_out (1) are float64; centers_old may alias nothing.
 * ---------------------------------------------------------------------------------------- */
int b200km_finalize(const double *acc, int k, int d, const double *centers_old,
                    double *centers_new, double *shift_out, void *stream);

/* ------------------------------------------------------------------------------------------
 * K4  k-means|| running minimum:  mind2[i] = min(mind2[i], min_j ||x_i - cand_j||^2), and
 *     cost_out += sum_i mind2[i] (float64; caller zeroes).  init != 0 ignores the old mind2.
 * Replaces: evaluate_cost (R/dask_ml/cluster/k_means.py:516-519) and the first half of
 *           _sample_points (:533-534); `cand` (m x d, X dtype, m <= B200KM_MINDIST_MAX_M).
 * float32, 16-byte aligned rows: 1-2 candidates stream X through a TMA ring (contiguous rows, >= 1024 rows per SM; the
 * register-path kernel otherwise -- bit-identical results), 3-8 candidates run one thread per row behind cp.async
 * buffers (rows of 128 / 256 / 512 bytes, any n).  A row's result depends on the row and the candidates alone, never
 * on n, the launch shape or where the row sits in X.  B200KM_OPT_K4_VARIANT = 1: the register-path kernel always.
 * ---------------------------------------------------------------------------------------- */
int b200km_mindist_update_f32(const float *X, int64_t n, int d, int64_t ldx, const float *cand,
                              int m, float *mind2, int init, double *cost_out, void *stream);
int b200km_mindist_update_f64(const double *X, int64_t n, int d, int64_t ldx, const double *cand,
                              int m, double *mind2, int init, double *cost_out, void *stream);
/* K4 against up to 128 new candidates in ONE pass, on the tensor cores (float32, d <= 128): the candidates, packed by
 * b200km_pack_centers (one-CTA layout), are the M side of K7's 3xTF32 kernel (b200km_cdist_pack_f32), whose epilogue
 * here keeps only the minimum over the candidates and merges it into mind2; distances in the centred expanded form
 * (error ~2^-21 of the data's spread squared).  The FFMA kernel above takes 8 candidates per launch; this one costs the
 * same whatever their number (11 ms per 6e7 x 64 rows: the 128-lane MMA is the bound) and wins from 25 candidates on. */
int b200km_mindist_tc_f32(const float *X, int64_t n, int d, int64_t ldx, const void *pack, int m, float *mind2,
                          int init, double *cost_out, void *stream);
/* The same update when the distances to the new candidates come from somewhere else (K1 against MANY candidates at
 * once -- an oversampling factor in the thousands -- instead of m / 8 passes of the kernel above):
 * mind2[i] = min(mind2[i], other[i]) (init != 0: mind2[i] = other[i]); cost_out += sum_i mind2[i]. */
int b200km_mindist_merge_f32(const float *other, int64_t n, float *mind2, int init, double *cost_out, void *stream);
int b200km_mindist_merge_f64(const double *other, int64_t n, double *mind2, int init, double *cost_out, void *stream);
/* k-means++ trial potentials: pots[j] += sum_i min(mind2[i], ||x_i - cand_j||^2) for m <= B200KM_MINDIST_MAX_M
 * candidates, mind2 untouched (pots must be zeroed by the caller).  Replaces, per k-means++ step, the
 * (trials x n) distance matrix + np.minimum + potential sums of scikit-learn's _kmeans_plusplus, which the
 * reference's init_pp calls on the fully materialised X (R/dask_ml/cluster/k_means.py:404-423). */
int b200km_mindist_pots_f32(const float *X, int64_t n, int d, int64_t ldx, const float *cand, int m,
                            const float *mind2, double *pots_out, void *stream);
int b200km_mindist_pots_f64(const double *X, int64_t n, int d, int64_t ldx, const double *cand, int m,
                            const double *mind2, double *pots_out, void *stream);

/* ------------------------------------------------------------------------------------------
 * K5  Bernoulli oversampling: row i is picked when  scale * mind2[i] > u_i,  scale = l / cost.
 *     u_i = Philox4x32-10(key = seed, counter = (row_offset + i, round)) -> 53-bit U[0,1), or
 *     draws[i] when `draws` is non-null (injected draws, for parity tests).
 *     Global indices (row_offset + i, int64) go to picked[0..min(*count, capacity)); *count
 *     (caller zeroes) always holds the true number picked.
 * Replaces: second half of _sample_points (R/dask_ml/cluster/k_means.py:535-541):
 *           random_state.uniform + `p > draws` + da.where.
 * ---------------------------------------------------------------------------------------- */
int b200km_bernoulli_pick_f32(const float *mind2, int64_t n, double scale, uint64_t seed,
                              uint32_t round, int64_t row_offset, const double *draws,
                              int64_t *picked, int32_t *count, int capacity, void *stream);
int b200km_bernoulli_pick_f64(const double *mind2, int64_t n, double scale, uint64_t seed,
                              uint32_t round, int64_t row_offset, const double *draws,
                              int64_t *picked, int32_t *count, int capacity, void *stream);

/* ------------------------------------------------------------------------------------------
 * K6  gather rows by LOCAL index: out[j,:] = X[idx[j],:]  (idx int64, 0 <= idx < n).
 * Replaces: X[sorted(c_idx)].compute() (R/dask_ml/cluster/k_means.py:479), X[locs] (:500),
 *           X[idx] (:431, :452).
 * ---------------------------------------------------------------------------------------- */
int b200km_gather_rows_f32(const float *X, int64_t n, int d, int64_t ldx, const int64_t *idx,
                           int m, float *out, void *stream);
int b200km_gather_rows_f64(const double *X, int64_t n, int d, int64_t ldx, const int64_t *idx,
                           int m, double *out, void *stream);

/* ------------------------------------------------------------------------------------------
 * K7  full distance matrix in the INPUT precision: out[i,j] = sqrt(max(xx_i - 2 x_i.y_j + yy_j, 0))
 *     (sqrt skipped when squared != 0).  xnorm2 / ynorm2 nullable (computed when null).
 * Replaces: dask_ml.metrics.euclidean_distances (R/dask_ml/metrics/pairwise.py:70-105), used
 *           by KMeans.transform (R/dask_ml/cluster/k_means.py:225-228).
 * ---------------------------------------------------------------------------------------- */
int b200km_cdist_f32(const float *X, int64_t n, int d, int64_t ldx, const float *Y, int k,
                     const float *xnorm2, const float *ynorm2, float *out, int squared, void *stream);
int b200km_cdist_f64(const double *X, int64_t n, int d, int64_t ldx, const double *Y, int k,
                     const double *xnorm2, const double *ynorm2, double *out, int squared, void *stream);

/* K7 on the tensor cores, for float32 rows of at most 128 features (else B200KM_ETOOBIG: use the entry points
 * above): out[i * ldo + j] = f(||x_i - c_j||^2), c_j the centres of a pack built by b200km_pack_centers (one-CTA
 * layout, i.e. not under B200KM_OPT_K1_VARIANT = 3), f by `mode`.  The op is bound by its output -- k floats per
 * row -- so the centres are the M side of the 3xTF32 tcgen05 MMA and 256 rows (128 beyond 64 features) the N side: a warp of the epilogue
 * holds 32 neighbouring centres of one row per register and every store writes 128 contiguous bytes.  Both squared
 * norms ride in an augmented K-step; rows and centres are centred on the pack's mu, so the cancellation error
 * scales with the spread of the data, not its distance from the origin (more accurate than the float32 expanded
 * form the reference evaluates, R/dask_ml/metrics/pairwise.py:101-105).
 * Replaces: euclidean_distances (R/dask_ml/metrics/pairwise.py:70-105) = KMeans.transform
 *           (R/dask_ml/cluster/k_means.py:225-228), pairwise_distances (:50-67), rbf_kernel (:141-151). */
#define B200KM_CDIST_SQUARED 0 /* ||x - c||^2                 */
#define B200KM_CDIST_DIST 1    /* ||x - c||                   */
#define B200KM_CDIST_RBF 2     /* exp(-gamma * ||x - c||^2)   */
int b200km_cdist_pack_f32(const float *X, int64_t n, int d, int64_t ldx, const void *pack, int k, float *out,
                          int64_t ldo, int mode, double gamma, void *stream);

/* Kernel matrices on the same tile kernel: out[i,j] = kernel(x_i, y_j) in the input precision.
 * Replaces: dask_ml.metrics.pairwise.{rbf,linear,polynomial,sigmoid}_kernel / pairwise_kernels
 *           (R/dask_ml/metrics/pairwise.py:135-226), the affinity blocks of SpectralClustering's Nystrom
 *           embedding (R/dask_ml/cluster/spectral.py:312-336) -- the only in-tree caller of KMeans.
 *   kind: RBF exp(-gamma ||x-y||^2), LINEAR x.y, POLY (gamma x.y + coef0)^degree, SIGMOID tanh(gamma x.y + coef0) */
#define B200KM_KERNEL_RBF 2
#define B200KM_KERNEL_LINEAR 3
#define B200KM_KERNEL_POLY 4
#define B200KM_KERNEL_SIGMOID 5
int b200km_pairwise_kernel_f32(const float *X, int64_t n, int d, int64_t ldx, const float *Y, int k, float *out,
                               int kind, double gamma, double coef0, int degree, void *stream);
int b200km_pairwise_kernel_f64(const double *X, int64_t n, int d, int64_t ldx, const double *Y, int k, double *out,
                               int kind, double gamma, double coef0, int degree, void *stream);

/* row_norms(X, squared) -- R/dask_ml/utils.py:83-88. */
int b200km_row_norms_f32(const float *X, int64_t n, int d, int64_t ldx, float *out, int squared, void *stream);
int b200km_row_norms_f64(const double *X, int64_t n, int d, int64_t ldx, double *out, int squared, void *stream);

/* *bad (int32, caller zeroes) becomes non-zero when X holds a NaN or +-inf.
 * Replaces: the da.isnull / da.isinf scan in KMeans._check_array (R/dask_ml/cluster/k_means.py:195-201). */
int b200km_count_nonfinite_f32(const float *X, int64_t n, int d, int64_t ldx, int32_t *bad, void *stream);
int b200km_count_nonfinite_f64(const double *X, int64_t n, int d, int64_t ldx, int32_t *bad, void *stream);

/* ------------------------------------------------------------------------------------------
 * Synthetic blobs on the device (benchmark input):
 *   X[i,:] = centers[class(row_offset+i), :] + cluster_std * N(0,1)   (Philox + Box-Muller; the class of a
 *   row is a Philox hash of its GLOBAL index, so any sharding generates the same array); y (nullable,
 *   int32) receives the class of every row.
 * Mirrors the distribution of dask_ml.datasets.make_blobs (R/dask_ml/datasets.py:78-205): equal
 * points per centre, isotropic Gaussian; NOT its bit stream.
 * ---------------------------------------------------------------------------------------- */
int b200km_make_blobs_f32(float *X, int64_t n, int d, int64_t ldx, const float *centers, int k_true,
                          float cluster_std, uint64_t seed, int64_t row_offset, int32_t *y, void *stream);
int b200km_make_blobs_f64(double *X, int64_t n, int d, int64_t ldx, const double *centers, int k_true,
                          double cluster_std, uint64_t seed, int64_t row_offset, int32_t *y, void *stream);

/* FP32 FFMA / FP64 DFMA pipe microbenchmarks (roofline denominators that MEASURED_PEAKS.json
 * lacks).  Runs `iters` dependent-chain FMA bundles on every SM; writes a checksum to `sink`
 * (>= 1 float/double).  flops = 2 * 64 * iters * blocks * threads; the caller times it. */
int b200km_probe_ffma(float *sink, int blocks, int threads, int iters, void *stream);
int b200km_probe_dfma(double *sink, int blocks, int threads, int iters, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* B200KM_H_ */
