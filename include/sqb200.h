/*
 * sqb200 — C ABI of the B200 (sm_100a) kernels behind sq-learn's quantum
 * simulation path.
 *
 * The reference (federicomegler/sq-learn) has no FFI seam on this path: the
 * estimators call sklearn/QuantumUtility/Utility.py by plain Python import
 * (sklearn/cluster/_dmeans.py:40, sklearn/decomposition/_qPCA.py:27).  The
 * boundary a maintainer binds is therefore this library, loaded with ctypes
 * from the functions that keep the reference's Python names (INTEGRATION.md
 * shows the stubs).  Each entry below cites the reference code it replaces.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless its name ends in _host;
 *   - matrices are row-major (C order) fp64, labels int32, counts fp64
 *     (exact integers < 2^53) so they share one all-reduce buffer;
 *   - `stream` is a cudaStream_t passed as void*; calls are asynchronous on
 *     it and never synchronise;
 *   - return value 0 = ok, otherwise see sq_status and sq_last_error();
 *   - nothing is allocated behind the caller's back: scratch comes in through
 *     (ws, ws_bytes), sized by the matching *_workspace_bytes();
 *   - randomness is counter-based Philox4x32-10 keyed by (seed, offset +
 *     element id, ...): results depend on (seed, offset) only, never on the
 *     launch geometry or on how rows/vectors are sharded across GPUs.
 */
#ifndef SQB200_H
#define SQB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum sq_status {
    SQ_STATUS_OK = 0,
    SQ_STATUS_INVALID = 1,     /* bad argument */
    SQ_STATUS_CUDA = 2,        /* a CUDA runtime call failed */
    SQ_STATUS_WORKSPACE = 3,   /* workspace missing or too small */
    SQ_STATUS_UNSUPPORTED = 4  /* shape outside what this build supports */
};

/* Message of the last failing call on this host thread. */
const char* sq_last_error(void);
int sq_abi_version(void);

/* ---- tomography --------------------------------------------------------- *
 * real_tomography + QuantumState.measure + estimate_wald
 * (sklearn/QuantumUtility/Utility.py:259-402, :51-55, :61-64), batched over B
 * vectors of length d.  sched_host[0..sched_len) are the attempt sizes N
 * (integers as doubles): one entry = single-shot mode (:365-398); the
 * check_measure(np.geomspace(1,N,100)) schedule (:317-319, :414-422) =
 * incremental mode, stopped at the first attempt with ||V-est|| <= delta when
 * `stop` is set (:355-363).  norm_inf: 0 = L2, 1 = L-infinity (:356-359).
 * Outputs: est[B,d], n_used[B] (N of the returned attempt), err[B]; optional
 * (NULL to skip) trace[B,sched_len,d] = the estimate of every attempt made
 * (the {N: estimate} dict real_tomography returns) and attempts[B].
 * Vector b draws from the Philox stream (seed, vec_offset + b).           */
int sq_tomography_max_fused_dim(void);
int sq_tomography_batched(const double* V, long long B, int d, double delta, int norm_inf, int stop,
                          const double* sched_host, int sched_len, unsigned long long seed,
                          unsigned long long vec_offset, double* est, double* n_used, double* err, double* trace,
                          int* attempts, void* stream);

/* Same routine for any d up to 4M, with scratch in (ws, ws_bytes) sized by sq_tomography_workspace_bytes:
 *  - d <= sq_tomography_max_fused_dim(): one attempt (single-shot mode, or a schedule without the stop rule) of
 *    296 vectors or more runs in the group kernel, which keeps the sum trees of a group of vectors in ws (2.5
 *    doubles per leaf and vector in flight); everything else - and any call whose ws is too small - goes to
 *    sq_tomography_batched.  The two give the same bits.  workspace_bytes is 0 when the group kernel would not
 *    be used.
 *  - longer vectors are cut into blocks of 1024 leaves that the whole grid works on (B <= 65535).           */
size_t sq_tomography_workspace_bytes(long long B, int d);
int sq_tomography_batched_ws(const double* V, long long B, int d, double delta, int norm_inf, int stop,
                             const double* sched_host, int sched_len, unsigned long long seed,
                             unsigned long long vec_offset, double* est, double* n_used, double* err, void* ws,
                             size_t ws_bytes, void* stream);

/* One long vector (or a few) over several GPUs: the leaf blocks of the long-vector path are sharded, rank r
 * splitting blocks [blk_lo, blk_hi) (reference: the single-process loop of real_tomography, Utility.py:259-402;
 * SURVEY 8(e) "single huge vector").  Every rank holds the whole V and the same (seed, vec_offset).  Per attempt
 * the host runs phase 1 (stage-1 counts of the own blocks), sums the block masses bm_a | bm_b over the ranks
 * (entries of foreign blocks zeroed first), phase 2 (stage-2 counts, signs, error partials), sums errp the same
 * way, phase 3 (error, stop rule); phase 0 once before the first attempt.  After the last attempt est is complete
 * on the own blocks only.  Block draws are keyed by (vector, block, node), so the result equals
 * sq_tomography_batched_ws bit for bit whatever the number of ranks.  sq_tomography_workspace_layout fills
 * out[0] = number of blocks, out[1] = leaves per block, out[2..6] = byte offsets in ws of bm_a, bm_b, bcnt,
 * errp ([B, blocks] doubles each) and done ([B] int32), out[7] = total bytes.                             */
int sq_tomography_workspace_layout(long long B, int d, long long out[8]);
int sq_tomography_sharded_phase(const double* V, long long B, int d, double delta, int norm_inf, int stop,
                                double n_measure, int attempt, unsigned long long seed,
                                unsigned long long vec_offset, double* est, double* n_used, double* err, void* ws,
                                size_t ws_bytes, int blk_lo, int blk_hi, int phase, void* stream);

/* Incremental mode with the stop rule, lower latency: `lanes` CTAs per vector try different attempts of the
 * schedule at the same time and record the first (in schedule order) that meets the accuracy; that attempt
 * is then replayed to produce the outputs.  Attempts are fresh draws keyed by their index, so est / n_used /
 * err equal those of sq_tomography_batched with stop = 1, bit for bit.  stop_idx[B] is int scratch.
 * d <= sq_tomography_max_fused_dim(), B <= 65535.                                                        */
int sq_tomography_batched_spec(const double* V, long long B, int d, double delta, int norm_inf,
                               const double* sched_host, int sched_len, int lanes, unsigned long long seed,
                               unsigned long long vec_offset, double* est, double* n_used, double* err,
                               int* attempts, int* stop_idx, void* stream);

/* The same in ONE launch: the attempt lane of a vector that finishes last replays the recorded attempt itself.
 * state[2 B] ints carry the search from lane to lane; prepare them once with sq_tomography_spec_state_init -
 * every call leaves them prepared for the next one (calls that share a state must not overlap on the GPU).
 * Results are those of sq_tomography_batched_spec / sq_tomography_batched with stop = 1, bit for bit.
 * (The q-means loop calls this once per Lloyd iteration for the centroid tomography, _dmeans.py:825-828.)   */
int sq_tomography_spec_state_init(int* state, long long B, void* stream);
int sq_tomography_batched_spec_fused(const double* V, long long B, int d, double delta, int norm_inf,
                                     const double* sched_host, int sched_len, int lanes, unsigned long long seed,
                                     unsigned long long vec_offset, double* est, double* n_used, double* err,
                                     int* attempts, int* state, void* stream);

/* Measurement hook (no reference counterpart; process-wide and NOT thread-safe - meant for tools/ and the
 * A/B tests, not for callers).  None of the switches changes the law of a result, and only `what = 0` changes
 * which draws are used:
 *   0  factor f of the rule "an attempt with N <= f * clamp(L, 32, 512) measurements draws them one by one instead
 *      of splitting counts down the tree" (L = d rounded up to a power of two; both samplers are exact);
 *   1, 2  divisor and cap of the per-vector kernel's threads-per-vector rule for large batches;
 *   3, 4  group kernel on / off, and the batch size from which it is used (default 296 vectors);
 *   6  largest L for which the per-vector kernel keeps the input and its stage-1 CDF in shared memory (256);
 *   7  skipping of attempts that cannot succeed on / off (results are bit-identical either way);
 *   8  block-group kernel for long vectors on / off (off: round 1's one-CTA-per-block kernel).              */
int sq_tomography_tune(int what, double value);

/* ---- q-means E-step ------------------------------------------------------ *
 * labels_estimation (sklearn/cluster/_dmeans.py:732-777) with ipe /
 * amplitude_estimation / median_evaluation (Utility.py:697-737, :442-531,
 * :534-572) and select_labels (_dmeans.py:2252-2257) fused in.
 * mode 0: exact argmin (delta == 0); 1: uniform pick in the delta window
 * (true_distance_estimate=False); 2: IPE-estimated distances (True) - for every pair the median of the 13
 * amplitude-estimation outcomes is drawn from its exact law (csrc/ipe.cuh).
 * xnorm[n] = ||x_i||^2 (sq_row_sqnorms).  Outputs: labels[n] int32,
 * mindist[n] (may be NULL), inertia[1] = sum_i min_j D_ij, n_fixed[1] (may be
 * NULL) = rows re-evaluated with the reference's exact arithmetic.          */
/* workspace_bytes = the fixed part + 60 MiB for the ipe tiles (640 tiles of 96 KB, one per resident CTA);
 * sq_qmeans_assign accepts less in modes 0 and 1 (fixed part only) and runs mode 2 on a smaller grid when fewer
 * tiles fit. */
size_t sq_qmeans_assign_workspace_bytes(long long n, int d, int k);
int sq_qmeans_assign(const double* X, const double* xnorm, long long n, int d, const double* C, int k, int mode,
                     double delta, unsigned long long seed, unsigned long long offset, int* labels,
                     double* mindist, double* inertia, int* n_fixed, void* ws, size_t ws_bytes, void* stream);

/* Test hook (no reference counterpart).  on = 1: every pair of later mode-2 calls is sampled by the draw-by-draw
 * replay path of csrc/ipe.cuh (the path taken when the median of the 13 outcomes may lie outside the enumerated
 * window), so that its law can be tested.  on = 2: every pair is sampled by the inversion of csrc/ipe_lean.cuh,
 * also the ones the kernel normally leaves out because their estimate certainly exceeds one already sampled in
 * the same row - labels, distances and inertia must come out bit for bit as with on = 0.  on = 0: normal.      */
int sq_ipe_debug_force_replay(int on);
/* Test hook (no reference counterpart).  on = 0: modes 0 and 1 of later sq_qmeans_assign calls run the fp64
 * tensor-core kernel for every shape; on = 1 (the default): shapes with k <= 256, d <= 128, d even and
 * 16-byte aligned X / C first narrow each row down to a few columns on the bf16 tensor cores, with a rigorous error
 * bound, and take every decision on fp64 distances of those columns (csrc/qmeans_filter.cuh).  Same labels; the
 * distances agree to the rounding of two fp64 summation orders.  Process-wide, not thread-safe.                  */
int sq_qmeans_debug_filter(int on);
/* Test hook: B draws of the amplitude-estimation outcome j conditioned on lying outside the 8-point window
 * around the grid position x (offsets -3..4 from floor(x)), as the tail sampler of csrc/ipe.cuh makes them.   */
int sq_ipe_debug_tail_sample(double M, double x, long long B, unsigned long long seed, double* out_j, void* stream);

/* Same E-step as sq_qmeans_assign for callers that cannot give the library a workspace
 * (no counterpart in the reference).  Meant for hosts that are not allowed to allocate on
 * the device: a plugin inside a process whose GPU memory belongs to another framework's
 * pool, or a serving loop that cannot afford the device-wide synchronisation cudaMalloc
 * implies.  The scratch is a 2 MiB array built into the library (resident on every GPU
 * that loads it, until exit), so a call is limited to n <= 65536 rows and k <= 1024
 * centroids; larger sizes return SQ_STATUS_UNSUPPORTED with a message.  Nothing is
 * allocated or freed by this call.  The first call on each device looks up the array's
 * device address (that device must be current) - make that first call before starting any
 * stream capture.  All calls share the one array and it keeps the previous call's contents:
 * two calls whose kernels can run at the same time on one GPU (different streams, no
 * dependency) race on it; serialise them.  Inputs, outputs and results are those of
 * sq_qmeans_assign.  In mode 2 every resident CTA parks a tile of dot products (96 KB) in
 * the scratch, so what the fixed part leaves of the 2 MiB bounds the grid (7 CTAs at the
 * largest shape, 65536 x 1024, whose fixed part is 1.33 MB; 21 for a small batch): same
 * results, a fraction of the speed - give sq_qmeans_assign a workspace if you can.  The
 * per-device address of the array is looked up once and cached without a lock (two first
 * calls racing on one device both store the same value); the cache does not survive
 * cudaDeviceReset.                                                                          */
int sq_qmeans_assign_noalloc(const double* X, const double* xnorm, long long n, int d, const double* C, int k,
                             int mode, double delta, unsigned long long seed, unsigned long long offset, int* labels,
                             double* mindist, double* inertia, int* n_fixed, void* stream);

/* Timing hook for the roofline measurement (no reference counterpart): enable = 2 resets
 * the counters and turns per-launch CUDA-event timing of the main E-step kernel on (each
 * sq_qmeans_assign then blocks until its kernel ends); 1 = on, 0 = off.  The totals seen
 * before the call are returned through total_ms / launches (either may be NULL).
 * Like sq_ipe_debug_force_replay and sq_tomography_tune this is a process-wide switch read
 * without synchronisation: NOT thread-safe, and while it is on the "calls never
 * synchronise" convention above does not hold.  Tools and tests only.                    */
int sq_profile_assign(int enable, double* total_ms, long long* launches);

/* row_norms(X, squared=True) (sklearn/utils/extmath.py:49-78) */
int sq_row_sqnorms(const double* X, long long n, int d, double* out, void* stream);
/* deterministic sum of n doubles; ws needs 8192 bytes */
int sq_sum_f64(const double* v, long long n, double* out, void* ws, size_t ws_bytes, void* stream);

/* ---- q-means M-step ------------------------------------------------------ *
 * _centers_update (sklearn/cluster/_dmeans.py:814-823): sums[k,d] and
 * counts[k] of the rows carrying each label (this rank's rows only; the
 * caller all-reduces sums and counts before sq_centroid_finalize).
 * bad_label_flag[1] (may be NULL) is set to 1 if a label is outside [0,k); it is only ever set, never
 * cleared: zero it once and look at it when you like.                        */
size_t sq_centroid_workspace_bytes(long long n, int d, int k);
int sq_centroid_accumulate(const double* X, const int* labels, long long n, int d, int k, double* sums,
                           double* counts, int* bad_label_flag, void* ws, size_t ws_bytes, void* stream);
/* centres = sums / counts (np.true_divide, _dmeans.py:823); a cluster with
 * no rows keeps c_old (the reference leaves that row uninitialised).        */
int sq_centroid_finalize(const double* sums, const double* counts, const double* c_old, int k, int d,
                         double* c_new, void* stream);
/* k-means++ step on the device (reference: _kmeans_plusplus, _dmeans.py:153-247, the euclidean_distances call at
 * :222-224 and the np.minimum at :227): dots[n, T] = X . candidates^T comes from sq_project; this fills
 * d2[t, i] = min(closest[i], max(|x_i|^2 + |c_t|^2 - 2 dots[i, t], 0)) ([T, n], one contiguous row per candidate).
 * closest may be NULL (first centre).                                                                        */
int sq_kpp_update(const double* dots, const double* xnorm, const double* cnorm, const double* closest, long long n,
                  int T, double* d2, void* stream);
/* squared_norm(a - b) (_dmeans.py:651), single deterministic reduction      */
int sq_sqdiff_sum(const double* a, const double* b, long long n, double* out, void* stream);

/* ---- scalar estimators --------------------------------------------------- *
 * One outcome j in [0, M) per element of the law shared by
 * amplitude_estimation (Utility.py:494-508) and phase_estimation
 * (Utility.py:637-653): p(j) ~ 1/sin^2(pi (x - j)/M), x = M*omega, drawn in
 * O(1) expected time.  The _bruteforce variant is the reference's O(M)
 * inversion, kept as the device-side truth for tests.                       */
int sq_fejer_sample(const double* M, const double* x, long long B, unsigned long long seed,
                    unsigned long long offset, double* out_j, void* stream);
int sq_fejer_sample_bruteforce(const double* M, const double* x, long long B, unsigned long long seed,
                               unsigned long long offset, double* out_j, void* stream);
/* out = x + TruncNormal(0,1;[-halfwidth, halfwidth]) per element:
 * make_gaussian_est (Utility.py:88-104), introduce_error(_array) (:68-73).  */
int sq_truncnorm_add(const double* x, long long L, double halfwidth, unsigned long long seed,
                     unsigned long long offset, double* out, void* stream);
/* The same with one half-width per element (half-widths <= 0 leave the element unchanged): the error model of
 * QLSSVC.predict / get_P / get_h, introduce_error(p_i, err / (2 beta_i)) (sklearn/svm/_qSVM.py:199-214, 263-298). */
int sq_truncnorm_add_each(const double* x, long long L, const double* halfwidth, unsigned long long seed,
                          unsigned long long offset, double* out, void* stream);
/* K[n, m] = k(X[n, d], Y[m, d]) for QLSSVC.get_kernel (_qSVM.py:375-389 over scikit-learn's pairwise kernels):
 * kind 0 linear X Y^T, 1 poly (gamma X Y^T + coef0)^degree, 2 rbf exp(-gamma ||x - y||^2) (squared distances
 * clipped at 0, exact zeros on the diagonal when Y == X), 3 sigmoid tanh(gamma X Y^T + coef0).  The dot products
 * run on the fp64 tensor cores (sq_project).  ws from sq_kernel_matrix_workspace_bytes.                      */
size_t sq_kernel_matrix_workspace_bytes(long long n, int m);
int sq_kernel_matrix(const double* X, long long n, const double* Y, int m, int d, int kind, double gamma,
                     double coef0, double degree, double* K, void* ws, size_t ws_bytes, void* stream);
/* Binomial(n_i, pl_i/(pl_i+pr_i)) per element: the sampler the tomography
 * kernel splits its multinomial counts with (test hook; no reference
 * counterpart beyond RandomState.choice, Utility.py:54).                    */
int sq_binomial_sample(const double* n, const double* pl, const double* pr, long long B, unsigned long long seed,
                       unsigned long long offset, double* out, void* stream);

/* ---- mu-norm search -------------------------------------------------------- *
 * best_mu / linear_search / __mu (Utility.py:196-231) for a whole p-grid in
 * two streaming passes.  exp_row_host[g] = 2p_g, exp_col_host[g] = 2(1-p_g);
 * exponent 0 means count_nonzero (:198-199).  Outputs (this rank's rows):
 * rowmax[n_p], colsum[n_p, d], frob2[1].                                    */
size_t sq_mu_norms_workspace_bytes(long long n, int d, int n_p);
int sq_mu_norms(const double* X, long long n, int d, const double* exp_row_host, const double* exp_col_host, int n_p,
                double* out_rowmax, double* out_colsum, double* out_frob2, void* ws, size_t ws_bytes, void* stream);

/* ---- Gram route for qPCA / condition number ------------------------------ *
 * out_gram[d,d] = (X - mean)^T (X - mean) over this rank's rows on the fp64
 * tensor cores (mean may be NULL: no centring).  New algorithm for the
 * outputs of linalg.svd in qPCA._fit_full (sklearn/decomposition/_qPCA.py:
 * 578-583) and np.linalg.svd in qMeans_.fit (sklearn/cluster/_dmeans.py:1244):
 * sigma = sqrt(eig(G)), V = eigenvectors.                                   */
size_t sq_gram_workspace_bytes(long long n, int d);
int sq_gram_centered(const double* X, long long n, int d, const double* mean, double* out_gram, void* ws,
                     size_t ws_bytes, void* stream);
/* svd_flip's decision data (sklearn/utils/extmath.py:552-557, called at _qPCA.py:583) without
 * materialising U: for every right singular vector v_j (rows of Vt [r,d]) the row i that
 * maximises |x_i.v_j - mu_v[j]| over this rank's rows.  out_key[j] = bits(|y|) with bit 0
 * replaced by the sign of y (1 = negative); combine ranks with max.  mu_v[j] = mean.v_j
 * (pass zeros for centred X).                                                            */
int sq_project_absmax(const double* X, long long n, int d, const double* Vt, int r, const double* mu_v,
                      unsigned long long* out_key, void* stream);
/* (X - mean) @ components.T of _BasePCA.transform (sklearn/decomposition/_base.py:117-128) on the fp64
 * tensor cores: out[n, r] = X Vt^T - offset, offset[j] = mean.v_j (zeros for centred X).            */
int sq_project(const double* X, long long n, int d, const double* Vt, int r, const double* offset, double* out,
               void* stream);
/* column sums of X (np.mean(X, axis=0) * n: _qPCA.py:578, _dmeans.py:1264) */
size_t sq_col_sums_workspace_bytes(long long n, int d);
int sq_col_sums(const double* X, long long n, int d, double* out, void* ws, size_t ws_bytes, void* stream);
/* Y = X - mu row-wise (X -= X_mean, _dmeans.py:1266; _qPCA.py:579); Y may alias X */
int sq_sub_rowvec(const double* X, const double* mu, long long n, int d, double* Y, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SQB200_H */
