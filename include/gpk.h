/* libgpk -- C ABI of the B200-native GP-fitting hot path (drop-in for rferguson22/Gaussian-Process).
 *
 * The reference has no FFI: its boundary is the Python functions of GP_func.py, find_len_scales.py and
 * calc_prob_surf.py.  The Python drop-in modules in gaussian-process_b200/dropin/ keep those signatures
 * and call the entry points below through ctypes (see INTEGRATION.md for the binding).
 *
 * Conventions
 *   - every pointer is a DEVICE pointer to C-contiguous float64 / int32 / uint8 data unless it says HOST;
 *   - data layout is the reference's: points are dims-first, X is (d, N), ls is (P, d);
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream);
 *   - functions return 0 on success, non-zero on error (gpk_last_error_string() explains); they never
 *     throw, never allocate: scratch is sized by the *_workspace_bytes query and passed in;
 *   - there is no CPU fallback anywhere behind this interface.
 */
#ifndef GPK_H_
#define GPK_H_
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define GPK_API __attribute__((visibility("default")))
#else
#define GPK_API
#endif

#define GPK_VERSION 200
#define GPK_NB 128            /* tile edge; N is padded to a multiple of it internally (Np) */
#define GPK_SMALL_N_MAX 128   /* N <= this: one-CTA shared-memory path */
#define GPK_MAX_D 8            /* kinematic dimensions; more is an error (status 2), not a fallback */
#define GPK_MAX_S 4096         /* sigma-scale grid points; find_len_scales.py:66 uses 1000 */

/* `precision` argument: arithmetic of the dense contractions of the blocked path (N > 128) -- the Cholesky trailing updates,
 * GP_func.py:22, and the triangular-inverse updates that replace the N-right-hand-side solve of GP_func.py:38.  It is a
 * per-call argument: the library keeps no precision state, two host threads may use different modes concurrently.
 *   GPK_PREC_FP64   native FP64 tensor instructions (DMMA, mma.sync.m8n8k4.f64) -- sm_100a has no FP64 kind of tcgen05.mma;
 *   GPK_PREC_EMU34  emulated FP64: every operand is split error-free into 7 signed 8-bit digits (Ozaki scheme), the 34 digit
 *                   products of weight >= 2^-56 run on the int8 tcgen05 tensor cores with exact int32 accumulation in TMEM,
 *                   the FP64 result is rebuilt in the epilogue.  Rounding level of an FP64 dot product of the same length.
 *                   Requires e > 0 (the column bound |L^-1[i][k]| <= 1 / e_k is used) and N <= 16256;
 *   GPK_PREC_EMU28  the 28 digit products of weight >= 2^-48 (7 digits, truncation ~2^8 larger);
 *   GPK_PREC_EMU21  6 digits, 21 products (operands rounded to 47 bits + sign relative to their row scale);
 *   GPK_PREC_EMU15  5 digits, 15 products (39 bits + sign).
 * The reduced modes are meant to be used with the a-posteriori check (check_tol of gpk_loss_batched, probe_out of gpk_fit). */
#define GPK_PREC_FP64 0
#define GPK_PREC_EMU34 1
#define GPK_PREC_EMU28 2
#define GPK_PREC_EMU21 3
#define GPK_PREC_EMU15 4
/* mixed: 28 products in the Cholesky phase + 34 in the inverse phase (5), and the other way round (6); measured in
 * DESIGN.md 3.2 */
#define GPK_PREC_EMU28_34 5
#define GPK_PREC_EMU34_28 6

/* `flags` argument of gpk_loss_batched (bit mask).
 *   GPK_FLAG_REFERENCE_FORM  which algebraic form of the self-prediction the loss uses.  Clear: the reduced form,
 *                       mu - y = -e^2 alpha, var = e^2 - e^4 (K^-1)_ii (triangular inverse, 2/3 N^3 flop; every N).  Set:
 *                       the reference's own formulae, GP_func.py:36-40: mu = K_s^T alpha, v = L^-1 K_s, var = 1 - sum v^2
 *                       (4/3 N^3 flop, with the reference's cancellations); one-CTA path only (N <= GPK_SMALL_N_MAX).
 *   GPK_FLAG_KEEP_INACTIVE   particles with active[p] == 0 keep their previous counts / info / resid / sd instead of being
 *                       zeroed: re-evaluation of a subset of a swarm in place (blocked path, active != NULL). */
#define GPK_FLAG_REFERENCE_FORM 1
#define GPK_FLAG_KEEP_INACTIVE 2

GPK_API int gpk_version(void);
GPK_API const char* gpk_last_error_string(void);            /* HOST string, valid until the next failing call */
GPK_API unsigned long long gpk_launch_count(void);          /* kernels launched by this library so far (process-wide) */
GPK_API int gpk_padded_n(int N);                            /* Np = N rounded up to a multiple of GPK_NB */

/* (a) kernel_func(x1, x2, l), GP_func.py:49-65, for P length-scale vectors in one launch.
 *     K[p][i][j] = exp(-0.5 * max(|x1_i/l_p|^2 + |x2_j/l_p|^2 - 2 <x1_i/l_p, x2_j/l_p>, 0)); when e != NULL
 *     (x1 == x2 case of GP_func.py:21) e_i^2 is added on the diagonal.  Row stride ldk, matrix stride strideK. */
GPK_API int gpk_cov_batched(const double* X1, const double* X2, const double* e, const double* ls, int P, int N1,
                    int N2, int d, double* K_out, long long ldk, long long strideK, void* stream);

/* (b) batched blocked Cholesky, numpy.linalg.cholesky at GP_func.py:22, for P matrices [P][Np][Np]
 *     (lower triangle read, lower triangle overwritten by L, strict upper of the diagonal blocks zeroed;
 *     blocks above the diagonal are not touched).  info[p] = 0 or 1 + index of the first bad pivot. */
GPK_API size_t gpk_potrf_workspace_bytes(int P, int Np);
GPK_API int gpk_potrf_batched(double* K_inout, int P, int Np, int* info, void* workspace, size_t workspace_bytes,
                      void* stream);

/* (b) staged form of the fit / self-prediction, on factors left by gpk_potrf_batched ([P][Np][Np], padding =
 *     identity): alpha = L^-T (L^-1 y), GP_func.py:24, and the calibration counts of find_len_scales.py:159-163 from
 *     (L, alpha).  gpk_loss_batched is the fused form of cov + potrf + these two and is what the PSO calls; these entry
 *     points exist for callers that keep the factor.  alpha is [P][Np] (row stride Np, entries >= N are 0).  Native FP64. */
GPK_API size_t gpk_staged_workspace_bytes(int P, int Np);
GPK_API int gpk_fit_alpha_batched(const double* L, const double* y, int P, int N, int Np, double* alpha_out,
                          void* workspace, size_t workspace_bytes, void* stream);
GPK_API int gpk_selfpredict_counts(const double* L, const double* alpha, const double* y, const double* e,
                           const double* sigma_vals, const unsigned char* active, int P, int N, int Np, int S,
                           int* counts_out, double* resid_out, double* sig_out, void* workspace,
                           size_t workspace_bytes, void* stream);

/* (b) the PSO hot call: evaluate_loss for a whole swarm, find_len_scales.py:154-163 up to the counts.
 *     counts[p][s] = #{ i : |mu_i - y_i| / max(sd_i * sigma_vals[s], 1e-12) <= 1 } (int32); the host finishes
 *     :163-177 in NumPy.  sigma_vals must be ascending (the count is found by bisection: the predicate is monotone in s;
 *     the Python mirror sorts and un-permutes other grids).  active[p] == 0 marks an out-of-bounds particle
 *     (find_len_scales.py:156): it is not fitted and its counts are 0.  resid_out / sig_out ([P][N], optional, may be NULL)
 *     receive mu - y and sd.  info[p] != 0: covariance not positive definite (the reference raises LinAlgError).
 *
 *     Checked mixed precision (blocked path): when check_tol > 0 or probe_out != NULL every particle's factor is probed
 *     with one fixed pseudo-random vector v: rho_p = |W^T (W (K v)) - v|_inf / |v|_inf, W = L^-1 as computed, K as built
 *     (three N^2 matrix-vector products fused into passes that exist anyway).  rho_p -> probe_out[p] (optional).  With an
 *     emulated `precision` and check_tol > 0, every particle with rho_p > check_tol (or NaN) is evaluated again by the
 *     native FP64 kernels inside the same call and fallback_out[p] (optional, [P] bytes) is set to 1.  No host
 *     synchronisation: the call stays asynchronous and can be captured into a CUDA graph (when nothing failed, the kernels
 *     of the second pass exit at their first instruction).  check_tol < 0: the probe runs with tolerance -check_tol and
 *     fallback_out is set, but nothing is re-evaluated: the caller reads the flags and calls again with GPK_PREC_FP64,
 *     active = the flags and GPK_FLAG_KEEP_INACTIVE (what the Python mirror does where the second pass' ~6 T empty
 *     launches would show: CUDA-graph replay of launch-bound sizes). */
GPK_API size_t gpk_loss_workspace_bytes(int P, int N, int d, int precision);
GPK_API int gpk_loss_batched(const double* X, const double* y, const double* e, const double* ls,
                     const unsigned char* active, const double* sigma_vals, int P, int N, int d, int S,
                     int precision, int flags, double check_tol, int* counts_out, int* info, double* resid_out,
                     double* sig_out, double* probe_out, unsigned char* fallback_out, void* workspace,
                     size_t workspace_bytes, void* stream);

/* (b)+(c) GP(): fit part, GP_func.py:21-24.  Produces the cached state used by gpk_predict:
 *     W = L^-1 ([Np][Np], lower block triangle), alpha = K^-1 y ([Np]), xs = X / l ([d][Np]), sq ([Np]).
 *     probe_out (optional, one double): the residual probe rho of gpk_loss_batched for this factor; the caller refits with
 *     GPK_PREC_FP64 when it exceeds its tolerance (engine.GPFit does). */
GPK_API size_t gpk_fit_workspace_bytes(int N, int d, int precision);
GPK_API int gpk_fit(const double* X, const double* y, const double* e, const double* ls, int N, int d, int precision,
            double* W_out, double* alpha_out, double* xs_out, double* sq_out, int* info, double* probe_out,
            void* workspace, size_t workspace_bytes, void* stream);

/* (c) GP(): predict part, GP_func.py:26-45, for M grid points Xfit (d, M):
 *     mu = K_*^T alpha, sd = sqrt(clip(1 - |W k_*|^2, 1e-12)).  chunk = grid points per internal pass. */
GPK_API size_t gpk_predict_workspace_bytes(int N, int chunk);
GPK_API int gpk_predict(const double* W, const double* alpha, const double* xs, const double* sq, const double* ls,
                int N, int d, const double* Xfit, long long M, double* mu_out, double* sd_out, int chunk,
                void* workspace, size_t workspace_bytes, void* stream);

/* (c) gpk_predict for callers that cannot allocate.  No counterpart in the reference: added
 *     for code under CUDA-graph capture, tight serving loops and harnesses that forbid cudaMalloc.  Same
 *     arguments, kernels and results as gpk_predict with chunk = GPK_STATIC_CHUNK, but the scratch (K_*^T of one chunk and the
 *     partial sums, 4.03 MiB) is a file-scope `static __device__` array inside the library: the call allocates nothing (no
 *     cudaMalloc / cudaFree, no allocating library routine) and takes no workspace.  Every GPU that loads the library holds
 *     the array until the process exits.  Limit: N <= GPK_STATIC_MAX_N known points (status 2 beyond it, with an
 *     explanation in gpk_last_error_string(); use gpk_predict there); M is unlimited (chunks run one after the other).
 *     Every call on a device shares the one array, which keeps the previous call's contents (nothing in it has to start at
 *     zero: each element is written before it is read within a call).  Calls that overlap on the GPU -- two streams, two host
 *     threads, two graphs replayed concurrently -- RACE on it: serialise them on one stream. */
#define GPK_STATIC_MAX_N 2048
#define GPK_STATIC_CHUNK 256
GPK_API int gpk_predict_static(const double* W, const double* alpha, const double* xs, const double* sq, const double* ls,
                               int N, int d, const double* Xfit, long long M, double* mu_out, double* sd_out, void* stream);

/* (c) the same prediction with |W k_*|^2 on the int8 tcgen05 digit GEMM (emulated FP64, see GPK_PREC_*):
 *     gpk_predict_prepare slices the fitted W once into a caller-owned state buffer (needs e > 0), gpk_predict_emu then
 *     builds the digits of each K_* chunk on the fly (no FP64 K_* in HBM) and reduces the squares in the epilogue.
 *     Results do not depend on `chunk`.  precision: one of GPK_PREC_EMU* (digit products kept by the GEMM). */
GPK_API size_t gpk_predict_state_bytes(int N);
GPK_API int gpk_predict_prepare(const double* W, const double* e, int N, void* state, size_t state_bytes, void* stream);
GPK_API size_t gpk_predict_emu_workspace_bytes(int N, int chunk);
GPK_API int gpk_predict_emu(const void* state, const double* alpha, const double* xs, const double* sq, const double* ls,
                    int N, int d, int precision, const double* Xfit, long long M, double* mu_out, double* sd_out,
                    int chunk, void* workspace, size_t workspace_bytes, void* stream);

/* (N4) one Lloyd iteration of the KMeans subsample, find_len_scales.py:25-47 (sklearn.cluster.KMeans(...).fit_predict with
 *     sklearn's single-threaded arithmetic).  X is [N][d] (points first, as sklearn holds them), centers [k][d], csq [k] their
 *     squared norms.  E-step: labels_inout[i] = first j minimising csq_j - 2 <x_i, c_j>; *changed_out is set to 1 when a label
 *     differs from the one the buffer held (the caller clears it; labels start at -1).  When sums_out != NULL the M-step sums
 *     follow: sums_out[j] = sum of the members of cluster j in ascending sample order, counts_out[j] = their number (float64).
 *     Everything on the centres' side (norms, empty clusters, averaging, shift, tolerance, k-means++ seeding) is the host's:
 *     gaussian-process_b200/kmeans.py.  Limits: k <= 1024, d <= GPK_MAX_D. */
GPK_API int gpk_kmeans_step(const double* X, int N, int d, const double* centers, const double* csq, int k,
                            int* labels_inout, int* changed_out, double* sums_out, double* counts_out, void* stream);

/* (b) the swarm loop of len_scale_opt kept on the device, find_len_scales.py:76-141.  After every gpk_loss_batched of the
 *     swarm, gpk_pso_step (one small kernel, asynchronous) finishes the scores from the counts (find_len_scales.py:163-177,182:
 *     NumPy's operation order and pairwise summation, bit for bit what the host finish gives), absorbs them (personal and
 *     global bests, stagnation counter, :79-83,106-125), and prepares the next evaluation (:88-100 velocity / position update of
 *     the next iteration, or the soft restart of :123-141): it writes the next positions [P][d] and the in-bounds mask [P] that
 *     the next gpk_loss_batched reads.  No host round trip per iteration: the caller enqueues max_evals pairs of calls and
 *     reads the state back once.  Random numbers are taken from a stream the caller generated with NumPy's legacy generator
 *     (np.random.random_sample), in the reference's order (two rand(P, d) per iteration, P d uniforms per restart), so the
 *     trajectory is the host loop's.
 *     State buffer (caller-allocated, gpk_pso_layout bytes): int32 header[32] = {P, d, S, n_points, max_iter, patience,
 *     max_evals, iteration, pending (0 initial evaluation, 1 iteration, 2 restart re-evaluation, 3 done), stall, restarts,
 *     evaluations recorded, error (1: covariance not positive definite, 2: random stream exhausted), random cursor, particles
 *     that fell back to FP64, random stream length, iterations logged}, then the arrays at the byte offsets gpk_pso_layout reports, in this order:
 *     scalars (gbest score, inertia decay), lower, upper, v_max [d], positions, velocities, personal best positions [P][d],
 *     personal best scores [P], global best position [d], last scores [P], expected [S], diff(sigma_vals) [S] (S - 1 used),
 *     measured_of [n_points + 1] (k / n_points as NumPy divides), per-iteration log: best score (f64), stall (i32), restarted
 *     (i32) [max_iter], trace of every evaluation: positions [max_evals][P][d], scores [max_evals][P], random stream.
 *     The caller fills everything up to the logs before the first call (positions = the initial swarm, pending = 0); the
 *     header's P, d, S, n_points, max_iter, max_evals and random stream length must equal the arguments of every call. */
#define GPK_PSO_N_OFFSETS 19
GPK_API size_t gpk_pso_layout(int P, int d, int S, int n_points, int max_iter, int max_evals, long long rng_len,
                              unsigned long long* offsets_out);
GPK_API int gpk_pso_step(void* state, size_t state_bytes, int P, int d, int S, int n_points, int max_iter, int max_evals,
                         long long rng_len, const int* counts, const int* info, const unsigned char* fallback,
                         double* positions_out, unsigned char* active_out, void* stream);

/* (d) generate_prob_surf row loop, calc_prob_surf.py:67-81 + sum_gaussians :15-30.
 *     table: [R][ncols] payload columns of the merged frame (interleaved mean, std; non-finite = gap).
 *     valid[r] = 1 if the row emits output; y_out/p_out [R][npoints] are written for valid rows only. */
GPK_API int gpk_prob_surface(const double* table, long long R, int ncols, int npoints, double* y_out, double* p_out,
                     int* valid_out, void* stream);

/* (d) compaction of the emitting rows (the reference appends output only for rows that pass calc_prob_surf.py:71):
 *     gpk_surface_scan: dst_out[r] = number of emitting rows before r ([R] int64), count_out[0] = their total (device int64);
 *     gpk_surface_gather: copies the y / p rows of the emitting rows to rows dst[r] of y_compact / p_compact and writes
 *     row_index_out[dst[r]] = r.  The caller reads count_out between the two (it sizes the device-to-host copy) and skips
 *     the gather when every row emits. */
GPK_API int gpk_surface_scan(const int* valid, long long R, long long* dst_out, long long* count_out, void* stream);
GPK_API int gpk_surface_gather(const int* valid, const long long* dst, long long R, int npoints, const double* y,
                               const double* p, double* y_compact, double* p_compact, long long* row_index_out, void* stream);

/* (d) sum_gaussians(temp_y, temp_gaus), calc_prob_surf.py:15-30, at n caller-supplied points y; gaus holds
 *     ng interleaved (mean, std) values; half_dy = (|max(y)-min(y)| / n) / 2 as formed at :23,:26-27. */
GPK_API int gpk_sum_gaussians(const double* y, int n, const double* gaus, int ng, double half_dy, double* p_out,
                      void* stream);

/* Scheduling knob, process-wide, results do not depend on it: gpk_loss_batched splits a swarm on the blocked path into at
 * most `groups` particle groups on concurrent streams (default 2, 1..8), each of at least `min_particles_per_group`
 * particles (default 16: smaller groups leave the persistent digit-GEMM launches of late block columns with fewer tiles
 * than SMs).  (0, 0) restores the defaults. */
GPK_API int gpk_set_groups(int groups, int min_particles_per_group);

/* Emulated-FP64 primitives (used by the blocked path in the GPK_PREC_EMU* modes, exposed for tests and measurement).
 *   gpk_oz_slice: rows x cols FP64 (both multiples of 128, row stride ld) -> digit records; rowinv[r] (optional) is
 *     the reciprocal power-of-two row scale: |src[r][c] * rowinv[r]| <= 1/2 is the caller's contract.
 *   gpk_oz_gemm_tn: C[M][N] = sum_k A[m][k] B[n][k] over sliced operands, scaled by sa[m] * sb[n] (the row scales).
 *     groups = 8 (34 digit products), 7 (28), 6 (21) or 5 (15); dbg (optional) receives the raw int32 accumulators
 *     [M/128][N/64][groups][128][64]. */
GPK_API size_t gpk_oz_slice_bytes(int rows, int cols);
GPK_API int gpk_oz_slice(const double* src, int ld, int rows, int cols, const double* rowinv, void* dst, void* stream);
GPK_API int gpk_oz_gemm_tn(const void* A, const void* B, int M, int N, int K, const double* sa, const double* sb,
                   double* C, int ldc, int* dbg, int groups, void* stream);

/* Result tables as text (host code, no device work): rows x cols float64 (row stride ld) -> CSV bytes exactly as the
 * reference writes them with pandas.DataFrame.to_csv(index=False) (GP_fit.py:63,81,104; calc_prob_surf.py:84):
 * shortest round-trip digits, positional for 1e-4 <= |x| < 1e16, "1e-05"-style exponents otherwise, "inf"/"-inf",
 * NaN as an empty field, '\n' line ends; no header line.  `out` must hold gpk_csv_bound(rows, cols) bytes; rows are
 * formatted by `threads` host threads (0 = all); returns the number of bytes written, or -1 on bad arguments. */
GPK_API size_t gpk_csv_bound(long long rows, int cols);
GPK_API long long gpk_csv_format_f64(const double* table, long long rows, int cols, long long ld, char* out,
                                     size_t out_cap, int threads);

/* HOST function.  The long-format block generate_prob_surf writes (calc_prob_surf.py:80-84) from the reduction's outputs:
 * merged row r emits `points` lines "coords[r][0],...,coords[r][ndims-1],y[r][j],p[r][j]\n".  The coordinate text is
 * formatted once per merged row.  Same bytes as gpk_csv_format_f64 on the expanded table; bound from gpk_csv_surface_bound. */
GPK_API size_t gpk_csv_surface_bound(long long rows, int ndims, int points);
GPK_API long long gpk_csv_format_surface(const double* coords, const double* y, const double* p, long long rows, int ndims,
                                 int points, char* out, size_t out_cap, int threads);

/* HOST function.  Loss finish of find_len_scales.py:163-166 for P particles: W_p = np.trapz(|counts[p] / n_points - expected|,
 * sigma_vals) with d_sigma = diff(sigma_vals) ([S-1]) and measured_of[k] = k / n_points for k = 0..n_points (the caller's
 * table of the quotients); NumPy's operation order and pairwise summation, one pass per particle.  scratch: 2 S doubles.
 * Returns 0, 2 for bad arguments, 3 for a count outside 0..n_points. */
GPK_API int gpk_finish_wasserstein(const int* counts, long long P, int S, const double* measured_of, int n_points,
                           const double* expected, const double* d_sigma, double* scratch, double* out);

/* Test hook: launches a kernel that executes a trap and waits for it.  Returns non-zero with gpk_last_error_string() set; the
 * CUDA context of the process is unusable afterwards (every later entry point fails the same way) -- run it in a child
 * process (tests/test_gpu_faults.py).  It shows how the watchdog trap of the digit-GEMM pipeline would surface. */
GPK_API int gpk_debug_trap(void* stream);

/* Measurement aid (bench.py roofline): per-kernel-class device time by events recorded after every launch
 * on `stream` between begin and end.  Not thread-safe; name_out is a HOST buffer. */
GPK_API int gpk_timing_begin(void* stream);
GPK_API int gpk_timing_end(void);
GPK_API int gpk_timing_query(int idx, char* name_out, int name_cap, double* ms_out, long long* count_out);

#ifdef __cplusplus
}
#endif
#endif /* GPK_H_ */
