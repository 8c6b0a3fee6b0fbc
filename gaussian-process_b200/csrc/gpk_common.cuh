// Shared declarations for libgpk (B200 / sm_100a GP-fitting hot path).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>

namespace gpk {

constexpr int NB = 128;           // tile edge of the blocked factorisation (doubles)
constexpr int MAX_D = 8;          // kinematic dimensions supported by the kernels
constexpr int MAX_S = 4096;       // sigma-scale grid points supported by the count kernels
constexpr int SMALL_N_MAX = 128;  // one-CTA (shared-memory resident) loss path up to this N

// ---- library state (gpk_api.cu) -------------------------------------------------------
void set_error(const char* fmt, ...);
void count_launch(int n = 1);
int check_launch(const char* what);   // cudaGetLastError -> status; also counts the launch

inline int round_up(int x, int m) { return (x + m - 1) / m * m; }

#ifdef __CUDACC__
// Fixed pseudo-random probe vector of the a-posteriori residual check (gpk_loss_batched / gpk_fit, include/gpk.h):
// v_i = +-[0.5, 1) from a hash of the index, the same for every particle and every call.
__device__ __forceinline__ double probe_v(int i) {
    unsigned long long z = (unsigned long long)(unsigned)i * 0x9E3779B97F4A7C15ull + 0xD1B54A32D192ED03ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= z >> 31;
    const double m = 0.5 + (double)(z >> 12) * 0x1p-53;
    return (z & 1ull) ? -m : m;
}
#endif

// ---- workspace layout of the blocked loss / fit path ------------------------------------
// All offsets in bytes, each section 256-byte aligned.
struct BlockedLayout {
    int P, N, Np, T, d;
    size_t mat;      // [P][Np][Np]   K -> L -> W = L^-1 (lower block triangle)
    size_t dinv;     // [P][T][NB][NB] inverse of each diagonal Cholesky block (zero upper)
    size_t dinvT;    // [P][T][NB][NB] its transpose
    size_t mt;       // [P][NB][Np]    phase-2 panel (L_KJ * D_J)^T
    size_t xs;       // [P][d][Np]     x / l
    size_t sq;       // [P][Np]        sum_d xs^2
    size_t cn;       // [P][Np]        column sums of squares of W  (= diag K^-1)
    size_t z;        // [P][Np]        W y
    size_t alpha;    // [P][Np]        W^T z = K^-1 y
    // a-posteriori residual probe (include/gpk.h, check_tol): rho = |W^T W K v - v|_inf / |v|_inf for the fixed vector probe_v
    size_t pkv;      // [P][T][Np]     per-tile partial sums of K v left by the covariance kernel
    size_t pu;       // [P][Np]        u = K v
    size_t pt;       // [P][Np]        t = W u
    size_t pw;       // [P][Np]        w = W^T t
    size_t probe;    // [P]            rho
    size_t fail;     // [P] bytes      rho > check_tol: re-evaluated by the native FP64 kernels
    // emulated-FP64 (int8 digit) mode only, gpk_oz.cu:
    size_t oz_slices;  // [P] digit records of L (phase 1) then W gamma (phase 2), oz_slice_bytes(Np) each
    size_t oz_mt;      // [P] digit records of the phase-2 panel, oz_mt_bytes(Np) each
    size_t oz_rows;    // [4][Np]   rowsig, rowinv, gamma, 1/gamma (particle independent)
    size_t oz_mtsc;    // [3][P][NB] row scales of the phase-2 panel, their inverses, and the row maxima they come from
    size_t oz_sdig;    // [P] digits of the Schur-complement block column (TRSM operand), oz_panel_bytes(Np) each
    size_t oz_ddig;    // [P] digits of D_J and D_J^T for every J, oz_ddig_bytes(Np) each
    size_t oz_dsc;     // [P][T][2][NB] their measured row scales
    size_t oz_cnt;     // [4 T + 8] int tile counters, one per digit-GEMM launch (zeroed once per evaluation)
    size_t total;
};
BlockedLayout blocked_layout(int P, int N, int d, int emulated);

// ---- kernels launchers (implemented across the .cu files) ---------------------------------
// gpk_cov.cu
int launch_scale_points(const double* X, const double* ls, int P, int N, int Np, int d,
                        double* xs, double* sq, const unsigned char* active, cudaStream_t st);
int launch_cov_lower(const double* xs, const double* sq, const double* e, int P, int N, int Np, int d,
                     double* mat, double* kv, const unsigned char* active, cudaStream_t st);
int launch_cov_full(const double* X1, const double* X2, const double* e, const double* ls, int P,
                    int N1, int N2, int d, double* K, long long ldk, long long strideK, cudaStream_t st);
int launch_cross_cov(const double* xs, const double* sq, const double* ls, const double* alpha, int N, int Np,
                     int d, const double* Xfit, long long M, long long m0, int B, int Bp, double* KsT,
                     double* mu_out, cudaStream_t st);

// gpk_diag.cu
enum DiagMode { DIAG_WRITE_INVERSE = 0, DIAG_WRITE_FACTOR = 1 };
int launch_diag_block(double* mat, int Np, int J, int P, double* dinv, double* dinvT, int T,
                      int* info, const unsigned char* active, int mode, cudaStream_t st);

int launch_tri_inv_diag(const double* Lsrc, double* dst, int Np, int T, int P, double* dinv, double* dinvT, int* info,
                        const unsigned char* active, cudaStream_t st);

// gpk_gemm.cu
int launch_syrk_update(double* mat, int Np, int T, int J, int P, const unsigned char* active, cudaStream_t st);
int launch_trsm_panel(double* mat, int Np, int T, int J, int P, const double* dinv,
                      const unsigned char* active, cudaStream_t st);
int launch_inv_panel(const double* mat, int Np, int T, int J, int P, const double* dinvT, double* mt,
                     const unsigned char* active, cudaStream_t st);
int launch_inv_update(double* mat, int Np, int T, int J, int P, const double* mt,
                      const unsigned char* active, cudaStream_t st);
int launch_predict_gemm(const double* W, int Np, int T, const double* KsT, int Bp, double* partial,
                        cudaStream_t st);

// gpk_oz.cu -- emulated FP64 on tcgen05 int8 tensor cores
size_t oz_slice_bytes(int Np);
size_t oz_mt_bytes(int Np);
int launch_oz_prep(const double* e, int N, int Np, double* rowsig, double* rowinv, double* gam, double* gaminv,
                   cudaStream_t st);
size_t oz_panel_bytes(int Np);
size_t oz_ddig_bytes(int Np);
int launch_oz_syrk(double* mat, int Np, int T, int J, int P, const uint8_t* slices, const double* rowsig,
                   const double* rowinv, uint8_t* sdig, int* counter, const unsigned char* active, cudaStream_t st);
int launch_oz_slice_d(const double* dinv, const double* dinvT, int T, int J, int P, const double* rowsig, uint8_t* ddig,
                      double* dsc, const unsigned char* active, cudaStream_t st);
int launch_oz_trsm(double* mat, int Np, int T, int J, int P, const uint8_t* sdig, const uint8_t* ddig, const double* dsc,
                   const double* rowsig, const double* rowinv, uint8_t* slices, int* counter, const unsigned char* active,
                   cudaStream_t st);
int launch_oz_invp(double* mt, int Np, int T, int J, int P, const uint8_t* ddig, const double* dsc, const uint8_t* slices,
                   const double* rowsig, unsigned long long* mtmax, const double* gaminv, int* counter,
                   const unsigned char* active, cudaStream_t st);
int launch_oz_slice_l(const double* mat, int Np, int T, int J, int P, const double* rowinv, uint8_t* slices,
                      const unsigned char* active, cudaStream_t st);
int launch_oz_slice_w(const double* mat, int Np, int T, int J, int diag, int P, const double* gam, uint8_t* slices,
                      const unsigned char* active, cudaStream_t st);
int launch_oz_mt(const double* mt, int Np, int T, int J, int P, const double* gaminv, unsigned long long* mtmax,
                 size_t mtmax_stride, double* mtsig, uint8_t* mtslices, const unsigned char* active, cudaStream_t st);
int launch_oz_invu(double* mat, int Np, int T, int J, int P, uint8_t* slices, const uint8_t* mtslices,
                   const double* mtsig, const double* gam, int emit, int* counter, const unsigned char* active,
                   cudaStream_t st);
int launch_oz_slice_generic(const double* src, int ld, int rows, int cols, const double* rowinv, uint8_t* dst,
                            cudaStream_t st);
int launch_oz_gemm_tn(const uint8_t* A, const uint8_t* B, int M, int N, int K, const double* sa, const double* sb,
                      double* C, int ldc, int* dbg, cudaStream_t st);
size_t oz_pred_state_bytes(int Np);
size_t oz_pred_workspace_bytes(int Np, int Bp);
int launch_oz_pred_prepare(const double* W, const double* e, int N, int Np, uint8_t* state, cudaStream_t st);
int launch_oz_predict_chunk(const uint8_t* state, const double* xs, const double* sq, const double* ls, const double* alpha,
                            int N, int Np, int d, const double* Xfit, long long M, long long m0, int B, int Bp,
                            uint8_t* workspace, double* mu_out, double** partial_out, cudaStream_t st);
void oz_set_groups(int ng);
int oz_get_groups();

// gpk_loss.cu
int launch_small_loss(const double* X, const double* y, const double* e, const double* ls,
                      const unsigned char* active, const double* sigma_vals, int P, int N, int d, int S,
                      int* counts, int* info, double* resid_out, double* sig_out, int form, cudaStream_t st);
int launch_wy(const double* mat, int Np, int N, int P, const double* y, double* z, const double* pu, double* pt,
              const unsigned char* active, cudaStream_t st);
int launch_colnorm_alpha(const double* mat, int Np, int P, const double* z, double* cn, double* alpha,
                         const double* pt, double* pw, const unsigned char* active, cudaStream_t st);
int launch_probe_sum(const double* kv, int Np, int T, int P, double* u, const unsigned char* active, cudaStream_t st);
int launch_counts(const double* alpha, const double* cn, const double* e, const double* sigma_vals,
                  int P, int N, int Np, int S, int* counts, double* resid_out, double* sig_out, const double* pw,
                  double tol, double* probe, unsigned char* fail, int zero_inactive,
                  const unsigned char* active, cudaStream_t st);
int launch_probe_value(const double* pw, int N, int Np, int P, double* probe, unsigned char* fail, double tol,
                       const unsigned char* active, cudaStream_t st);
int launch_clear_info(int* info, const unsigned char* mask, int P, cudaStream_t st);
int launch_predict_finalize(const double* partial, int T, int Bp, int B, double* sigma_out, cudaStream_t st);

// gpk_pso.cu -- device-resident swarm loop: byte offsets of the state buffer (include/gpk.h gpk_pso_layout)
struct PsoLayout {
    size_t scalars, lower, upper, vmax, pos, vel, pbest_pos, pbest_sc, gbest_pos, scores, expected, d_sigma, measured_of,
           log_best, log_stall, log_restart, trace_pos, trace_sc, rng, total;
};
PsoLayout pso_layout(int P, int d, int S, int n_points, int max_iter, int max_evals, long long rng_len);
int launch_pso_step(unsigned char* state, const PsoLayout& L, int P, int S, const int* counts, const int* info,
                    const unsigned char* fallback, double* pos_out, unsigned char* active_out, cudaStream_t st);

// gpk_kmeans.cu -- E-step and cluster sums of one Lloyd iteration (include/gpk.h gpk_kmeans_step)
int launch_kmeans_step(const double* X, int N, int d, const double* centers, const double* csq, int k, int* labels,
                       int* changed, double* sums, double* counts, cudaStream_t st);

// gpk_surface.cu
int launch_prob_surface(const double* table, long long R, int ncols, int npoints,
                        double* y_out, double* p_out, int* valid_out, cudaStream_t st);

int launch_surface_scan(const int* valid, long long R, long long* dst, long long* count, cudaStream_t st);
int launch_surface_gather(const int* valid, const long long* dst, long long R, int npoints, const double* y, const double* p,
                          double* yc, double* pc, long long* idx, cudaStream_t st);
int launch_sum_gaussians(const double* y, int n, const double* gaus, int ng, double half_dy, double* p_out,
                         cudaStream_t st);

}  // namespace gpk
