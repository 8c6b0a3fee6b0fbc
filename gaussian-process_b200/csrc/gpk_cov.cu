// Kernel (a): ARD-RBF covariance builders.
// Reference arithmetic: GP_func.py:49-65 (kernel_func) and :21 (+ diag(e^2)).
//   xs = x / l ; sq_i = sum_d xs_i^2 ; dist2 = (sq_i + sq_j) - 2 * dot(xs_i, xs_j) ; clamp >= 0 ;
//   K = exp(-0.5 * dist2).
// The expanded form and the operation order of the reference are kept on purpose (its rounding
// is part of what the parity tests compare against); products and sums that NumPy evaluates as
// separate ufuncs use the non-contracting intrinsics, the dot product follows the BLAS FMA chain.
#include "gpk_common.cuh"

namespace gpk {

__device__ __forceinline__ double rbf_from_parts(double sqi, double sqj, double dot) {
    double s = __dsub_rn(__dadd_rn(sqi, sqj), __dmul_rn(2.0, dot));
    s = fmax(s, 0.0);
    return exp(__dmul_rn(-0.5, s));
}

// ---------------------------------------------------------------------------------------------
// xs[p][dd][i] = X[dd][i] / ls[p][dd]   (0 in the padding), sq[p][i] = sum_dd xs^2
__global__ void k_scale_points(const double* __restrict__ X, const double* __restrict__ ls, int N, int Np,
                               int d, double* __restrict__ xs, double* __restrict__ sq,
                               const unsigned char* __restrict__ active) {
    const int p = blockIdx.y;
    if (active && !active[p]) return;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= Np) return;
    double acc = 0.0;
    for (int dd = 0; dd < d; ++dd) {
        double v = 0.0;
        if (i < N) v = __ddiv_rn(X[(size_t)dd * N + i], ls[(size_t)p * d + dd]);
        xs[((size_t)p * d + dd) * Np + i] = v;
        const double v2 = __dmul_rn(v, v);
        acc = (dd == 0) ? v2 : __dadd_rn(acc, v2);
    }
    sq[(size_t)p * Np + i] = acc;
}

int launch_scale_points(const double* X, const double* ls, int P, int N, int Np, int d, double* xs,
                        double* sq, const unsigned char* active, cudaStream_t st) {
    dim3 grid((Np + 255) / 256, P);
    k_scale_points<<<grid, 256, 0, st>>>(X, ls, N, Np, d, xs, sq, active);
    return check_launch("k_scale_points");
}

// rows r = (threadIdx.x >> 7), +2, ... of tile (I, .) for column j: K_ij + delta_ij e_i^2, identity in the padding.
// PROBE: the tile's contributions to u = K v (v = probe_v, the residual probe of include/gpk.h) are formed on the way:
// the column sum sum_i K_ij v_i of this thread's 64 rows stays in a register (returned); the row sums sum_j K_ij v_j need
// all 128 columns = threads of a row and go through shared memory in batches of 8 rows (s_prod, then one warp per row
// adds 4 columns per lane and finishes with 5 shuffle steps) -> s_rowsum[r].  Keeping the 64 products of a thread in registers
// for a warp butterfly instead costs 66 registers and halves the occupancy of this FP64-pipe-bound kernel (measured: 2x time).
constexpr int CP_ROWS = 8;               // rows per batch (two staging buffers of CP_ROWS x 129 doubles: 16.5 KB, 5 CTAs per SM)
constexpr int CP_PITCH = NB + 1;
template <int D, bool PROBE>
__device__ __forceinline__ double cov_lower_rows(const double* __restrict__ xsp, const double* __restrict__ sqp,
                                                 const double* __restrict__ e, int N, int Np, int I, int j,
                                                 const double (*s_xi)[NB], const double* s_sqi, double* __restrict__ out,
                                                 double* s_prod, double* s_rowsum, const double* s_pvi) {
    double xj[D];
#pragma unroll
    for (int dd = 0; dd < D; ++dd) xj[dd] = xsp[(size_t)dd * Np + j];
    const double sqj = sqp[j];
    const double vj = (PROBE && j < N) ? probe_v(j) : 0.0;
    const int h = threadIdx.x >> 7, c = threadIdx.x & (NB - 1);
    double colsum = 0.0;
    for (int b = 0; b < NB / CP_ROWS; ++b) {
#pragma unroll
        for (int k = 0; k < CP_ROWS / 2; ++k) {
            const int lr = h + 2 * k;
            const int r = b * CP_ROWS + lr;
            const int i = I * NB + r;
            double v;
            if (i < N && j < N) {
                double dot = __dmul_rn(s_xi[0][r], xj[0]);
#pragma unroll
                for (int dd = 1; dd < D; ++dd) dot = fma(s_xi[dd][r], xj[dd], dot);
                v = rbf_from_parts(s_sqi[r], sqj, dot);
                if (i == j) { const double ee = e[i]; v = __dadd_rn(v, __dmul_rn(ee, ee)); }
            } else {
                v = (i == j) ? 1.0 : 0.0;
            }
            out[(size_t)i * Np + j] = v;
            if (PROBE) {
                s_prod[(b & 1) * CP_ROWS * CP_PITCH + lr * CP_PITCH + c] = v * vj;
                colsum = fma(v, s_pvi[r], colsum);
            }
        }
        if (PROBE) {
            // two staging buffers: one barrier per batch (the barrier of batch b + 1 separates these reads from the
            // writes of batch b + 2 into the same buffer)
            __syncthreads();
            const int rr = threadIdx.x >> 5, q = threadIdx.x & 31;          // one warp per row of the batch
            const double* src = s_prod + (b & 1) * CP_ROWS * CP_PITCH + rr * CP_PITCH + q;
            double a = 0.0;
#pragma unroll
            for (int m = 0; m < NB / 32; ++m) a += src[32 * m];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
            if (q == 0) s_rowsum[b * CP_ROWS + rr] = a;
        }
    }
    return colsum;
}

// ---------------------------------------------------------------------------------------------
// Lower block triangle of K + diag(e^2) for every particle, written straight into the factorisation
// buffer.  One CTA per 128x128 tile (I >= J); padding rows/cols become the identity.
// kv != nullptr ([P][T][Np]): slot [B][i] receives the contribution to (K v)_i of the tile that pairs row block of i with
// block B -- tile (I, J) writes its row sums to [J][rows of I] and, for I > J, its column sums (K is symmetric) to
// [I][columns of J]: every slot is written exactly once, k_probe_sum adds them in a fixed order.
template <bool PROBE>
__global__ void __launch_bounds__(256, 5) k_cov_lower(const double* __restrict__ xs, const double* __restrict__ sq,
                                                   const double* __restrict__ e, int N, int Np, int d,
                                                   double* __restrict__ mat, double* __restrict__ kv,
                                                   const unsigned char* __restrict__ active) {
    const int p = blockIdx.y;
    if (active && !active[p]) return;
    // decode tile index -> (I, J), I >= J
    const int t = blockIdx.x;
    int I = (int)((sqrt(8.0 * t + 1.0) - 1.0) * 0.5);
    while ((I + 1) * (I + 2) / 2 <= t) ++I;
    while (I * (I + 1) / 2 > t) --I;
    const int J = t - I * (I + 1) / 2;

    __shared__ double s_xi[MAX_D][NB];
    __shared__ double s_sqi[NB];
    __shared__ double s_prod[PROBE ? 2 * CP_ROWS * CP_PITCH : 1];
    __shared__ double s_rowsum[PROBE ? NB : 1];
    __shared__ double s_col[PROBE ? 2 : 1][NB];
    __shared__ double s_pvi[PROBE ? NB : 1];
    const double* xsp = xs + (size_t)p * d * Np;
    const double* sqp = sq + (size_t)p * Np;
    for (int k = threadIdx.x; k < d * NB; k += blockDim.x) {
        const int dd = k / NB, r = k % NB;
        s_xi[dd][r] = xsp[(size_t)dd * Np + I * NB + r];
    }
    if (threadIdx.x < NB) {
        s_sqi[threadIdx.x] = sqp[I * NB + threadIdx.x];
        if (PROBE) s_pvi[threadIdx.x] = (I * NB + threadIdx.x < N) ? probe_v(I * NB + threadIdx.x) : 0.0;
    }
    __syncthreads();

    const int c = threadIdx.x & (NB - 1);
    const int j = J * NB + c;
    double* out = mat + (size_t)p * Np * Np;
    // the dimension loop is unrolled for the actual d (a run-time bound costs MAX_D predicated load / FMA pairs per
    // element in issue slots); same operations in the same order for every D
    double colsum;
    switch (d) {
        case 1: colsum = cov_lower_rows<1, PROBE>(xsp, sqp, e, N, Np, I, j, s_xi, s_sqi, out, s_prod, s_rowsum, s_pvi); break;
        case 2: colsum = cov_lower_rows<2, PROBE>(xsp, sqp, e, N, Np, I, j, s_xi, s_sqi, out, s_prod, s_rowsum, s_pvi); break;
        case 3: colsum = cov_lower_rows<3, PROBE>(xsp, sqp, e, N, Np, I, j, s_xi, s_sqi, out, s_prod, s_rowsum, s_pvi); break;
        case 4: colsum = cov_lower_rows<4, PROBE>(xsp, sqp, e, N, Np, I, j, s_xi, s_sqi, out, s_prod, s_rowsum, s_pvi); break;
        case 5: colsum = cov_lower_rows<5, PROBE>(xsp, sqp, e, N, Np, I, j, s_xi, s_sqi, out, s_prod, s_rowsum, s_pvi); break;
        case 6: colsum = cov_lower_rows<6, PROBE>(xsp, sqp, e, N, Np, I, j, s_xi, s_sqi, out, s_prod, s_rowsum, s_pvi); break;
        case 7: colsum = cov_lower_rows<7, PROBE>(xsp, sqp, e, N, Np, I, j, s_xi, s_sqi, out, s_prod, s_rowsum, s_pvi); break;
        default: colsum = cov_lower_rows<8, PROBE>(xsp, sqp, e, N, Np, I, j, s_xi, s_sqi, out, s_prod, s_rowsum, s_pvi); break;
    }
    if (PROBE) {
        s_col[threadIdx.x >> 7][c] = colsum;
        __syncthreads();
        if (threadIdx.x < NB) {
            const int r = threadIdx.x;
            const int T = Np / NB;
            double* slot = kv + (size_t)p * T * Np;
            slot[(size_t)J * Np + I * NB + r] = s_rowsum[r];
            if (I > J) slot[(size_t)I * Np + J * NB + r] = s_col[0][r] + s_col[1][r];
        }
    }
}

int launch_cov_lower(const double* xs, const double* sq, const double* e, int P, int N, int Np, int d,
                     double* mat, double* kv, const unsigned char* active, cudaStream_t st) {
    const int T = Np / NB;
    dim3 grid(T * (T + 1) / 2, P);
    if (kv) k_cov_lower<true><<<grid, 256, 0, st>>>(xs, sq, e, N, Np, d, mat, kv, active);
    else k_cov_lower<false><<<grid, 256, 0, st>>>(xs, sq, e, N, Np, d, mat, kv, active);
    return check_launch("k_cov_lower");
}

// ---------------------------------------------------------------------------------------------
// Stand-alone kernel_func(x1, x2, l) for P length-scale vectors: K[p][i][j], i over x1, j over x2.
// e != nullptr adds e_i^2 on the diagonal (only meaningful when x1 is x2).
__global__ void __launch_bounds__(256) k_cov_full(const double* __restrict__ X1, const double* __restrict__ X2,
                                                  const double* __restrict__ e, const double* __restrict__ ls,
                                                  int N1, int N2, int d, double* __restrict__ K, long long ldk,
                                                  long long strideK) {
    const int p = blockIdx.z;
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int i0 = blockIdx.y * 32;
    double l[MAX_D], xj[MAX_D];
    double sqj = 0.0;
#pragma unroll
    for (int dd = 0; dd < MAX_D; ++dd) {
        l[dd] = (dd < d) ? ls[(size_t)p * d + dd] : 1.0;
        xj[dd] = 0.0;
        if (dd < d && j < N2) {
            xj[dd] = __ddiv_rn(X2[(size_t)dd * N2 + j], l[dd]);
            const double v2 = __dmul_rn(xj[dd], xj[dd]);
            sqj = (dd == 0) ? v2 : __dadd_rn(sqj, v2);
        }
    }
    if (j >= N2) return;
    for (int r = 0; r < 32; ++r) {
        const int i = i0 + r;
        if (i >= N1) break;
        double sqi = 0.0, dot = 0.0;
#pragma unroll
        for (int dd = 0; dd < MAX_D; ++dd) {
            if (dd < d) {
                const double xi = __ddiv_rn(X1[(size_t)dd * N1 + i], l[dd]);
                const double v2 = __dmul_rn(xi, xi);
                sqi = (dd == 0) ? v2 : __dadd_rn(sqi, v2);
                dot = (dd == 0) ? __dmul_rn(xi, xj[0]) : fma(xi, xj[dd], dot);
            }
        }
        double v = rbf_from_parts(sqi, sqj, dot);
        if (e != nullptr && i == j) { const double ee = e[i]; v = __dadd_rn(v, __dmul_rn(ee, ee)); }
        K[(size_t)p * strideK + (size_t)i * ldk + j] = v;
    }
}

int launch_cov_full(const double* X1, const double* X2, const double* e, const double* ls, int P, int N1,
                    int N2, int d, double* K, long long ldk, long long strideK, cudaStream_t st) {
    dim3 grid((N2 + 255) / 256, (N1 + 31) / 32, P);
    k_cov_full<<<grid, 256, 0, st>>>(X1, X2, e, ls, N1, N2, d, K, ldk, strideK);
    return check_launch("k_cov_full");
}

// ---------------------------------------------------------------------------------------------
// Prediction-side cross covariance, transposed: KsT[b][j] = k(x_j, xfit_b)  (GP_func.py:32), zero in
// the padding, with the predictive mean mu_b = sum_j KsT[b][j] * alpha_j fused in (GP_func.py:36).
// One warp per grid point, lanes sweep j (coalesced 256-byte stores).
__global__ void __launch_bounds__(256) k_cross_cov(const double* __restrict__ xs, const double* __restrict__ sq,
                                                   const double* __restrict__ ls, const double* __restrict__ alpha,
                                                   int N, int Np, int d, const double* __restrict__ Xfit,
                                                   long long M, long long m0, int B, int Bp,
                                                   double* __restrict__ KsT, double* __restrict__ mu_out) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int b = blockIdx.x * 8 + warp;
    if (b >= Bp) return;
    double* row = KsT + (size_t)b * Np;
    if (b >= B) {   // padding points: zero rows
        for (int j = lane; j < Np; j += 32) row[j] = 0.0;
        return;
    }
    double xb[MAX_D];
    double sqb = 0.0;
#pragma unroll
    for (int dd = 0; dd < MAX_D; ++dd) {
        xb[dd] = 0.0;
        if (dd < d) {
            xb[dd] = __ddiv_rn(Xfit[(size_t)dd * M + m0 + b], ls[dd]);
            const double v2 = __dmul_rn(xb[dd], xb[dd]);
            sqb = (dd == 0) ? v2 : __dadd_rn(sqb, v2);
        }
    }
    double acc = 0.0;
    for (int j = lane; j < Np; j += 32) {
        double v = 0.0;
        if (j < N) {
            double dot = __dmul_rn(xs[j], xb[0]);
#pragma unroll
            for (int dd = 1; dd < MAX_D; ++dd)
                if (dd < d) dot = fma(xs[(size_t)dd * Np + j], xb[dd], dot);
            v = rbf_from_parts(sq[j], sqb, dot);
            acc = fma(v, alpha[j], acc);
        }
        row[j] = v;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) mu_out[m0 + b] = acc;
}

int launch_cross_cov(const double* xs, const double* sq, const double* ls, const double* alpha, int N, int Np,
                     int d, const double* Xfit, long long M, long long m0, int B, int Bp, double* KsT,
                     double* mu_out, cudaStream_t st) {
    dim3 grid((Bp + 7) / 8);
    k_cross_cov<<<grid, 256, 0, st>>>(xs, sq, ls, alpha, N, Np, d, Xfit, M, m0, B, Bp, KsT, mu_out);
    return check_launch("k_cross_cov");
}

}  // namespace gpk
