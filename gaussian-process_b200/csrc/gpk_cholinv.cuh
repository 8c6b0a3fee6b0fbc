// Shared-memory resident Cholesky + triangular inverse of one n x n block (n <= 128), used by the
// diagonal-block kernel of the blocked factorisation and by the one-CTA small-N loss kernel.
//
// Storage: one 128 x 129 array `Lm` (row stride CI_LD).  The lower triangle (incl. diagonal) holds
// A on entry and L on exit.  X = L^-1 (lower triangular) is built alongside by running the same
// forward elimination on the identity:  X[i][c] (c <= i) lives at Lm[c][i + 1], i.e. transposed in the
// strictly-upper part shifted one column right.  The caller initialises that part to the identity
// (Lm[c][c+1] = 1, Lm[c][i+1] = 0 for i > c).
//
// Right-looking, one column per step, two block barriers per step; the rank-1 updates of the trailing
// triangle and of X are spread over the 8 warps (n "rows" of work per step in total).
#pragma once
#include "gpk_common.cuh"

namespace gpk {

constexpr int CI_LD = NB + 1;
constexpr size_t CI_SMEM_BYTES = (size_t)NB * CI_LD * sizeof(double) + 2 * NB * sizeof(double);

// All 256 threads of the CTA must call.  col/xr: NB doubles each.  Returns 0 or (1 + index of the
// first non-positive / NaN pivot), identical in every thread.
__device__ __forceinline__ int chol_inv_inplace(double* __restrict__ Lm, int n, double* __restrict__ col,
                                                double* __restrict__ xr) {
    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    int bad = 0;
    for (int j = 0; j < n; ++j) {
        const double dj = Lm[j * CI_LD + j];
        if (!(dj > 0.0) && bad == 0) bad = j + 1;
        const double s = sqrt(dj);
        const double inv = 1.0 / s;
        if (tid < NB) {
            const int i = tid;
            if (i > j && i < n) {
                const double v = Lm[i * CI_LD + j] * inv;
                col[i] = v;
                Lm[i * CI_LD + j] = v;
            }
        } else {
            const int c = tid - NB;
            if (c <= j) {
                const double xv = Lm[c * CI_LD + j + 1] * inv;
                xr[c] = xv;
                Lm[c * CI_LD + j + 1] = xv;
            }
        }
        __syncthreads();
        if (tid == 0) Lm[j * CI_LD + j] = s;
        // trailing triangle: rows i in (j, n), columns k in (j, i]
        for (int i = j + 1 + warp; i < n; i += 8) {
            const double ci = -col[i];
            double* row = Lm + i * CI_LD;
            for (int k = j + 1 + lane; k <= i; k += 32) row[k] = fma(ci, col[k], row[k]);
        }
        // X rows c in [0, j]: X[i][c] -= L[i][j] * X[j][c] for i in (j, n)
        for (int c = warp; c <= j; c += 8) {
            const double xc = -xr[c];
            double* row = Lm + c * CI_LD + 1;
            for (int i = j + 1 + lane; i < n; i += 32) row[i] = fma(xc, col[i], row[i]);
        }
        __syncthreads();
    }
    return bad;
}

}  // namespace gpk
