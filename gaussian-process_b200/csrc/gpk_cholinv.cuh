// Register-resident Cholesky + triangular inverse of one n x n block (n <= 128), used by the diagonal-block
// kernel of the blocked factorisation and by the one-CTA small-N loss kernel (256 threads).
//
// The 128 x 128 working array E is spread over the CTA as 16 x 16 register tiles of 8 x 8 doubles
// (thread t owns rows 8*(t/16).., columns 8*(t%16)..).  Lower triangle incl. diagonal: A on entry, L on exit.
// Strictly upper part: X = L^-1 stored transposed, E(r,c) = X[c][r] for c > r (zero on entry); the diagonal
// of X (1 / L_jj) goes to xd[].  X is produced by running the same forward elimination on the identity.
//
// Step j (right-looking): the threads owning column j scale it by 1/sqrt(E(j,j)) and publish it as v[];
// then every element with column c > j that is either in the lower triangle (c <= r) or in the X part with
// r <= j receives E(r,c) -= v[r] * v[c]  (trailing Cholesky update and X update are the same product).
// All operands of a step come from registers plus 16 broadcast loads of v[]; two block barriers per step.
#pragma once
#include "gpk_common.cuh"

namespace gpk {

constexpr int CI_LD = NB + 1;    // row stride of the shared staging array used around the register kernel
constexpr size_t CI_SMEM_BYTES = (size_t)NB * CI_LD * sizeof(double) + 3 * NB * sizeof(double);

// All 256 threads must call.  v, xd: NB doubles each in shared memory; piv: 2 doubles in shared memory.
// Returns 0 or (1 + index of the first non-positive / NaN pivot), identical in every thread.
__device__ __forceinline__ int chol_inv_regs(double (&E)[8][8], int n, double* v,
                                             double* xd, double* piv) {
    const int tid = threadIdx.x;
    const int R0 = (tid >> 4) * 8, C0 = (tid & 15) * 8;
    int bad = 0;
    if (tid == 0) piv[0] = E[0][0];
    __syncthreads();
    // steps are taken 8 at a time so that the column inside the owning tile (jj) is a compile-time index:
    // the register tile must never be indexed dynamically (it would be demoted to local memory)
    for (int jb = 0; jb * 8 < n; ++jb) {
#pragma unroll
        for (int jj = 0; jj < 8; ++jj) {
            const int j = jb * 8 + jj;
            if (j < n) {                               // uniform over the CTA
                const double dj = piv[j & 1];
                if (!(dj > 0.0) && bad == 0) bad = j + 1;
                const double s = sqrt(dj);
                const double inv = 1.0 / s;
                if (jb == (tid & 15)) {                // this thread owns part of column j
#pragma unroll
                    for (int rr = 0; rr < 8; ++rr) {
                        const int r = R0 + rr;
                        if (r == j) {
                            E[rr][jj] = s;
                            v[r] = inv;
                            xd[r] = inv;
                        } else {
                            const double t = E[rr][jj] * inv;
                            E[rr][jj] = t;
                            v[r] = t;
                        }
                    }
                }
                __syncthreads();
                if (C0 + 7 > j) {                      // tiles left of / at column j are final
                    double vr[8], vc[8];
#pragma unroll
                    for (int k = 0; k < 8; ++k) { vr[k] = v[R0 + k]; vc[k] = v[C0 + k]; }
#pragma unroll
                    for (int rr = 0; rr < 8; ++rr) {
                        const int r = R0 + rr;
#pragma unroll
                        for (int cc = 0; cc < 8; ++cc) {
                            const int c = C0 + cc;
                            if (c > j && (c <= r || r <= j)) E[rr][cc] = fma(-vr[rr], vc[cc], E[rr][cc]);
                        }
                    }
                }
                // owner of the next pivot publishes it (after its update)
                if (jj < 7) {
                    if (jb == (tid >> 4) && jb == (tid & 15)) piv[(j + 1) & 1] = E[(jj + 1) & 7][(jj + 1) & 7];
                } else {
                    if (jb + 1 == (tid >> 4) && jb + 1 == (tid & 15)) piv[(j + 1) & 1] = E[0][0];
                }
                __syncthreads();
            }
        }
    }
    return bad;
}

// Dump the register tiles into the shared staging array in the layout the consumers read:
//   L[r][c] (c <= r) at Lm[r][c];   X[i][c] (c < i) at Lm[c][i + 1];   X[i][i] at Lm[i][i + 1].
__device__ __forceinline__ void chol_inv_dump(const double (&E)[8][8], const double* __restrict__ xd,
                                              double* __restrict__ Lm) {
    const int tid = threadIdx.x;
    const int R0 = (tid >> 4) * 8, C0 = (tid & 15) * 8;
#pragma unroll
    for (int rr = 0; rr < 8; ++rr) {
        const int r = R0 + rr;
#pragma unroll
        for (int cc = 0; cc < 8; ++cc) {
            const int c = C0 + cc;
            if (c <= r) Lm[r * CI_LD + c] = E[rr][cc];
            else Lm[r * CI_LD + c + 1] = E[rr][cc];        // X[c][r] -> Lm[r][c + 1]
        }
    }
    if (tid < NB) Lm[tid * CI_LD + tid + 1] = xd[tid];
}

}  // namespace gpk
