// Register-resident Cholesky + triangular inverse of one n x n block (n <= 128), used by the diagonal-block
// kernel of the blocked factorisation and by the one-CTA small-N loss kernel (256 threads).
//
// The 128 x 128 working array E is spread over the CTA as 16 x 16 register tiles of 8 x 8 doubles
// (thread t owns rows 8*(t/16).., columns 8*(t%16)..).  Lower triangle incl. diagonal: A on entry, L on exit.
// Strictly upper part: X = L^-1 stored transposed, E(r,c) = X[c][r] for c > r (zero on entry); the diagonal
// of X (1 / L_jj) goes to xd[].  X is produced by running the same forward elimination on the identity.
//
// Step j (right-looking): the threads owning column j scale it by 1/sqrt(E(j,j)) and publish it three times:
//   v[r]   = the scaled column (L part for r > j, X part for r < j, 1/L_jj at r = j),
//   vcm[c] = v[c] for c > j, else 0      (columns <= j are final; stored as [c % 8][c / 8] so that the 16 tile
//                                         columns of a warp read 16 consecutive doubles: no bank conflicts),
//   vup[r] = v[r] for r <= j, else 0     (rows of the X part that have not started yet must stay zero).
// Every element with column c > j that is either in the lower triangle (c <= r) or in the X part with r <= j then
// receives E(r,c) -= v[r] * v[c]: with the masks folded into the published vectors this is 64 unconditional FMAs per
// thread, E(r,c) -= (c <= r ? v : vup)[r] * vcm[c], the row source being a per-thread constant for tiles off the
// diagonal (v below it, vup above it) and a compile-time choice inside a diagonal tile.  Same products, same order and
// same rounding as the element-wise predicated form (a masked term adds an exact zero).  sqrt and the reciprocal of
// the next pivot (both correctly rounded, as sqrt() and 1.0 / x are) are computed by the one thread that owns it.
// All operands of a step come from registers plus 24 broadcast loads; two block barriers per step.
#pragma once
#include "gpk_common.cuh"

namespace gpk {

constexpr int CI_LD = NB + 1;    // row stride of the shared staging array used around the register kernel
constexpr int CI_WORK = 3 * NB + 8;      // doubles of shared scratch chol_inv_regs needs (v, vcm, vup, pivot slots)
constexpr size_t CI_SMEM_BYTES = (size_t)NB * CI_LD * sizeof(double) + (size_t)(CI_WORK + NB) * sizeof(double);

// All 256 threads must call.  work: CI_WORK doubles, xd: NB doubles, both in shared memory.
// Returns 0 or (1 + index of the first non-positive / NaN pivot), identical in every thread.
__device__ __forceinline__ int chol_inv_regs(double (&E)[8][8], int n, double* work, double* xd) {
    double* v = work;
    double* vcm = work + NB;
    double* vup = work + 2 * NB;
    double* piv = work + 3 * NB;            // two slots of (pivot, sqrt, 1 / sqrt)
    const int tid = threadIdx.x;
    const int tr = tid >> 4, tc = tid & 15;
    const int R0 = tr * 8, C0 = tc * 8;
    const double* row_lo = (tc > tr) ? vup : v;          // row multiplier of the elements with c <= r
    const double* row_up = (tc < tr) ? v : vup;          // ... and of those with c > r
    int bad = 0;
    if (tid == 0) {
        const double d0 = E[0][0], s0 = __dsqrt_rn(d0);
        piv[0] = d0; piv[1] = s0; piv[2] = __drcp_rn(s0);
    }
    __syncthreads();
    // steps are taken 8 at a time so that the column inside the owning tile (jj) is a compile-time index:
    // the register tile must never be indexed dynamically (it would be demoted to local memory)
    for (int jb = 0; jb * 8 < n; ++jb) {
#pragma unroll
        for (int jj = 0; jj < 8; ++jj) {
            const int j = jb * 8 + jj;
            if (j < n) {                               // uniform over the CTA
                const double* slot = piv + 3 * (j & 1);
                if (!(slot[0] > 0.0) && bad == 0) bad = j + 1;
                if (jb == tc) {                        // this thread owns part of column j
                    const double s = slot[1], inv = slot[2];
#pragma unroll
                    for (int rr = 0; rr < 8; ++rr) {
                        const int r = R0 + rr;
                        double t;
                        if (r == j) {
                            E[rr][jj] = s;
                            t = inv;
                            xd[r] = inv;
                        } else {
                            t = E[rr][jj] * inv;
                            E[rr][jj] = t;
                        }
                        v[r] = t;
                        vcm[rr * 16 + tr] = (r > j) ? t : 0.0;      // [r % 8][r / 8]: see the load below
                        vup[r] = (r <= j) ? t : 0.0;
                    }
                }
                __syncthreads();
                // tiles left of column j are final; tiles above the diagonal whose rows are all > j have not started
                if (C0 + 7 > j && !(tc > tr && R0 > j)) {
                    double lo[8], up[8], vc[8];
#pragma unroll
                    for (int k = 0; k < 8; ++k) { lo[k] = row_lo[R0 + k]; up[k] = row_up[R0 + k]; vc[k] = vcm[k * 16 + tc]; }
                    // the owner of the next pivot updates that element first and publishes it with its square root and
                    // reciprocal, so that the serial chain of the factorisation does not wait for its other 63 products
                    const int jn = j + 1;
                    if ((jn >> 3) == tr && tr == tc && jn < n) {
                        const int q = (jj + 1) & 7;                 // compile-time after unrolling
                        const double dn = fma(-lo[q], vc[q], E[q][q]);
                        const double sn = __dsqrt_rn(dn);
                        double* nslot = piv + 3 * (jn & 1);
                        nslot[0] = dn; nslot[1] = sn; nslot[2] = __drcp_rn(sn);
                    }
#pragma unroll
                    for (int rr = 0; rr < 8; ++rr)
#pragma unroll
                        for (int cc = 0; cc < 8; ++cc)
                            E[rr][cc] = fma(-((cc <= rr) ? lo[rr] : up[rr]), vc[cc], E[rr][cc]);
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
