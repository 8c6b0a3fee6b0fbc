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
// All operands of a step come from registers plus 24 broadcast loads; one block barrier per step (see chol_inv_regs).
#pragma once
#include "gpk_common.cuh"

namespace gpk {

constexpr int CI_LD = NB + 1;    // row stride of the shared staging array used around the register kernel
// doubles of shared scratch chol_inv_regs needs: two generations of (v, vcm, vup), the initial diagonal, two pivot slots
constexpr int CI_WORK = 7 * NB + 8;
constexpr size_t CI_SMEM_BYTES = (size_t)NB * CI_LD * sizeof(double) + (size_t)(CI_WORK + NB) * sizeof(double);

// All 256 threads must call.  work: CI_WORK doubles, xd: NB doubles, both in shared memory.
// Returns 0 or (1 + index of the first non-positive / NaN pivot), identical in every thread.
//
// One block barrier per step: the published vectors alternate between two generations (step j writes generation j & 1,
// which was last read in step j - 2, i.e. before the barrier of step j - 1 that every writer has passed).  The hand-over
// of the pivot between the thread that owns E(j,j) and the 16 threads that own column j costs no barrier either: every
// thread tracks the pivots of its own 8 tile columns, pv[c] = A(c,c) - sum_{m<j} v_m[c]^2, with the very FMAs the
// diagonal tile applies to E(c,c) (same operands, same order: bit-identical), so the owners of column j + 1 start its
// square root and reciprocal right after the step-j loads, ahead of their 64 update FMAs.
__device__ __forceinline__ int chol_inv_regs(double (&E)[8][8], int n, double* work, double* xd) {
    double* dgs = work + 6 * NB;             // initial diagonal
    double* piv = work + 7 * NB;             // pivot of step j at piv[j & 1] (for the positivity check of all threads)
    const int tid = threadIdx.x;
    const int tr = tid >> 4, tc = tid & 15;
    const int R0 = tr * 8, C0 = tc * 8;
    int bad = 0;
    if (tr == tc) {
#pragma unroll
        for (int k = 0; k < 8; ++k) dgs[R0 + k] = E[k][k];
    }
    __syncthreads();
    double pv[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) pv[k] = dgs[C0 + k];
    // square root and reciprocal of the pivot of the column this thread scales next (valid in the owners of that column)
    double s_next = 0.0, inv_next = 0.0;
    if (tc == 0) { s_next = __dsqrt_rn(pv[0]); inv_next = __drcp_rn(s_next); }
    // steps are taken 8 at a time so that the column inside the owning tile (jj) is a compile-time index:
    // the register tile must never be indexed dynamically (it would be demoted to local memory)
    for (int jb = 0; jb * 8 < n; ++jb) {
#pragma unroll
        for (int jj = 0; jj < 8; ++jj) {
            const int j = jb * 8 + jj;
            if (j < n) {                               // uniform over the CTA
                double* v = work + (j & 1) * 3 * NB;
                double* vcm = v + NB;
                double* vup = v + 2 * NB;
                const double* row_lo = (tc > tr) ? vup : v;          // row multiplier of the elements with c <= r
                const double* row_up = (tc < tr) ? v : vup;          // ... and of those with c > r
                if (jb == tc) {                        // this thread owns part of column j
                    const double s = s_next, inv = inv_next;
                    if (tr == 0) piv[j & 1] = pv[jj];
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
                if (!(piv[j & 1] > 0.0) && bad == 0) bad = j + 1;
                if (C0 + 7 > j) {                      // columns <= j are final
                    double vc[8];
#pragma unroll
                    for (int k = 0; k < 8; ++k) vc[k] = vcm[k * 16 + tc];
                    // the thread's copy of its columns' pivots, then (owners of column j + 1) that pivot's root and
                    // reciprocal, so that the serial chain of the factorisation does not wait for the 64 products below
#pragma unroll
                    for (int k = 0; k < 8; ++k) pv[k] = fma(-vc[k], vc[k], pv[k]);
                    const int jn = j + 1;
                    if ((jn >> 3) == tc && jn < n) {
                        const double dn = pv[(jj + 1) & 7];           // compile-time index after unrolling
#ifdef CI_EXP_NOSQRT                          /* timing experiment only: no square-root / reciprocal chain */
                        s_next = dn;
                        inv_next = 0.5;
#else
                        s_next = __dsqrt_rn(dn);
                        inv_next = __drcp_rn(s_next);
#endif
                    }
                    // tiles above the diagonal whose rows are all > j have not started
#ifdef CI_EXP_NOFMA                           /* timing experiment only: no trailing update */
                    if (false) {
#else
                    if (!(tc > tr && R0 > j)) {
#endif
                        double lo[8], up[8];
#pragma unroll
                        for (int k = 0; k < 8; ++k) { lo[k] = row_lo[R0 + k]; up[k] = row_up[R0 + k]; }
#pragma unroll
                        for (int rr = 0; rr < 8; ++rr)
#pragma unroll
                            for (int cc = 0; cc < 8; ++cc)
                                E[rr][cc] = fma(-((cc <= rr) ? lo[rr] : up[rr]), vc[cc], E[rr][cc]);
                    }
                }
            }
        }
    }
    __syncthreads();                                   // the callers reuse the scratch and read xd
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
