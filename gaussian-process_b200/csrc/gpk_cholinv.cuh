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
// root and reciprocal root of a pivot.  FAST = false: the correctly rounded sqrt() and 1.0 / x of the reference's LAPACK path
// (two dependent software sequences, ~530 cycles: 36 of the 106 us of a 128-column block, profiles/r01 knock-out timing).
// FAST = true: one reciprocal-square-root sequence (rsqrt, <= 1 ulp) and one Newton correction each for the root and its
// reciprocal (results within 1 ulp of the correctly rounded ones, ~3x shorter dependency chain); used by the diagonal blocks
// of the blocked path, not by the one-CTA loss whose seeded trajectories are pinned bit for bit.
template <bool FAST>
__device__ __forceinline__ void pivot_root(double piv, double& s, double& inv) {
    if (FAST) {
        const double r = rsqrt(piv);
        const double s0 = piv * r;
        s = fma(fma(-s0, s0, piv), 0.5 * r, s0);
        inv = fma(fma(-s, r, 1.0), r, r);
    } else {
        s = __dsqrt_rn(piv);
        inv = __drcp_rn(s);
    }
}

template <bool FAST = false>
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
    if (tc == 0) pivot_root<FAST>(pv[0], s_next, inv_next);
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
                        pivot_root<FAST>(dn, s_next, inv_next);
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

// ---------------------------------------------------------------------------------------------------------------------
// Thread -> tile maps.  MAP 0 (chol_inv_regs, chol_inv_blocked): tile row = tid / 16, tile column = tid % 16.  MAP 1
// (chol_inv_lookahead): the 16 tiles of a tile column belong to ONE warp -- tile column = 2 * warp + lane / 16, tile row =
// lane % 16 -- so that everything only the owners of a column do (scaling, pivot root, publishing) is executed by one warp
// instead of by two lanes of each of the eight.
template <int MAP>
__device__ __forceinline__ void ci_tile_of(int tid, int& tr, int& tc) {
    if (MAP == 0) { tr = tid >> 4; tc = tid & 15; }
    else { tr = tid & 15; tc = ((tid >> 5) << 1) + ((tid >> 4) & 1); }
}

// Look-ahead form of chol_inv_regs: same contract, same operations on every element in the same order (bit-identical
// results), different schedule.  ncu source counters of chol_inv_regs (profiles/r02_ncu_diag_source.md): of 242 warp
// instructions per warp and column only 64 are the update FMAs; 80 are the publish block and 30 the pivot's root and
// reciprocal, executed by all eight warps with two active lanes each, on the critical path before the block barrier.  Here
//   * the owners of a column are one warp (MAP 1): the other seven skip that code;
//   * the owners of column j + 1 apply step j to that column only, take the root of its pivot and publish it (into the other
//     generation of the published vectors) BEFORE their remaining 56 FMAs of step j: when the other warps arrive at the next
//     barrier the column is there.
// The published vectors are stored as [r % 8][r / 8] so that the 16 tile rows of a warp read and write consecutive doubles.
template <bool FAST = false>
__device__ __forceinline__ int chol_inv_lookahead(double (&E)[8][8], int n, double* work, double* xd) {
    double* dgs = work + 6 * NB;             // initial diagonal
    double* piv = work + 7 * NB;             // pivot of step j at piv[j & 1]
    const int tid = threadIdx.x;
    int tr, tc;
    ci_tile_of<1>(tid, tr, tc);
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
    // column jn (in-tile index jjn) of this thread's tile, already updated by every step < jn: scale it and publish it
    auto publish = [&](int jn, int jjn_rt, double pivot) {
        double s, inv;
        pivot_root<FAST>(pivot, s, inv);
        double* v = work + (jn & 1) * 3 * NB;
        double* vcm = v + NB;
        double* vup = v + 2 * NB;
        if (tr == 0) piv[jn & 1] = pivot;
#pragma unroll
        for (int rr = 0; rr < 8; ++rr) {
            const int r = R0 + rr;
            double t = 0.0;
#pragma unroll
            for (int cc = 0; cc < 8; ++cc) {              // jjn_rt is a compile-time constant at every call site
                if (cc == jjn_rt) {
                    if (r == jn) { E[rr][cc] = s; t = inv; xd[r] = inv; }
                    else { t = E[rr][cc] * inv; E[rr][cc] = t; }
                }
            }
            const int ti = rr * 16 + tr;
            v[ti] = t;
            vcm[ti] = (r > jn) ? t : 0.0;
            vup[ti] = (r <= jn) ? t : 0.0;
        }
    };
    if (tc == 0) publish(0, 0, pv[0]);
    for (int jb = 0; jb * 8 < n; ++jb) {
#pragma unroll
        for (int jj = 0; jj < 8; ++jj) {
            const int j = jb * 8 + jj;
            if (j < n) {                               // uniform over the CTA
                const double* v = work + (j & 1) * 3 * NB;
                const double* vcm = v + NB;
                const double* vup = v + 2 * NB;
                __syncthreads();                       // column j is published
                if (!(piv[j & 1] > 0.0) && bad == 0) bad = j + 1;
                if (C0 + 7 > j) {                      // columns <= j are final
                    double vc[8];
#pragma unroll
                    for (int k = 0; k < 8; ++k) vc[k] = vcm[k * 16 + tc];
#pragma unroll
                    for (int k = 0; k < 8; ++k) pv[k] = fma(-vc[k], vc[k], pv[k]);
                    const bool live = !(tc > tr && R0 > j);      // tiles above the diagonal whose rows are all > j have not started
                    const double* row_lo = (tc > tr) ? vup : v;  // row multiplier of the elements with c <= r
                    const double* row_up = (tc < tr) ? v : vup;  // ... and of those with c > r
                    double lo[8], up[8];
#pragma unroll
                    for (int k = 0; k < 8; ++k) { lo[k] = row_lo[k * 16 + tr]; up[k] = row_up[k * 16 + tr]; }
                    const int jn = j + 1;
                    const int jjn = (jj + 1) & 7;                // compile-time after unrolling
                    const bool own_next = ((jn >> 3) == tc) && jn < n;
                    if (own_next) {                              // one warp (16 lanes): look ahead
#pragma unroll
                        for (int rr = 0; rr < 8; ++rr)
                            E[rr][jjn] = fma(-((jjn <= rr) ? lo[rr] : up[rr]), vc[jjn], E[rr][jjn]);
                        publish(jn, jjn, pv[jjn]);
                    }
                    if (live) {
                        // column jjn was updated ahead by its owners; every other column unconditionally (jjn is a
                        // compile-time constant here: no per-element test)
#pragma unroll
                        for (int rr = 0; rr < 8; ++rr)
#pragma unroll
                            for (int cc = 0; cc < 8; ++cc)
                                if (cc != jjn)
                                    E[rr][cc] = fma(-((cc <= rr) ? lo[rr] : up[rr]), vc[cc], E[rr][cc]);
                        if (!own_next) {
#pragma unroll
                            for (int rr = 0; rr < 8; ++rr)
                                E[rr][jjn] = fma(-((jjn <= rr) ? lo[rr] : up[rr]), vc[jjn], E[rr][jjn]);
                        }
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
template <int MAP = 0>
__device__ __forceinline__ void chol_inv_dump(const double (&E)[8][8], const double* __restrict__ xd,
                                              double* __restrict__ Lm) {
    const int tid = threadIdx.x;
    int tr, tc;
    ci_tile_of<MAP>(tid, tr, tc);
    const int R0 = tr * 8, C0 = tc * 8;
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

// ---------------------------------------------------------------------------------------------------------------------
// Blocked form of the same factorisation for a full 128 x 128 block (the diagonal-block kernel of the blocked path).
//
// chol_inv_regs pays one block barrier and one square root -> reciprocal -> scaling chain per column: 128 x ~0.8 us.  Here
// the columns are taken in panels of CB = 8: the panel (128 rows x 8 columns: L part below the diagonal, X part above) is
// handed to ONE warp, which runs the 8 column steps on registers with shuffles only (no block barrier inside the serial
// chain), publishes the 8 scaled columns, and all 256 threads then apply them to their 8 x 8 register tiles as 8 rank-1
// updates in a row.  2 block barriers per panel = 32 per block instead of 128.
//
// Every element still receives exactly the FMAs of chol_inv_regs, with the same operands, in the same order (column k
// ascending; a masked term is skipped instead of adding an exact zero), and the pivots' square roots and reciprocals are the
// same correctly rounded operations: the results are bit-identical to chol_inv_regs.
constexpr int CB = 8;                        // panel width
constexpr int CB_PITCH = CB + 1;             // row pitch of the panel staging (odd: the lanes of the panel warp own rows l + 32 q)
// doubles of shared scratch: two panel staging buffers, published columns (row operands, masked row operands, column
// operands in [k][c % 8][c / 8] order), the status word
constexpr int CBK_PN = NB * CB_PITCH;
constexpr int CBK_WORK = 2 * CBK_PN + 3 * NB * CB + 8;

// the panel warp: lane l owns rows l, l + 32, l + 64, l + 96; Q = (first panel column) / 32 = register index of the rows
// ps .. ps + 7 that the pivots and the column operands come from
template <int Q>
__device__ __forceinline__ int chol_panel_warp(double* __restrict__ pn, int ps, double* __restrict__ vr,
                                               double* __restrict__ vup, double* __restrict__ vc, double* __restrict__ xd) {
    const int l = threadIdx.x & 31;
    const int src0 = ps & 31;
    double P[4][CB], T[4];
    int bad = 0;
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int cc = 0; cc < CB; ++cc) P[q][cc] = pn[(l + 32 * q) * CB_PITCH + cc];
#pragma unroll
    for (int jj = 0; jj < CB; ++jj) {
        const int j = ps + jj;
        const double piv = __shfl_sync(0xffffffffu, P[Q][jj], src0 + jj);
        if (!(piv > 0.0) && bad == 0) bad = j + 1;
        const double s = __dsqrt_rn(piv);
        const double inv = __drcp_rn(s);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int r = l + 32 * q;
            if (r == j) { T[q] = inv; P[q][jj] = s; }
            else { T[q] = P[q][jj] * inv; P[q][jj] = T[q]; }
            vr[r * CB + jj] = T[q];
            vup[r * CB + jj] = (r <= j) ? T[q] : 0.0;
            vc[(jj * 8 + (r & 7)) * 16 + (r >> 3)] = T[q];
        }
        if (l == src0 + jj) xd[j] = inv;
#pragma unroll
        for (int cc = jj + 1; cc < CB; ++cc) {
            const int c = ps + cc;
            const double vcol = __shfl_sync(0xffffffffu, T[Q], src0 + cc);       // scaled column j at row c (c > j)
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int r = l + 32 * q;
                if (c <= r || r <= j) P[q][cc] = fma(-T[q], vcol, P[q][cc]);
            }
        }
    }
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int cc = 0; cc < CB; ++cc) pn[(l + 32 * q) * CB_PITCH + cc] = P[q][cc];
    return bad;
}

// All 256 threads must call.  work: CBK_WORK doubles, xd: NB doubles, both in shared memory.  Same contract as
// chol_inv_regs with n = NB.
__device__ __forceinline__ int chol_inv_blocked(double (&E)[8][8], double* work, double* xd) {
    double* vr = work + 2 * CBK_PN;
    double* vup = vr + NB * CB;
    double* vc = vup + NB * CB;
    int* status = reinterpret_cast<int*>(vc + NB * CB);
    const int tid = threadIdx.x;
    const int tr = tid >> 4, tc = tid & 15;
    const int R0 = tr * 8;
    if (tid == 0) status[0] = 0;
#pragma unroll 1
    for (int jb = 0; jb < NB / CB; ++jb) {
        const int ps = jb * CB;
        double* pn = work + (jb & 1) * CBK_PN;
        if (tc == jb) {                                   // this thread owns 8 rows of the panel's columns
#pragma unroll
            for (int rr = 0; rr < 8; ++rr)
#pragma unroll
                for (int cc = 0; cc < 8; ++cc) pn[(R0 + rr) * CB_PITCH + cc] = E[rr][cc];
        }
        __syncthreads();
        if (tid < 32) {
            int bad;
            switch (jb >> 2) {
                case 0: bad = chol_panel_warp<0>(pn, ps, vr, vup, vc, xd); break;
                case 1: bad = chol_panel_warp<1>(pn, ps, vr, vup, vc, xd); break;
                case 2: bad = chol_panel_warp<2>(pn, ps, vr, vup, vc, xd); break;
                default: bad = chol_panel_warp<3>(pn, ps, vr, vup, vc, xd); break;
            }
            if (bad && tid == 0 && status[0] == 0) status[0] = bad;
        }
        __syncthreads();
        if (tc == jb) {
#pragma unroll
            for (int rr = 0; rr < 8; ++rr)
#pragma unroll
                for (int cc = 0; cc < 8; ++cc) E[rr][cc] = pn[(R0 + rr) * CB_PITCH + cc];
        } else if (tc > jb && (tr >= tc || tr <= jb)) {
            // trailing tiles: the L part (on or below the diagonal) and the rows of the X part that have started
            const double* row_lo = (tc > tr) ? vup : vr;          // row multiplier of the elements with c <= r
            const double* row_up = (tc < tr) ? vr : vup;          // ... and of those with c > r
#pragma unroll
            for (int k = 0; k < CB; ++k) {
                double vcol[8], lo[8], up[8];
#pragma unroll
                for (int cc = 0; cc < 8; ++cc) vcol[cc] = vc[(k * 8 + cc) * 16 + tc];
#pragma unroll
                for (int rr = 0; rr < 8; ++rr) { lo[rr] = row_lo[(R0 + rr) * CB + k]; up[rr] = row_up[(R0 + rr) * CB + k]; }
#pragma unroll
                for (int rr = 0; rr < 8; ++rr)
#pragma unroll
                    for (int cc = 0; cc < 8; ++cc)
                        E[rr][cc] = fma(-((cc <= rr) ? lo[rr] : up[rr]), vcol[cc], E[rr][cc]);
            }
        }
    }
    __syncthreads();
    return status[0];
}

}  // namespace gpk
