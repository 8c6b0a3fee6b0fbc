// Diagonal-block step of the blocked batched Cholesky (kernel (b), panel part):
// factor the 128x128 block A_JJ (already updated by the left-looking SYRK) and invert the factor.
// One CTA per particle.  Reference step being replaced: numpy.linalg.cholesky at GP_func.py:22.
#include "gpk_cholinv.cuh"

#include <cstdlib>

namespace gpk {

constexpr int DIAG_DEFAULT_VARIANT = 3;
constexpr int DIAG_WORK = (CBK_WORK > CI_WORK) ? CBK_WORK : CI_WORK;
constexpr size_t DIAG_SMEM_BYTES = (size_t)NB * CI_LD * sizeof(double) + (size_t)(DIAG_WORK + NB) * sizeof(double);

// VARIANT 0: one column per step, owners of a column spread over the eight warps (round 1); 1: panels of 8 columns in one
// warp; 2: as 0 with the short reciprocal-square-root chain; 3 (default): look-ahead schedule with warp-owned columns
// (chol_inv_lookahead), correctly rounded pivot roots as the reference's LAPACK path; 4: as 3 with the short root chain.
// All five give the same bits on the test problems (0, 1 and 3 by construction: same operations in the same order per
// element).  Measured on C2 (N = 1024, 64 particles, profiles/r02_ab_diag_la_v*.json), real launches: 107 / 152 / 103 /
// 83 / 77 us.  What the ncu source counters showed for variant 0 (profiles/r02_ncu_diag_source.md): 242 warp instructions
// per warp and column, of which 64 are the update FMAs, 80 the publish block and 30 the pivot's root and reciprocal --
// executed by every warp with two active lanes, before the barrier.  In variant 3 one warp owns a column; it is then the
// critical path by itself (root ~400 cycles, its own 56 FMAs, the publish: ~900 cycles per column).
template <int VARIANT>
__global__ void __launch_bounds__(256, 1)
k_diag_block(double* __restrict__ mat, int Np, int J, double* __restrict__ dinv, double* __restrict__ dinvT,
             int T, int* __restrict__ info, const unsigned char* __restrict__ active, int mode) {
    extern __shared__ double smem[];
    double* Lm = smem;
    double* work = smem + NB * CI_LD;
    double* xr = work + DIAG_WORK;
    const int p = blockIdx.x;
    if (active && !active[p]) return;
    const int tid = threadIdx.x;
    double* blk = mat + (size_t)p * Np * Np + (size_t)J * NB * Np + (size_t)J * NB;

    // each thread loads its 8x8 register tile of the lower triangle (the X part starts at zero)
    double* xd = xr;
    double E[8][8];
    constexpr int MAP = (VARIANT >= 3) ? 1 : 0;
    {
        int tr, tc;
        ci_tile_of<MAP>(tid, tr, tc);
        const int R0 = tr * 8, C0 = tc * 8;
#pragma unroll
        for (int rr = 0; rr < 8; ++rr)
#pragma unroll
            for (int cc = 0; cc < 8; ++cc) {
                const int r = R0 + rr, c = C0 + cc;
                E[rr][cc] = (c <= r) ? blk[(size_t)r * Np + c] : 0.0;
            }
    }
    int bad;
    if (VARIANT == 1) bad = chol_inv_blocked(E, work, xd);
    else if (VARIANT == 2) bad = chol_inv_regs<true>(E, NB, work, xd);
    else if (VARIANT == 3) bad = chol_inv_lookahead<false>(E, NB, work, xd);
    else if (VARIANT == 4) bad = chol_inv_lookahead<true>(E, NB, work, xd);
    else bad = chol_inv_regs<false>(E, NB, work, xd);
    if (bad && tid == 0) atomicCAS(&info[p], 0, J * NB + bad);
    chol_inv_dump<MAP>(E, xd, Lm);
    __syncthreads();

    double* dI = dinv + ((size_t)p * T + J) * NB * NB;
    double* dT = dinvT + ((size_t)p * T + J) * NB * NB;
    for (int idx = tid; idx < NB * NB; idx += 256) {
        const int r = idx >> 7, c = idx & (NB - 1);
        const double x_rc = (c <= r) ? Lm[c * CI_LD + r + 1] : 0.0;   // D[r][c]
        const double x_cr = (r <= c) ? Lm[r * CI_LD + c + 1] : 0.0;   // D[c][r]
        dI[idx] = x_rc;
        dT[idx] = x_cr;
        if (mode == DIAG_WRITE_INVERSE) blk[(size_t)r * Np + c] = x_rc;
        else blk[(size_t)r * Np + c] = (c <= r) ? Lm[r * CI_LD + c] : 0.0;
    }
}

int launch_diag_block(double* mat, int Np, int J, int P, double* dinv, double* dinvT, int T, int* info,
                      const unsigned char* active, int mode, cudaStream_t st) {
    static int variant = -1;          // GPK_DIAG_VARIANT = 0 .. 4 (A/B runs and tests)
    if (variant < 0) {
        const char* v = getenv("GPK_DIAG_VARIANT");
        const int want = (v && atoi(v) >= 0 && atoi(v) <= 4) ? atoi(v) : DIAG_DEFAULT_VARIANT;
        cudaError_t e = cudaSuccess;
#define GPK_DIAG_ATTR(V) case V: e = cudaFuncSetAttribute(k_diag_block<V>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DIAG_SMEM_BYTES); break
        switch (want) { GPK_DIAG_ATTR(0); GPK_DIAG_ATTR(1); GPK_DIAG_ATTR(2); GPK_DIAG_ATTR(3); default: GPK_DIAG_ATTR(4); }
#undef GPK_DIAG_ATTR
        if (e != cudaSuccess) { set_error("k_diag_block smem attr: %s", cudaGetErrorString(e)); return 1; }
        variant = want;
    }
#define GPK_DIAG_GO(V) case V: k_diag_block<V><<<P, 256, DIAG_SMEM_BYTES, st>>>(mat, Np, J, dinv, dinvT, T, info, active, mode); break
    switch (variant) { GPK_DIAG_GO(0); GPK_DIAG_GO(1); GPK_DIAG_GO(2); GPK_DIAG_GO(3); default: GPK_DIAG_GO(4); }
#undef GPK_DIAG_GO
    return check_launch("k_diag_block");
}

// Staged interface only (gpk_fit_alpha_batched / gpk_selfpredict_counts): the diagonal blocks of an already computed
// factor L (gpk_potrf_batched output) are inverted by forward substitution on the identity, one column per thread, and
// written where the fused path's diagonal kernel leaves them: dst diagonal block, dinv, dinvT.
constexpr int TI_TRI = NB * (NB + 1) / 2;                       // packed lower triangle: (r, c) at r (r + 1) / 2 + c
constexpr size_t TI_SMEM_BYTES = (size_t)2 * TI_TRI * sizeof(double);

__global__ void __launch_bounds__(NB, 1)
k_tri_inv_diag(const double* __restrict__ Lsrc, double* __restrict__ dst, int Np, int T, double* __restrict__ dinv,
               double* __restrict__ dinvT, int* __restrict__ info, const unsigned char* __restrict__ active) {
    extern __shared__ double smem[];
    double* Ls = smem;                        // L_JJ, packed
    double* Xs = smem + TI_TRI;               // X = L_JJ^-1, packed
    const int J = blockIdx.x, p = blockIdx.y;
    if (active && !active[p]) return;
    const int t = threadIdx.x;
    const double* blk = Lsrc + (size_t)p * Np * Np + (size_t)J * NB * Np + (size_t)J * NB;
    for (int idx = t; idx < NB * NB; idx += NB) {
        const int r = idx >> 7, c = idx & (NB - 1);
        if (c <= r) Ls[r * (r + 1) / 2 + c] = blk[(size_t)r * Np + c];
    }
    __syncthreads();
    if (!(Ls[t * (t + 1) / 2 + t] > 0.0)) atomicCAS(&info[p], 0, J * NB + t + 1);
    // column t of X: x_i = (delta_it - sum_{k=t}^{i-1} L_ik x_k) / L_ii; a thread reads only its own column of Xs
    for (int i = t; i < NB; ++i) {
        const double* Li = Ls + i * (i + 1) / 2;
        double acc = (i == t) ? 1.0 : 0.0;
        for (int k = t; k < i; ++k) acc = fma(-Li[k], Xs[k * (k + 1) / 2 + t], acc);
        Xs[i * (i + 1) / 2 + t] = acc / Li[i];
    }
    __syncthreads();
    double* dI = dinv + ((size_t)p * T + J) * NB * NB;
    double* dT = dinvT + ((size_t)p * T + J) * NB * NB;
    double* out = dst + (size_t)p * Np * Np + (size_t)J * NB * Np + (size_t)J * NB;
    for (int idx = t; idx < NB * NB; idx += NB) {
        const int r = idx >> 7, c = idx & (NB - 1);
        const double x_rc = (c <= r) ? Xs[r * (r + 1) / 2 + c] : 0.0;
        dI[idx] = x_rc;
        dT[idx] = (r <= c) ? Xs[c * (c + 1) / 2 + r] : 0.0;
        out[(size_t)r * Np + c] = x_rc;
    }
}

int launch_tri_inv_diag(const double* Lsrc, double* dst, int Np, int T, int P, double* dinv, double* dinvT, int* info,
                        const unsigned char* active, cudaStream_t st) {
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(k_tri_inv_diag, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TI_SMEM_BYTES);
        if (e != cudaSuccess) { set_error("k_tri_inv_diag smem attr: %s", cudaGetErrorString(e)); return 1; }
        configured = true;
    }
    k_tri_inv_diag<<<dim3(T, P), NB, TI_SMEM_BYTES, st>>>(Lsrc, dst, Np, T, dinv, dinvT, info, active);
    return check_launch("k_tri_inv_diag");
}

}  // namespace gpk
