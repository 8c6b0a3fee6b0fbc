// Diagonal-block step of the blocked batched Cholesky (kernel (b), panel part):
// factor the 128x128 block A_JJ (already updated by the left-looking SYRK) and invert the factor.
// One CTA per particle.  Reference step being replaced: numpy.linalg.cholesky at GP_func.py:22.
#include "gpk_cholinv.cuh"

namespace gpk {

__global__ void __launch_bounds__(256, 1)
k_diag_block(double* __restrict__ mat, int Np, int J, double* __restrict__ dinv, double* __restrict__ dinvT,
             int T, int* __restrict__ info, const unsigned char* __restrict__ active, int mode) {
    extern __shared__ double smem[];
    double* Lm = smem;
    double* work = smem + NB * CI_LD;
    double* xr = work + CI_WORK;
    const int p = blockIdx.x;
    if (active && !active[p]) return;
    const int tid = threadIdx.x;
    double* blk = mat + (size_t)p * Np * Np + (size_t)J * NB * Np + (size_t)J * NB;

    // each thread loads its 8x8 register tile of the lower triangle (the X part starts at zero)
    double* xd = xr;
    double E[8][8];
    {
        const int R0 = (tid >> 4) * 8, C0 = (tid & 15) * 8;
#pragma unroll
        for (int rr = 0; rr < 8; ++rr)
#pragma unroll
            for (int cc = 0; cc < 8; ++cc) {
                const int r = R0 + rr, c = C0 + cc;
                E[rr][cc] = (c <= r) ? blk[(size_t)r * Np + c] : 0.0;
            }
    }
    const int bad = chol_inv_regs(E, NB, work, xd);
    if (bad && tid == 0) atomicCAS(&info[p], 0, J * NB + bad);
    chol_inv_dump(E, xd, Lm);
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
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(k_diag_block, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)CI_SMEM_BYTES);
        if (e != cudaSuccess) { set_error("k_diag_block smem attr: %s", cudaGetErrorString(e)); return 1; }
        configured = true;
    }
    k_diag_block<<<P, 256, CI_SMEM_BYTES, st>>>(mat, Np, J, dinv, dinvT, T, info, active, mode);
    return check_launch("k_diag_block");
}

}  // namespace gpk
