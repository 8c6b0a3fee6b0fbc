// FP64 tensor-core (DMMA) tile engine for the blocked batched factorisation, inverse and prediction.
//
// Every dense contraction on the path has the "TN" shape  C[m][n] (+)= sum_k A[m][k] * B[n][k]  with
// both operands K-contiguous in HBM, so one main loop serves them all:
//   SYRK   (Cholesky, left-looking)  A_IJ -= sum_{K<J} L_IK L_JK^T                  GP_func.py:22
//   TRSM   (as GEMM with D_J = L_JJ^-1)  L_IJ = S_IJ D_J^T
//   INVP   Mt = D_J^T-panel product (L_KJ D_J)^T                                      phase 2 (L^-1)
//   INVU   W_IJ = - sum_{K=J+1..I} W_IK M_K                                           GP_func.py:38 (all N rhs)
//   PRED   || W k_* ||^2 column sums for a tile of grid points                        GP_func.py:38-39
// CTA tile 128x128, 8 warps (2x4), warp tile 64x32 = 8x4 mma.m8n8k4.f64 fragments (sm_100a has no
// tcgen05 f64 kind; DMMA.8x8x4 is the FP64 tensor instruction, 37.1 TFLOP/s measured peak).
// Operands are staged through a 4-deep cp.async ring of 16-wide K chunks; rows are padded to 20
// doubles so the fragment loads (row = lane/4, k = lane%4) are bank-conflict free.
#include "gpk_common.cuh"

namespace gpk {

constexpr int KC = 16;                 // K chunk per ring stage (one block barrier per chunk); 32 x 3 stages measured 1.5 % slower
constexpr int SLD = KC + 4;            // padded row stride: 2*SLD = 8 (mod 32) -> conflict-free fragment loads
constexpr int STAGES = 4;
constexpr int KSTEPS = KC / 4;         // m8n8k4 steps per chunk (even)
constexpr int ROWCH = KC / 2;          // 16-byte chunks per operand row
constexpr int GEMM_THREADS = 512;      // 16 warps as 4 (M) x 4 (N), warp tile 32 x 32 = 4 x 4 DMMA fragments
constexpr int MI = 4;                  // 8-row fragments per warp tile
constexpr int WARP_M = MI * 8;
constexpr int LD_ITERS = NB * ROWCH / GEMM_THREADS;
constexpr int OPER_ELEMS = NB * SLD;                                  // one operand, one stage
constexpr size_t GEMM_SMEM_BYTES = (size_t)STAGES * 2 * OPER_ELEMS * sizeof(double);

__device__ __forceinline__ void cp_async16(double* dst_smem, const double* src) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(dst_smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(src));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N)); }

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
    asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
        : "+d"(c0), "+d"(c1)
        : "d"(a), "d"(b));
}

struct Tile {
    const double* A;   // &A[row0][k0]
    const double* B;   // &B[row0][k0]
    int lda, ldb;
    int kchunks;       // number of KC-wide K chunks
};

__device__ __forceinline__ void load_stage(const Tile& t, int kc, double* As, double* Bs, int tid) {
#pragma unroll
    for (int i = 0; i < LD_ITERS; ++i) {
        const int cid = tid + GEMM_THREADS * i;
        const int row = cid / ROWCH, ch = cid % ROWCH;
        cp_async16(As + row * SLD + ch * 2, t.A + (size_t)row * t.lda + kc * KC + ch * 2);
    }
#pragma unroll
    for (int i = 0; i < LD_ITERS; ++i) {
        const int cid = tid + GEMM_THREADS * i;
        const int row = cid / ROWCH, ch = cid % ROWCH;
        cp_async16(Bs + row * SLD + ch * 2, t.B + (size_t)row * t.ldb + kc * KC + ch * 2);
    }
}

// acc[mi][ni][h]: row = warp_m*64 + mi*8 + lane/4, col = warp_n*32 + ni*8 + (lane%4)*2 + h
//
// Software pipeline: the fragments of k4-step t+1 are loaded into a second register set while the 32 DMMAs of
// step t issue, also across the chunk boundary -- the ring barrier sits *before* the last k4-step of a chunk,
// whose DMMAs then cover the barrier, the cp.async refill and the first fragment loads of the next chunk
// (both warps of a scheduler run in lockstep, so nothing else would hide those bubbles; ncu r01: 85 % -> v2).
__device__ __forceinline__ void load_frags(const double* As, const double* Bs, int kk, double (&a)[MI], double (&b)[4]) {
#pragma unroll
    for (int mi = 0; mi < MI; ++mi) a[mi] = As[mi * 8 * SLD + kk * 4];
#pragma unroll
    for (int ni = 0; ni < 4; ++ni) b[ni] = Bs[ni * 8 * SLD + kk * 4];
}

__device__ __forceinline__ void mma_step(double (&acc)[MI][4][2], const double (&a)[MI], const double (&b)[4]) {
#pragma unroll
    for (int mi = 0; mi < MI; ++mi)
#pragma unroll
        for (int ni = 0; ni < 4; ++ni) dmma884(acc[mi][ni][0], acc[mi][ni][1], a[mi], b[ni]);
}

__device__ __forceinline__ void mainloop(const Tile& t, double (&acc)[MI][4][2], double* smem) {
    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int warp_m = warp >> 2, warp_n = warp & 3;
    const int g = lane >> 2, q = lane & 3;
#pragma unroll
    for (int mi = 0; mi < MI; ++mi)
#pragma unroll
        for (int ni = 0; ni < 4; ++ni) acc[mi][ni][0] = acc[mi][ni][1] = 0.0;

    // prologue: fill the whole ring
#pragma unroll
    for (int s = 0; s < STAGES; ++s) {
        if (s < t.kchunks) load_stage(t, s, smem + (size_t)s * 2 * OPER_ELEMS, smem + (size_t)s * 2 * OPER_ELEMS + OPER_ELEMS, tid);
        cp_async_commit();
    }
    const int a_off = (warp_m * WARP_M + g) * SLD + q;
    const int b_off = (warp_n * 32 + g) * SLD + q;
    double fa[2][MI], fb[2][4];
    cp_async_wait<STAGES - 1>();          // chunk 0 has landed
    __syncthreads();
    load_frags(smem + a_off, smem + OPER_ELEMS + b_off, 0, fa[0], fb[0]);

    for (int kc = 0; kc < t.kchunks; ++kc) {
        const double* As = smem + (size_t)(kc % STAGES) * 2 * OPER_ELEMS + a_off;
        const double* Bs = As - a_off + OPER_ELEMS + b_off;
#pragma unroll
        for (int kk = 0; kk < KSTEPS; ++kk) {
            const int cur = kk & 1, nxt = cur ^ 1;
            if (kk < KSTEPS - 1) {
                load_frags(As, Bs, kk + 1, fa[nxt], fb[nxt]);
            } else if (kc + 1 < t.kchunks) {
                cp_async_wait<STAGES - 2>();  // chunk kc+1 has landed (this thread's copies) ...
                __syncthreads();              // ... everybody's; and nobody reads the slot of chunk kc any more
                const double* An = smem + (size_t)((kc + 1) % STAGES) * 2 * OPER_ELEMS + a_off;
                load_frags(An, An - a_off + OPER_ELEMS + b_off, 0, fa[nxt], fb[nxt]);
                if (kc + STAGES < t.kchunks) {
                    double* base = smem + (size_t)(kc % STAGES) * 2 * OPER_ELEMS;
                    load_stage(t, kc + STAGES, base, base + OPER_ELEMS, tid);
                }
                cp_async_commit();
            }
            mma_step(acc, fa[cur], fb[cur]);
        }
    }
    cp_async_wait<0>();
    __syncthreads();
}

enum Epi { EPI_SUB = 0, EPI_STORE = 1, EPI_STORE_NEG = 2 };

template <int EPI>
__device__ __forceinline__ void store_tile(double* C, int ldc, const double (&acc)[MI][4][2]) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int warp_m = warp >> 2, warp_n = warp & 3;
    const int g = lane >> 2, q = lane & 3;
#pragma unroll
    for (int mi = 0; mi < MI; ++mi) {
        const int row = warp_m * WARP_M + mi * 8 + g;
#pragma unroll
        for (int ni = 0; ni < 4; ++ni) {
            const int colx = warp_n * 32 + ni * 8 + q * 2;
            double2* ptr = reinterpret_cast<double2*>(C + (size_t)row * ldc + colx);
            double2 v;
            if (EPI == EPI_SUB) {
                v = *ptr;
                v.x -= acc[mi][ni][0];
                v.y -= acc[mi][ni][1];
            } else if (EPI == EPI_STORE) {
                v.x = acc[mi][ni][0];
                v.y = acc[mi][ni][1];
            } else {
                v.x = -acc[mi][ni][0];
                v.y = -acc[mi][ni][1];
            }
            *ptr = v;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Left-looking Cholesky update of block column J:  A_IJ -= sum_{K<J} L_IK L_JK^T,  I = J + blockIdx.x
__global__ void __launch_bounds__(GEMM_THREADS, 1)
k_syrk_update(double* __restrict__ mat, int Np, int J, const unsigned char* __restrict__ active) {
    extern __shared__ double smem[];
    const int p = blockIdx.y;
    if (active && !active[p]) return;
    const int I = J + blockIdx.x;
    double* M = mat + (size_t)p * Np * Np;
    Tile t{M + (size_t)I * NB * Np, M + (size_t)J * NB * Np, Np, Np, J * (NB / KC)};
    double acc[MI][4][2];
    mainloop(t, acc, smem);
    store_tile<EPI_SUB>(M + (size_t)I * NB * Np + (size_t)J * NB, Np, acc);
}

// L_IJ = S_IJ * D_J^T  (in place),  I = J + 1 + blockIdx.x
__global__ void __launch_bounds__(GEMM_THREADS, 1)
k_trsm_panel(double* __restrict__ mat, int Np, int T, int J, const double* __restrict__ dinv,
             const unsigned char* __restrict__ active) {
    extern __shared__ double smem[];
    const int p = blockIdx.y;
    if (active && !active[p]) return;
    const int I = J + 1 + blockIdx.x;
    double* C = mat + (size_t)p * Np * Np + (size_t)I * NB * Np + (size_t)J * NB;
    Tile t{C, dinv + ((size_t)p * T + J) * NB * NB, Np, NB, NB / KC};
    double acc[MI][4][2];
    mainloop(t, acc, smem);     // ends with a block barrier: every read of S_IJ has landed
    store_tile<EPI_STORE>(C, Np, acc);
}

// Phase 2 panel: Mt[n][r] = sum_j D_J^T[n][j] * L[r][J*NB + j],  r in block J + 1 + blockIdx.x
__global__ void __launch_bounds__(GEMM_THREADS, 1)
k_inv_panel(const double* __restrict__ mat, int Np, int T, int J, const double* __restrict__ dinvT,
            double* __restrict__ mt, const unsigned char* __restrict__ active) {
    extern __shared__ double smem[];
    const int p = blockIdx.y;
    if (active && !active[p]) return;
    const int R = J + 1 + blockIdx.x;
    const double* M = mat + (size_t)p * Np * Np;
    Tile t{dinvT + ((size_t)p * T + J) * NB * NB, M + (size_t)R * NB * Np + (size_t)J * NB, NB, Np, NB / KC};
    double acc[MI][4][2];
    mainloop(t, acc, smem);
    store_tile<EPI_STORE>(mt + (size_t)p * NB * Np + (size_t)R * NB, Np, acc);
}

// Phase 2 update: W_IJ = - sum_{k in [(J+1)NB, (I+1)NB)} W[I-block][k] * Mt[:, k]^T,  deepest I first
__global__ void __launch_bounds__(GEMM_THREADS, 1)
k_inv_update(double* __restrict__ mat, int Np, int T, int J, const double* __restrict__ mt,
             const unsigned char* __restrict__ active) {
    extern __shared__ double smem[];
    const int p = blockIdx.y;
    if (active && !active[p]) return;
    const int I = T - 1 - blockIdx.x;
    double* M = mat + (size_t)p * Np * Np;
    Tile t{M + (size_t)I * NB * Np + (size_t)(J + 1) * NB, mt + (size_t)p * NB * Np + (size_t)(J + 1) * NB, Np,
           Np, (I - J) * (NB / KC)};
    double acc[MI][4][2];
    mainloop(t, acc, smem);
    store_tile<EPI_STORE_NEG>(M + (size_t)I * NB * Np + (size_t)J * NB, Np, acc);
}

// Prediction: partial[I][b*NB + n] = sum_{m in block I} ( sum_k W[m][k] * KsT[b*NB + n][k] )^2
__global__ void __launch_bounds__(GEMM_THREADS, 1)
k_predict_gemm(const double* __restrict__ W, int Np, int T, const double* __restrict__ KsT, int Bp,
               double* __restrict__ partial) {
    extern __shared__ double smem[];
    const int I = T - 1 - blockIdx.x;
    const int b = blockIdx.y;
    Tile t{W + (size_t)I * NB * Np, KsT + (size_t)b * NB * Np, Np, Np, (I + 1) * (NB / KC)};
    double acc[MI][4][2];
    mainloop(t, acc, smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int warp_m = warp >> 2, warp_n = warp & 3;
    const int q = lane & 3;
    double* red = smem;   // [NB / WARP_M][NB]
#pragma unroll
    for (int ni = 0; ni < 4; ++ni) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            double s = 0.0;
#pragma unroll
            for (int mi = 0; mi < MI; ++mi) s = fma(acc[mi][ni][h], acc[mi][ni][h], s);
            s += __shfl_xor_sync(0xffffffffu, s, 4);
            s += __shfl_xor_sync(0xffffffffu, s, 8);
            s += __shfl_xor_sync(0xffffffffu, s, 16);
            if (lane < 4) red[warp_m * NB + warp_n * 32 + ni * 8 + q * 2 + h] = s;
        }
    }
    __syncthreads();
    if (threadIdx.x < NB) {
        double tot = 0.0;
#pragma unroll
        for (int w = 0; w < NB / WARP_M; ++w) tot += red[w * NB + threadIdx.x];
        partial[(size_t)I * Bp + (size_t)b * NB + threadIdx.x] = tot;
    }
}

// ---------------------------------------------------------------------------------------------
template <typename K>
static int ensure_smem(K kernel, bool& done) {
    if (!done) {
        cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GEMM_SMEM_BYTES);
        if (e != cudaSuccess) { set_error("gemm smem attr: %s", cudaGetErrorString(e)); return 1; }
        done = true;
    }
    return 0;
}

int launch_syrk_update(double* mat, int Np, int T, int J, int P, const unsigned char* active, cudaStream_t st) {
    static bool done = false;
    if (J == 0) return 0;
    if (ensure_smem(k_syrk_update, done)) return 1;
    k_syrk_update<<<dim3(T - J, P), GEMM_THREADS, GEMM_SMEM_BYTES, st>>>(mat, Np, J, active);
    return check_launch("k_syrk_update");
}

int launch_trsm_panel(double* mat, int Np, int T, int J, int P, const double* dinv, const unsigned char* active,
                      cudaStream_t st) {
    static bool done = false;
    if (J >= T - 1) return 0;
    if (ensure_smem(k_trsm_panel, done)) return 1;
    k_trsm_panel<<<dim3(T - 1 - J, P), GEMM_THREADS, GEMM_SMEM_BYTES, st>>>(mat, Np, T, J, dinv, active);
    return check_launch("k_trsm_panel");
}

int launch_inv_panel(const double* mat, int Np, int T, int J, int P, const double* dinvT, double* mt,
                     const unsigned char* active, cudaStream_t st) {
    static bool done = false;
    if (J >= T - 1) return 0;
    if (ensure_smem(k_inv_panel, done)) return 1;
    k_inv_panel<<<dim3(T - 1 - J, P), GEMM_THREADS, GEMM_SMEM_BYTES, st>>>(mat, Np, T, J, dinvT, mt, active);
    return check_launch("k_inv_panel");
}

int launch_inv_update(double* mat, int Np, int T, int J, int P, const double* mt, const unsigned char* active,
                      cudaStream_t st) {
    static bool done = false;
    if (J >= T - 1) return 0;
    if (ensure_smem(k_inv_update, done)) return 1;
    k_inv_update<<<dim3(T - 1 - J, P), GEMM_THREADS, GEMM_SMEM_BYTES, st>>>(mat, Np, T, J, mt, active);
    return check_launch("k_inv_update");
}

int launch_predict_gemm(const double* W, int Np, int T, const double* KsT, int Bp, double* partial,
                        cudaStream_t st) {
    static bool done = false;
    if (ensure_smem(k_predict_gemm, done)) return 1;
    k_predict_gemm<<<dim3(T, Bp / NB), GEMM_THREADS, GEMM_SMEM_BYTES, st>>>(W, Np, T, KsT, Bp, partial);
    return check_launch("k_predict_gemm");
}

}  // namespace gpk
