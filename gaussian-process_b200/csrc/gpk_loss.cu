// Kernel (b), loss part: self-prediction at the training points and the count reduction of the
// calibration loss (find_len_scales.py:159-163), plus the one-CTA small-N path that fuses everything.
//
// With D = diag(e^2), K = k(X,X) + D and K_s = K - D (GP_func.py:21,32 with x_fit = x_known):
//     mu - y        = K_s alpha - y            = -D alpha                (alpha = K^-1 y)
//     var_i         = 1 - k_i^T K^-1 k_i       = d_i - d_i^2 (K^-1)_ii
//     (K^-1)_ii     = sum_k W_ki^2,  W = L^-1  (column sums of squares of the triangular inverse)
// so the N x N right-hand-side solve of GP_func.py:38 reduces to the triangular inverse (N^3/3 flop).
//
// Counts: |pull| <= 1 with pull = r / max(sd * s, 1e-12) (find_len_scales.py:161-163) is, for IEEE
// division, exactly |r| <= max(fl(sd * s), 1e-12); the predicate is monotone in s, so each point
// contributes to all scales from its first passing index on: histogram of first indices + prefix sum.
#include "gpk_cholinv.cuh"

#include <cstdlib>

namespace gpk {

// first index k in [0, S) with  a <= max(fl(sd * s_k), 1e-12);  S if none
__device__ __forceinline__ int first_inside(double a, double sd, const double* __restrict__ sv, int S) {
    int lo = 0, hi = S;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        const double thr = fmax(__dmul_rn(sd, sv[mid]), 1e-12);
        if (a <= thr) hi = mid; else lo = mid + 1;
    }
    return lo;
}

// hist[0..S] (int, shared) -> counts_out[0..S) inclusive prefix; called by all threads after a barrier
__device__ __forceinline__ void prefix_and_store(int* hist, int S, int* __restrict__ counts_out) {
    const int tid = threadIdx.x;
    if (tid < 32) {
        const int chunk = (S + 31) / 32;
        const int b = tid * chunk, e = min(S, b + chunk);
        int s = 0;
        for (int k = b; k < e; ++k) s += hist[k];
        int incl = s;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (tid >= o) incl += v;
        }
        int run = incl - s;
        for (int k = b; k < e; ++k) { run += hist[k]; hist[k] = run; }
    }
    __syncthreads();
    for (int k = tid; k < S; k += blockDim.x) counts_out[k] = hist[k];
}

__device__ __forceinline__ void resid_sigma(double d, double alpha, double cn, double& resid, double& sd) {
    resid = -(d * alpha);
    const double var = fma(-(d * cn), d, d);
    sd = sqrt(fmax(var, 1e-12));                  // np.clip(var, 1e-12, None); sqrt   GP_func.py:39-40
}

// ---------------------------------------------------------------------------------------------
// Small-N fused loss: covariance, Cholesky, inverse, alpha, counts -- one CTA per particle, all in
// shared memory (the reference's real PSO size: N <= 100 after the KMeans cap, find_len_scales.py:24-47).
constexpr int SMALL_DEFAULT_VARIANT = 1;
struct SmallSmem {
    static constexpr size_t lm = 0;
    static constexpr size_t col = lm + (size_t)NB * CI_LD * 8;
    static constexpr size_t xr = col + (size_t)CI_WORK * 8;
    static constexpr size_t xs = xr + NB * 8;
    static constexpr size_t sq = xs + (size_t)MAX_D * NB * 8;
    static constexpr size_t y = sq + NB * 8;
    static constexpr size_t d = y + NB * 8;
    static constexpr size_t z = d + NB * 8;
    static constexpr size_t a = z + NB * 8;
    static constexpr size_t sg = a + NB * 8;
    static constexpr size_t piv = sg + NB * 8;
    static constexpr size_t hist = piv + 16;
    static constexpr size_t total = hist + (size_t)(MAX_S + 1) * 4;
};

// FORM 0: reduced form (above).  FORM 1: the reference's own formulae (GP_func.py:36-40, find_len_scales.py:161-163):
// mu = K_s^T alpha, v = L^-1 K_s, var = 1 - sum_rows v^2, pull = (mu - y) / max(sd s, 1e-12) -- the cancellations of the
// reference are reproduced (K_s is rebuilt over the factor in shared memory, L^-1 is applied as the explicit inverse).
// Lower triangle of kernel_func(x, x) (+ diag(e^2) when dv != nullptr) into the shared staging array, Lm[r][c], c <= r < N
// (GP_func.py:49-65, :21), shared evenly by the 256 threads of the CTA: rows i and N - 1 - i are folded into one row of
// N + 1 elements, a thread takes every 256th element of the folded array, four independent chains at a time.  (Each thread
// evaluating the 64 elements of the 8 x 8 tile it owns instead runs 64 exp() latency chains back to back in the threads of the
// diagonal tiles while most of the CTA idles: 64 k cycles of a 227 k cycle kernel at N = 100, profiles/r02_small_profile.log.)
// Same expressions per element as before: bit-identical values.
__device__ __forceinline__ void small_cov_to_smem(const double (*xs)[NB], const double* __restrict__ sq,
                                                  const double* __restrict__ dv, int N, int d, double* __restrict__ Lm) {
    const int W = N + 1, total = ((N + 1) >> 1) * W;
    for (int e0 = threadIdx.x; e0 < total; e0 += 4 * 256) {
        int r[4], c[4];
        bool ok[4];
        double dot[4], val[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int e = e0 + 256 * u;
            ok[u] = e < total;
            const int i = ok[u] ? e / W : 0;
            const int j = ok[u] ? e - i * W : 0;
            if (j <= i) { r[u] = i; c[u] = j; }
            else { r[u] = N - 1 - i; c[u] = j - i - 1; ok[u] = ok[u] && (r[u] != i); }      // odd N: the middle row once
            dot[u] = __dmul_rn(xs[0][r[u]], xs[0][c[u]]);
        }
        for (int dd = 1; dd < d; ++dd) {
#pragma unroll
            for (int u = 0; u < 4; ++u) dot[u] = fma(xs[dd][r[u]], xs[dd][c[u]], dot[u]);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            double sd = __dsub_rn(__dadd_rn(sq[r[u]], sq[c[u]]), __dmul_rn(2.0, dot[u]));
            sd = fmax(sd, 0.0);
            val[u] = exp(__dmul_rn(-0.5, sd));
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (ok[u]) {
                if (dv != nullptr && r[u] == c[u]) val[u] = __dadd_rn(val[u], dv[r[u]]);
                Lm[r[u] * CI_LD + c[u]] = val[u];
            }
        }
    }
}

#ifdef GPK_SMALL_PROFILE
// instrumented build only (tools/small_profile.py): clock64 at the phase boundaries of CTA 0
__device__ long long g_small_prof[16];
#define SMALL_PROF(i) do { __syncthreads(); if (blockIdx.x == 0 && threadIdx.x == 0) g_small_prof[i] = clock64(); } while (0)
#else
#define SMALL_PROF(i)
#endif

// LA: the factorisation's schedule -- 0 chol_inv_regs (one column per step, owners spread over the warps), 1 chol_inv_lookahead
// (warp-owned columns, next column published ahead of the bulk update; bit-identical results)
template <int FORM, int LA>
__global__ void __launch_bounds__(256, 1)
k_small_loss(const double* __restrict__ X, const double* __restrict__ y, const double* __restrict__ e,
             const double* __restrict__ ls, const unsigned char* __restrict__ active,
             const double* __restrict__ sigma_vals, int N, int d, int S, int* __restrict__ counts,
             int* __restrict__ info, double* __restrict__ resid_out, double* __restrict__ sig_out) {
    extern __shared__ __align__(16) unsigned char raw[];
    double* Lm = reinterpret_cast<double*>(raw + SmallSmem::lm);
    double* col = reinterpret_cast<double*>(raw + SmallSmem::col);
    double* xr = reinterpret_cast<double*>(raw + SmallSmem::xr);
    double (*xs)[NB] = reinterpret_cast<double (*)[NB]>(raw + SmallSmem::xs);
    double* sq = reinterpret_cast<double*>(raw + SmallSmem::sq);
    double* yv = reinterpret_cast<double*>(raw + SmallSmem::y);
    double* dv = reinterpret_cast<double*>(raw + SmallSmem::d);
    double* zv = reinterpret_cast<double*>(raw + SmallSmem::z);
    double* av = reinterpret_cast<double*>(raw + SmallSmem::a);
    double* sg = reinterpret_cast<double*>(raw + SmallSmem::sg);
    int* hist = reinterpret_cast<int*>(raw + SmallSmem::hist);

    const int p = blockIdx.x;
    const int tid = threadIdx.x;
    if (active && !active[p]) {
        for (int k = tid; k < S; k += 256) counts[(size_t)p * S + k] = 0;
        return;
    }
    SMALL_PROF(0);
    if (tid < N) {
        double acc = 0.0;
        for (int dd = 0; dd < d; ++dd) {
            const double v = __ddiv_rn(X[(size_t)dd * N + tid], ls[(size_t)p * d + dd]);
            xs[dd][tid] = v;
            const double v2 = __dmul_rn(v, v);
            acc = (dd == 0) ? v2 : __dadd_rn(acc, v2);
        }
        sq[tid] = acc;
        yv[tid] = y[tid];
        const double ee = e[tid];
        dv[tid] = __dmul_rn(ee, ee);
    }
    for (int k = tid; k <= S; k += 256) hist[k] = 0;
    __syncthreads();

    SMALL_PROF(1);
    // K + diag(e^2): built by all threads into the staging array, then each thread loads its 8x8 register tile of the lower
    // triangle (zero elsewhere)
    small_cov_to_smem(xs, sq, dv, N, d, Lm);
    __syncthreads();
    double E[8][8];
    {
        int tr, tc;
        ci_tile_of<LA>(tid, tr, tc);
        const int R0 = tr * 8, C0 = tc * 8;
#pragma unroll
        for (int rr = 0; rr < 8; ++rr)
#pragma unroll
            for (int cc = 0; cc < 8; ++cc) {
                const int r = R0 + rr, c = C0 + cc;
                E[rr][cc] = (c <= r && r < N) ? Lm[r * CI_LD + c] : 0.0;
            }
    }
    __syncthreads();                     // the staging array is rewritten by chol_inv_dump
    SMALL_PROF(2);
    const int bad = LA ? chol_inv_lookahead<false>(E, N, col, xr) : chol_inv_regs<false>(E, N, col, xr);
    if (bad && tid == 0) info[p] = bad;
    SMALL_PROF(3);
    chol_inv_dump<LA>(E, xr, Lm);
    __syncthreads();
    SMALL_PROF(4);

    // z = W y  (W[k][c] = Lm[c][k+1])
    if (tid < N) {
        const int k = tid;
        double acc = 0.0;
        for (int c = 0; c <= k; ++c) acc = fma(Lm[c * CI_LD + k + 1], yv[c], acc);
        zv[k] = acc;
    }
    __syncthreads();
    SMALL_PROF(5);
    double cn = 0.0, al = 0.0;
    if (tid < N) {
        const int i = tid;
        const double* row = Lm + i * CI_LD + 1;
        for (int k = i; k < N; ++k) {
            const double w = row[k];
            cn = fma(w, w, cn);
            al = fma(w, zv[k], al);
        }
    }
    if constexpr (FORM == 0) {
        if (tid < N) {
            const int i = tid;
            double resid, sd;
            resid_sigma(dv[i], al, cn, resid, sd);
            av[i] = fabs(resid);
            sg[i] = sd;
            if (resid_out) resid_out[(size_t)p * N + i] = resid;
            if (sig_out) sig_out[(size_t)p * N + i] = sd;
            const int k0 = first_inside(av[i], sd, sigma_vals, S);
            atomicAdd(&hist[k0], 1);
        }
    } else {
        // K_s = kernel_func(x, x) (GP_func.py:32, no noise term) over the L part of the staging array: lower triangle at
        // Lm[r][c], c <= r; the inverse W[k][c] = Lm[c][k + 1] (c <= k) stays where it is
        double* alv = col;                          // scratch of the factorisation, free now
        double* muv = col + NB;
        if (tid < N) alv[tid] = al;
        small_cov_to_smem(xs, sq, nullptr, N, d, Lm);
        __syncthreads();
        if (tid >= NB && tid - NB < N) {            // mu_i = sum_c K_s[c][i] alpha_c                       GP_func.py:36
            const int i = tid - NB;
            double m = 0.0;
            for (int c = 0; c < N; ++c) m = fma((c >= i) ? Lm[c * CI_LD + i] : Lm[i * CI_LD + c], alv[c], m);
            muv[i] = m;
        }
        double ssq = 0.0;
        if (tid < N) {                              // sum_k (sum_{c<=k} W[k][c] K_s[c][i])^2              GP_func.py:38-39
            const int i = tid;
            for (int k = 0; k < N; ++k) {
                double vk = 0.0;
                for (int c = 0; c <= k; ++c)
                    vk = fma(Lm[c * CI_LD + k + 1], (c >= i) ? Lm[c * CI_LD + i] : Lm[i * CI_LD + c], vk);
                ssq = __dadd_rn(ssq, __dmul_rn(vk, vk));
            }
        }
        __syncthreads();
        if (tid < N) {
            const int i = tid;
            const double resid = __dsub_rn(muv[i], yv[i]);
            const double sd = sqrt(fmax(__dsub_rn(1.0, ssq), 1e-12));                     // GP_func.py:34,39-40
            if (resid_out) resid_out[(size_t)p * N + i] = resid;
            if (sig_out) sig_out[(size_t)p * N + i] = sd;
            atomicAdd(&hist[first_inside(fabs(resid), sd, sigma_vals, S)], 1);
        }
    }
    __syncthreads();
    SMALL_PROF(6);
    prefix_and_store(hist, S, counts + (size_t)p * S);
    SMALL_PROF(7);
}

int launch_small_loss(const double* X, const double* y, const double* e, const double* ls,
                      const unsigned char* active, const double* sigma_vals, int P, int N, int d, int S,
                      int* counts, int* info, double* resid_out, double* sig_out, int form, cudaStream_t st) {
    static int la = -1;               // GPK_SMALL_VARIANT = 0 | 1 (A/B runs): schedule of the factorisation
    if (la < 0) {
        const char* v = getenv("GPK_SMALL_VARIANT");
        const int want = (v && (atoi(v) == 0 || atoi(v) == 1)) ? atoi(v) : SMALL_DEFAULT_VARIANT;
        cudaError_t er = cudaFuncSetAttribute(k_small_loss<0, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SmallSmem::total);
        if (er == cudaSuccess) er = cudaFuncSetAttribute(k_small_loss<1, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SmallSmem::total);
        if (er == cudaSuccess) er = cudaFuncSetAttribute(k_small_loss<0, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SmallSmem::total);
        if (er == cudaSuccess) er = cudaFuncSetAttribute(k_small_loss<1, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SmallSmem::total);
        if (er != cudaSuccess) { set_error("k_small_loss smem attr: %s", cudaGetErrorString(er)); return 1; }
        la = want;
    }
#define GPK_SMALL_GO(F, L) k_small_loss<F, L><<<P, 256, SmallSmem::total, st>>>(X, y, e, ls, active, sigma_vals, N, d, S, counts, info, resid_out, sig_out)
    if (form) { if (la) GPK_SMALL_GO(1, 1); else GPK_SMALL_GO(1, 0); }
    else { if (la) GPK_SMALL_GO(0, 1); else GPK_SMALL_GO(0, 0); }
#undef GPK_SMALL_GO
    return check_launch(form ? "k_small_loss_ref_form" : "k_small_loss");
}

// ---------------------------------------------------------------------------------------------
// z = W y : one warp per row, W lower block-triangular with zero upper parts in the diagonal blocks.
// pu / pt (optional, per particle [Np]): the residual probe's second right-hand side, t = W u, in the same pass over W.
__global__ void __launch_bounds__(256) k_wy(const double* __restrict__ mat, int Np, int N,
                                            const double* __restrict__ y, double* __restrict__ z,
                                            const double* __restrict__ pu, double* __restrict__ pt,
                                            const unsigned char* __restrict__ active) {
    const int p = blockIdx.y;
    if (active && !active[p]) return;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int k = blockIdx.x * 8 + warp;
    if (k >= Np) return;
    const double* row = mat + (size_t)p * Np * Np + (size_t)k * Np;
    const int cend = min(k, N - 1);
    double acc = 0.0;
    if (pu == nullptr) {
        for (int c = lane; c <= cend; c += 32) acc = fma(row[c], y[c], acc);
    } else {
        const double* u = pu + (size_t)p * Np;
        double acc2 = 0.0;
        for (int c = lane; c <= cend; c += 32) {
            const double w = row[c];
            acc = fma(w, y[c], acc);
            acc2 = fma(w, u[c], acc2);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc2 += __shfl_xor_sync(0xffffffffu, acc2, o);
        if (lane == 0) pt[(size_t)p * Np + k] = acc2;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) z[(size_t)p * Np + k] = acc;
}

int launch_wy(const double* mat, int Np, int N, int P, const double* y, double* z, const double* pu, double* pt,
              const unsigned char* active, cudaStream_t st) {
    k_wy<<<dim3(Np / 8, P), 256, 0, st>>>(mat, Np, N, y, z, pu, pt, active);
    return check_launch("k_wy");
}

// cn_i = sum_{k>=i} W[k][i]^2,  alpha_i = sum_{k>=i} W[k][i] z_k : 32 columns per CTA, 8 warps sweep rows.
// pt / pw (optional, per particle [Np]): the residual probe, w = W^T t, in the same pass over W.
// One CTA per 128-column strip (= one block column: the strip starts at its diagonal block), 8 warps sweep the rows, a lane
// owns 4 consecutive columns: every load is one full 32-byte sector per lane, 1 KB contiguous per warp and row.
template <bool PROBE>
__global__ void __launch_bounds__(256) k_colnorm_alpha(const double* __restrict__ mat, int Np,
                                                       const double* __restrict__ z, double* __restrict__ cn,
                                                       double* __restrict__ alpha, const double* __restrict__ pt,
                                                       double* __restrict__ pw,
                                                       const unsigned char* __restrict__ active) {
    const int p = blockIdx.y;
    if (active && !active[p]) return;
    __shared__ double s_red[8][3][NB];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int c0 = blockIdx.x * NB + 4 * lane;
    const int kstart = blockIdx.x * NB;                    // first row of the diagonal block
    const double* M = mat + (size_t)p * Np * Np + c0;
    const double* zp = z + (size_t)p * Np;
    const double* tp = PROBE ? pt + (size_t)p * Np : nullptr;
    double a[4] = {0.0, 0.0, 0.0, 0.0}, b[4] = {0.0, 0.0, 0.0, 0.0}, c[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll 2
    for (int k = kstart + warp; k < Np; k += 8) {
        double w[4];
        asm volatile("ld.global.nc.v4.f64 {%0, %1, %2, %3}, [%4];\n"
                     : "=d"(w[0]), "=d"(w[1]), "=d"(w[2]), "=d"(w[3]) : "l"(M + (size_t)k * Np));
        const double zk = zp[k];
        const double tk = PROBE ? tp[k] : 0.0;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            a[q] = fma(w[q], w[q], a[q]);
            b[q] = fma(w[q], zk, b[q]);
            if (PROBE) c[q] = fma(w[q], tk, c[q]);
        }
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        s_red[warp][0][4 * lane + q] = a[q];
        s_red[warp][1][4 * lane + q] = b[q];
        if (PROBE) s_red[warp][2][4 * lane + q] = c[q];
    }
    __syncthreads();
    if (threadIdx.x < NB) {
        const int i = threadIdx.x;
        double ta = 0.0, tb = 0.0, tc = 0.0;
#pragma unroll
        for (int w = 0; w < 8; ++w) {
            ta += s_red[w][0][i];
            tb += s_red[w][1][i];
            if (PROBE) tc += s_red[w][2][i];
        }
        const size_t o = (size_t)p * Np + blockIdx.x * NB + i;
        cn[o] = ta;
        alpha[o] = tb;
        if (PROBE) pw[o] = tc;
    }
}

int launch_colnorm_alpha(const double* mat, int Np, int P, const double* z, double* cn, double* alpha,
                         const double* pt, double* pw, const unsigned char* active, cudaStream_t st) {
    if (pt) k_colnorm_alpha<true><<<dim3(Np / NB, P), 256, 0, st>>>(mat, Np, z, cn, alpha, pt, pw, active);
    else k_colnorm_alpha<false><<<dim3(Np / NB, P), 256, 0, st>>>(mat, Np, z, cn, alpha, pt, pw, active);
    return check_launch("k_colnorm_alpha");
}

// u[p][i] = sum_B kv[p][B][i]: the probe's K v from the per-tile partial sums k_cov_lower leaves (fixed order)
__global__ void __launch_bounds__(256) k_probe_sum(const double* __restrict__ kv, int Np, int T, double* __restrict__ u,
                                                   const unsigned char* __restrict__ active) {
    const int p = blockIdx.y;
    if (active && !active[p]) return;
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= Np) return;
    const double* src = kv + (size_t)p * T * Np + i;
    double s = 0.0;
    for (int b = 0; b < T; ++b) s += src[(size_t)b * Np];
    u[(size_t)p * Np + i] = s;
}

int launch_probe_sum(const double* kv, int Np, int T, int P, double* u, const unsigned char* active, cudaStream_t st) {
    k_probe_sum<<<dim3((Np + 255) / 256, P), 256, 0, st>>>(kv, Np, T, u, active);
    return check_launch("k_probe_sum");
}

// counts[p][s] from alpha / cn: one CTA per particle.
// pw != nullptr: also the residual probe rho_p = max_i |w_i - v_i| / max_i |v_i| (w = W^T W K v for the fixed vector
// probe_v) -> probe[p], and fail[p] = !(rho_p <= tol) (a NaN fails).  zero_inactive = 0: particles outside `active` are
// left untouched (the FP64 fallback pass recomputes only the failed ones).
__global__ void __launch_bounds__(256) k_counts(const double* __restrict__ alpha, const double* __restrict__ cn,
                                                const double* __restrict__ e, const double* __restrict__ sigma_vals,
                                                int N, int Np, int S, int* __restrict__ counts,
                                                double* __restrict__ resid_out, double* __restrict__ sig_out,
                                                const double* __restrict__ pw, double tol, double* __restrict__ probe,
                                                unsigned char* __restrict__ fail, int zero_inactive,
                                                const unsigned char* __restrict__ active) {
    __shared__ int hist[MAX_S + 1];
    __shared__ double s_max[8][2];
    const int p = blockIdx.x;
    const int tid = threadIdx.x;
    if (active && !active[p]) {
        if (zero_inactive) {
            for (int k = tid; k < S; k += 256) counts[(size_t)p * S + k] = 0;
            if (pw != nullptr && tid == 0) { probe[p] = 0.0; fail[p] = 0; }
        }
        return;
    }
    if (pw != nullptr) {
        double dmax = 0.0, vmax = 0.0;
        bool nan = false;
        for (int i = tid; i < N; i += 256) {
            const double v = probe_v(i);
            const double dlt = fabs(pw[(size_t)p * Np + i] - v);
            nan |= !(dlt == dlt);
            dmax = fmax(dmax, dlt);
            vmax = fmax(vmax, fabs(v));
        }
        if (nan) dmax = __longlong_as_double(0x7ff0000000000000LL);        // +inf: fmax would drop a NaN
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            dmax = fmax(dmax, __shfl_xor_sync(0xffffffffu, dmax, o));
            vmax = fmax(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
        }
        if ((tid & 31) == 0) { s_max[tid >> 5][0] = dmax; s_max[tid >> 5][1] = vmax; }
        __syncthreads();
        if (tid == 0) {
            for (int w = 1; w < 8; ++w) { dmax = fmax(dmax, s_max[w][0]); vmax = fmax(vmax, s_max[w][1]); }
            const double rho = dmax / vmax;
            probe[p] = rho;
            fail[p] = (rho <= tol) ? 0 : 1;
        }
    }
    for (int k = tid; k <= S; k += 256) hist[k] = 0;
    __syncthreads();
    for (int i = tid; i < N; i += 256) {
        const double ee = e[i];
        const double d = __dmul_rn(ee, ee);
        double resid, sd;
        resid_sigma(d, alpha[(size_t)p * Np + i], cn[(size_t)p * Np + i], resid, sd);
        if (resid_out) resid_out[(size_t)p * N + i] = resid;
        if (sig_out) sig_out[(size_t)p * N + i] = sd;
        atomicAdd(&hist[first_inside(fabs(resid), sd, sigma_vals, S)], 1);
    }
    __syncthreads();
    prefix_and_store(hist, S, counts + (size_t)p * S);
}

int launch_counts(const double* alpha, const double* cn, const double* e, const double* sigma_vals, int P, int N,
                  int Np, int S, int* counts, double* resid_out, double* sig_out, const double* pw, double tol,
                  double* probe, unsigned char* fail, int zero_inactive, const unsigned char* active, cudaStream_t st) {
    k_counts<<<P, 256, 0, st>>>(alpha, cn, e, sigma_vals, N, Np, S, counts, resid_out, sig_out, pw, tol, probe, fail,
                                zero_inactive, active);
    return check_launch("k_counts");
}

// rho_p of the residual probe for callers without a count stage (gpk_fit): one CTA per particle
__global__ void __launch_bounds__(256) k_probe_value(const double* __restrict__ pw, int N, int Np, double* __restrict__ probe,
                                                     unsigned char* __restrict__ fail, double tol,
                                                     const unsigned char* __restrict__ active) {
    __shared__ double s_max[8][2];
    const int p = blockIdx.x, tid = threadIdx.x;
    if (active && !active[p]) return;
    double dmax = 0.0, vmax = 0.0;
    bool nan = false;
    for (int i = tid; i < N; i += 256) {
        const double v = probe_v(i);
        const double dlt = fabs(pw[(size_t)p * Np + i] - v);
        nan |= !(dlt == dlt);
        dmax = fmax(dmax, dlt);
        vmax = fmax(vmax, fabs(v));
    }
    if (nan) dmax = __longlong_as_double(0x7ff0000000000000LL);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        dmax = fmax(dmax, __shfl_xor_sync(0xffffffffu, dmax, o));
        vmax = fmax(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
    }
    if ((tid & 31) == 0) { s_max[tid >> 5][0] = dmax; s_max[tid >> 5][1] = vmax; }
    __syncthreads();
    if (tid == 0) {
        for (int w = 1; w < 8; ++w) { dmax = fmax(dmax, s_max[w][0]); vmax = fmax(vmax, s_max[w][1]); }
        const double rho = dmax / vmax;
        probe[p] = rho;
        fail[p] = (rho <= tol) ? 0 : 1;
    }
}

int launch_probe_value(const double* pw, int N, int Np, int P, double* probe, unsigned char* fail, double tol,
                       const unsigned char* active, cudaStream_t st) {
    k_probe_value<<<P, 256, 0, st>>>(pw, N, Np, probe, fail, tol, active);
    return check_launch("k_probe_value");
}

// the FP64 fallback pass starts from a clean status for the particles it re-evaluates
__global__ void k_clear_info(int* __restrict__ info, const unsigned char* __restrict__ mask, int P) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p < P && mask[p]) info[p] = 0;
}

int launch_clear_info(int* info, const unsigned char* mask, int P, cudaStream_t st) {
    k_clear_info<<<(P + 255) / 256, 256, 0, st>>>(info, mask, P);
    return check_launch("k_clear_info");
}

// sigma_b = sqrt(clip(1 - sum_I partial[I][b], 1e-12))     GP_func.py:34,39-40
__global__ void k_predict_finalize(const double* __restrict__ partial, int T, int Bp, int B,
                                   double* __restrict__ sigma_out) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    double s = 0.0;
    for (int I = 0; I < T; ++I) s += partial[(size_t)I * Bp + b];
    sigma_out[b] = sqrt(fmax(1.0 - s, 1e-12));
}

int launch_predict_finalize(const double* partial, int T, int Bp, int B, double* sigma_out, cudaStream_t st) {
    k_predict_finalize<<<(B + 255) / 256, 256, 0, st>>>(partial, T, Bp, B, sigma_out);
    return check_launch("k_predict_finalize");
}

}  // namespace gpk

#ifdef GPK_SMALL_PROFILE
extern "C" __attribute__((visibility("default"))) int gpk_debug_small_prof(long long* out16) {
    if (cudaDeviceSynchronize() != cudaSuccess) return 1;
    return cudaMemcpyFromSymbol(out16, gpk::g_small_prof, sizeof(long long) * 16) == cudaSuccess ? 0 : 1;
}
#endif
