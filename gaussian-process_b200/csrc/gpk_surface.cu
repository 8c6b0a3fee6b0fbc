// Kernel (d): fused inf-masked experiment combination + probability-surface reduction.
// Reference: calc_prob_surf.py:67-81 (row loop of generate_prob_surf) and :15-30 (sum_gaussians).
// One warp per merged-table row:
//   * keep the finite payload entries in column order (np.isfinite mask, :69), require an even
//     count >= 2 (:71), de-interleave the *compacted* list into means / stds (:72-73);
//   * y-axis = np.linspace(min(mu-3sd), max(mu+3sd), 100): y_i = fl(fl(i*step) + start), last = stop;
//   * dy = |y_max - y_min| / 100 (the reference divides by the point count, not by 99, :23);
//   * p(y) = (1/E') sum_e [ Phi((y+dy/2-mu_e)/sd_e) - Phi((y-dy/2-mu_e)/sd_e) ], accumulated in the
//     reference's order (+upper, -lower per experiment, :25-27).
// Phi follows scipy.special.ndtr (Cephes): 0.5 + 0.5 erf(x/sqrt2) near 0, erfc in the tails.
#include "gpk_common.cuh"

namespace gpk {

// payload columns per row = 2 * experiments.  The compacted finite entries of a row live in shared memory, 8 rows (warps)
// per CTA: static up to 128 columns, dynamic (opt-in, up to 200 KB) beyond; 3072 columns = 1536 experiments is the limit.
constexpr int SURF_STATIC_COLS = 128;
constexpr int SURF_MAX_COLS = 3072;

__device__ __forceinline__ double ndtr(double a) {
    if (isnan(a)) return a;
    const double x = __dmul_rn(a, 0.70710678118654752440);
    const double z = fabs(x);
    if (z < 0.70710678118654752440) return 0.5 + 0.5 * erf(x);
    double yv = 0.5 * erfc(z);
    if (x > 0) yv = 1.0 - yv;
    return yv;
}

// scipy's rv_continuous.cdf wrapper: NaN for scale <= 0 or NaN arguments
__device__ __forceinline__ double norm_cdf(double x, double loc, double scale) {
    if (!(scale > 0.0)) return nan("");
    return ndtr(__ddiv_rn(__dsub_rn(x, loc), scale));
}

__global__ void __launch_bounds__(256) k_prob_surface(const double* __restrict__ table, long long R, int ncols,
                                                      int npoints, double* __restrict__ y_out,
                                                      double* __restrict__ p_out, int* __restrict__ valid_out) {
    extern __shared__ double s_vals[];           // [8][pitch], pitch = ncols rounded up to 32
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long r = (long long)blockIdx.x * 8 + warp;
    if (r >= R) return;
    const double* row = table + r * ncols;
    double* sv = s_vals + (size_t)warp * ((ncols + 31) & ~31);

    int cnt = 0;
    for (int base = 0; base < ncols; base += 32) {
        const int c = base + lane;
        const double v = (c < ncols) ? row[c] : INFINITY;
        const bool fin = isfinite(v);
        const unsigned m = __ballot_sync(0xffffffffu, fin);
        if (fin) sv[cnt + __popc(m & ((1u << lane) - 1u))] = v;
        cnt += __popc(m);
    }
    __syncwarp();
    const bool valid = (cnt >= 2) && ((cnt & 1) == 0);
    if (lane == 0) valid_out[r] = valid ? 1 : 0;
    if (!valid) return;
    const int ne = cnt >> 1;

    double lo = INFINITY, hi = -INFINITY;
    for (int i = lane; i < ne; i += 32) {
        const double mu = sv[2 * i], t3 = __dmul_rn(3.0, sv[2 * i + 1]);
        lo = fmin(lo, __dsub_rn(mu, t3));
        hi = fmax(hi, __dadd_rn(mu, t3));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        lo = fmin(lo, __shfl_xor_sync(0xffffffffu, lo, o));
        hi = fmax(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    const double delta = __dsub_rn(hi, lo);
    const double step = __ddiv_rn(delta, (double)(npoints - 1));
    const double dy = __ddiv_rn(fabs(__dsub_rn(hi, lo)), (double)npoints);
    const double half = __ddiv_rn(dy, 2.0);
    const double inv_ne = (double)ne;

    for (int i = lane; i < npoints; i += 32) {
        double yi;
        if (i == npoints - 1 && npoints > 1) yi = hi;
        else yi = __dadd_rn(__dmul_rn((double)i, step), lo);
        const double yu = __dadd_rn(yi, half), yl = __dsub_rn(yi, half);
        double z = 0.0;
        for (int k = 0; k < ne; ++k) {
            const double mu = sv[2 * k], sd = sv[2 * k + 1];
            z = __dadd_rn(z, norm_cdf(yu, mu, sd));
            z = __dsub_rn(z, norm_cdf(yl, mu, sd));
        }
        y_out[r * npoints + i] = yi;
        p_out[r * npoints + i] = __ddiv_rn(z, inv_ne);
    }
}

// Stand-alone sum_gaussians(temp_y, temp_gaus) (calc_prob_surf.py:15-30) at caller-supplied points;
// half_dy = (|max(y) - min(y)| / len(y)) / 2 is formed by the caller exactly as :23 does.
__global__ void k_sum_gaussians(const double* __restrict__ y, int n, const double* __restrict__ gaus, int ng,
                                double half_dy, double* __restrict__ p_out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int ne = ng / 2;                      // int(len(temp_gaus) / 2)
    const double yu = __dadd_rn(y[i], half_dy), yl = __dsub_rn(y[i], half_dy);
    double z = 0.0;
    for (int k = 0; k < ne; ++k) {
        z = __dadd_rn(z, norm_cdf(yu, gaus[2 * k], gaus[2 * k + 1]));
        z = __dsub_rn(z, norm_cdf(yl, gaus[2 * k], gaus[2 * k + 1]));
    }
    p_out[i] = __ddiv_rn(z, (double)ne);
}

int launch_sum_gaussians(const double* y, int n, const double* gaus, int ng, double half_dy, double* p_out,
                         cudaStream_t st) {
    if (n <= 0) return 0;
    k_sum_gaussians<<<(n + 127) / 128, 128, 0, st>>>(y, n, gaus, ng, half_dy, p_out);
    return check_launch("k_sum_gaussians");
}

// ---- compaction of the emitting rows (calc_prob_surf.py:71 drops rows; the output is compacted) ----------------
// dst[r] = number of emitting rows before r (exclusive prefix sum of valid), count[0] = their total.  One CTA: a pass of the
// surface reduction holds at most a few 10^5 rows.
__global__ void __launch_bounds__(1024) k_surface_scan(const int* __restrict__ valid, long long R, long long* __restrict__ dst,
                                                       long long* __restrict__ count) {
    __shared__ long long s_warp[32];
    __shared__ long long s_base;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_base = 0;
    __syncthreads();
    for (long long r0 = 0; r0 < R; r0 += 1024) {
        const long long r = r0 + tid;
        const int v = (r < R && valid[r]) ? 1 : 0;
        int incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            long long w = s_warp[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const long long t = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += t;
            }
            s_warp[lane] = w;                 // inclusive over warps
        }
        __syncthreads();
        const long long before = s_base + (warp ? s_warp[warp - 1] : 0) + (incl - v);
        if (r < R) dst[r] = before;
        __syncthreads();
        if (tid == 0) s_base += s_warp[31];
        __syncthreads();
    }
    if (tid == 0) count[0] = s_base;
}

// emitting row r -> row dst[r] of the compacted outputs; one warp per row
__global__ void __launch_bounds__(256) k_surface_gather(const int* __restrict__ valid, const long long* __restrict__ dst,
                                                        long long R, int npoints, const double* __restrict__ y,
                                                        const double* __restrict__ p, double* __restrict__ yc,
                                                        double* __restrict__ pc, long long* __restrict__ idx) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long r = (long long)blockIdx.x * 8 + warp;
    if (r >= R || !valid[r]) return;
    const long long q = dst[r];
    for (int i = lane; i < npoints; i += 32) {
        yc[q * npoints + i] = y[r * npoints + i];
        pc[q * npoints + i] = p[r * npoints + i];
    }
    if (lane == 0) idx[q] = r;
}

int launch_surface_scan(const int* valid, long long R, long long* dst, long long* count, cudaStream_t st) {
    k_surface_scan<<<1, 1024, 0, st>>>(valid, R, dst, count);
    return check_launch("k_surface_scan");
}

int launch_surface_gather(const int* valid, const long long* dst, long long R, int npoints, const double* y, const double* p,
                          double* yc, double* pc, long long* idx, cudaStream_t st) {
    if (R <= 0) return 0;
    const long long blocks = (R + 7) / 8;
    if (blocks > 0x7fffffffLL) { set_error("surface_gather: too many rows"); return 2; }
    k_surface_gather<<<(unsigned)blocks, 256, 0, st>>>(valid, dst, R, npoints, y, p, yc, pc, idx);
    return check_launch("k_surface_gather");
}

int launch_prob_surface(const double* table, long long R, int ncols, int npoints, double* y_out, double* p_out,
                        int* valid_out, cudaStream_t st) {
    if (ncols > SURF_MAX_COLS || ncols < 0) {
        set_error("prob_surface: %d payload columns (%d experiments); at most %d columns (%d experiments) fit the per-row shared-memory staging",
                  ncols, ncols / 2, SURF_MAX_COLS, SURF_MAX_COLS / 2);
        return 2;
    }
    if (R <= 0) return 0;
    const long long blocks = (R + 7) / 8;
    if (blocks > 0x7fffffffLL) { set_error("prob_surface: too many rows"); return 2; }
    const int pitch = (ncols + 31) & ~31;
    const size_t smem = (size_t)8 * (pitch < SURF_STATIC_COLS ? SURF_STATIC_COLS : pitch) * sizeof(double);
    static size_t configured = 48 * 1024;
    if (smem > configured) {
        cudaError_t er = cudaFuncSetAttribute(k_prob_surface, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (er != cudaSuccess) { set_error("k_prob_surface smem attr: %s", cudaGetErrorString(er)); return 1; }
        configured = smem;
    }
    k_prob_surface<<<(unsigned)blocks, 256, smem, st>>>(table, R, ncols, npoints, y_out, p_out, valid_out);
    return check_launch("k_prob_surface");
}

}  // namespace gpk
