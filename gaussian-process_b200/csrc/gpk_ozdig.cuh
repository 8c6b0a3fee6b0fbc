// Error-free digit split of FP64 tiles into the int8 operand records of the digit GEMM (gpk_oz.cu) -- shared by the slicer
// kernels there and by the diagonal-block kernel (gpk_diag.cu), which slices the inverse diagonal blocks it has just written.
#pragma once
#include "gpk_common.cuh"

namespace gpk {

constexpr int OZ_NS = 7;                        // int8 digits per FP64 value
constexpr int OZ_DIGIT = 4096;                  // one digit of a 128-row block x 32 k: [16 row groups][256 B]
constexpr int OZ_DIGIT_HALF = 2048;             // 64 rows
constexpr int OZ_CHUNK = OZ_NS * OZ_DIGIT;      // 128 rows x 32 k x 7 digits = 28672 B
constexpr int OZ_HALF = OZ_NS * OZ_DIGIT_HALF;  // 64 rows, digits back to back (shared-memory image of B)
constexpr int OZ_STAGE = OZ_CHUNK + OZ_HALF;    // A (128 rows) + B (64 rows)

constexpr long long OZ_BIAS = 0x0000808080808080LL;

// 256-bit global accesses (32-byte aligned): one full sector per lane
__device__ __forceinline__ void stg_f64x4(double* p, double a, double b, double c, double d) {
    asm volatile("st.global.v4.f64 [%0], {%1, %2, %3, %4};\n" ::"l"(p), "d"(a), "d"(b), "d"(c), "d"(d) : "memory");
}
__device__ __forceinline__ void ldg_f64x4(const double* p, double& a, double& b, double& c, double& d) {
    asm volatile("ld.global.v4.f64 {%0, %1, %2, %3}, [%4];\n" : "=d"(a), "=d"(b), "=d"(c), "=d"(d) : "l"(p) : "memory");
}


// byte b of four 32-bit words -> one word (digit b of four consecutive k)
__device__ __forceinline__ uint32_t pack_digit(const uint32_t* w, int b) {
    const uint32_t sel = (uint32_t)b | ((uint32_t)(4 + b) << 4);
    const uint32_t t0 = __byte_perm(w[0], w[1], sel);
    const uint32_t t1 = __byte_perm(w[2], w[3], sel);
    return __byte_perm(t0, t1, 0x5410);
}


// ---- slicers: FP64 128 x 128 tile -> 4 chunk records ----------------------------------------------------------

// u = x * rs[row] * rconst * cs[col];  |u| <= 0.98 is the caller's contract (a-priori bounds / measured maxima)
__device__ __forceinline__ void oz_slice_tile(const double* __restrict__ src, int ld, const double* __restrict__ rs,
                                              double rconst, const double* __restrict__ cs, uint8_t* __restrict__ dst) {
    for (int unit = threadIdx.x; unit < 1024; unit += blockDim.x) {
        const int kcl = unit >> 8, rg = (unit >> 4) & 15, kg = (unit >> 3) & 1, r8 = unit & 7;
        const int row = rg * 8 + r8, col0 = kcl * 32 + kg * 16;
        const double* sp = src + (size_t)row * ld + col0;
        const double rsv = (rs ? rs[row] : 1.0) * rconst * 0x1p55;
        uint32_t lo[16], hi[16];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            double x[4];
            ldg_f64x4(sp + 4 * j, x[0], x[1], x[2], x[3]);          // one full 32-byte sector per lane
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                double u = x[i] * rsv;
                if (cs) u *= cs[col0 + 4 * j + i];
                const long long v = (__double2ll_rn(u) + OZ_BIAS) ^ OZ_BIAS;
                lo[4 * j + i] = (uint32_t)v;
                hi[4 * j + i] = (uint32_t)(v >> 32);
            }
        }
        uint8_t* d = dst + (size_t)kcl * OZ_CHUNK + rg * 256 + kg * 128 + r8 * 16;
#pragma unroll
        for (int s = 0; s < OZ_NS; ++s) {
            const int bytepos = 6 - s;                 // digit s = byte (6 - s) of the biased fixed-point word
            const uint32_t* w = (bytepos >= 4) ? hi : lo;
            const int b = bytepos & 3;
            uint4 o;
            o.x = pack_digit(w + 0, b);
            o.y = pack_digit(w + 4, b);
            o.z = pack_digit(w + 8, b);
            o.w = pack_digit(w + 12, b);
            *reinterpret_cast<uint4*>(d + s * OZ_DIGIT) = o;
        }
    }
}

// smallest power of two sigma with bound / sigma <= 0.98 (the top digit then stays inside [-126, 126])
__device__ __forceinline__ double oz_scale_for(double bound) {
    if (!(bound > 0.0) || !(bound < 1e300)) return 1.0;
    double sig = scalbn(1.0, ilogb(bound) + 1);
    if (bound > 0.98 * sig) sig *= 2.0;
    return sig;
}


// Digits of the inverse diagonal block D_J = L_JJ^-1 for the two panel contractions, with measured row scales:
//   blockIdx.x = 0: D_J[n][j] * rowsig_j   (B operand of the TRSM  L_IJ = S_IJ D_J^T, whose A digits carry 1/rowsig_j)
//   blockIdx.x = 1: D_J^T[n][j]            (A operand of the phase-2 panel  Mt = D_J^T L_RJ^T)
// dig: [P][T][2][4 chunks], dsc: [P][T][2][NB] row scales sigma_n.
// Device routine for one operand set (`which`), called by all 256 threads of a CTA; inv: NB doubles of shared scratch.  The
// diagonal-block kernel calls it for both sets right after it has written D_J and D_J^T (gpk_diag.cu).
// dinv / dinvT are NOT restrict-qualified: the calling kernel has just written them, the reads below must be ordinary
// coherent global loads (after the caller's block barrier), never the read-only data path.
__device__ __forceinline__ void oz_slice_d_block(const double* dinv, const double* dinvT, int T, int J,
                                                 int p, int which, const double* __restrict__ rowsig,
                                                 uint8_t* __restrict__ dig, double* __restrict__ dsc, double* inv) {
    const double* src = (which ? dinvT : dinv) + ((size_t)p * T + J) * NB * NB;
    const double* cs = which ? nullptr : rowsig + J * NB;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int n = warp; n < NB; n += 8) {
        double m = 0.0;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int j = lane + 32 * q;
            m = fmax(m, fabs(src[n * NB + j] * (cs ? cs[j] : 1.0)));
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
        if (lane == 0) {
            const double sig = oz_scale_for(m);
            inv[n] = 1.0 / sig;
            dsc[(((size_t)p * T + J) * 2 + which) * NB + n] = sig;
        }
    }
    __syncthreads();
    oz_slice_tile(src, NB, inv, 1.0, cs, dig + (((size_t)p * T + J) * 2 + which) * (4 * (size_t)OZ_CHUNK));
}


}  // namespace gpk
