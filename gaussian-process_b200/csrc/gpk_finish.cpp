// Host side of the loss finish (find_len_scales.py:163-166) for the PSO loop: from the int32 count matrix the GPU returns to
// the Wasserstein-style integral, one pass per particle instead of NumPy's seven passes with temporaries.
//   measured_s = counts[p][s] / n_points (looked up in measured_of[k] = k / n_points, k = 0..n_points: the quotient a
//   count can take is tabulated once by the caller, with NumPy's own division);  y_s = |measured_s - expected_s|;
//   W_p = np.trapz(y, sigma_vals) = sum_s d_s * (y_{s+1} + y_s) / 2.0,  d = diff(sigma_vals)
// The terms are formed with NumPy's operation order and summed with NumPy's pairwise summation (8 running sums per
// block of <= 128 terms, halves split at multiples of 8), so the result is the bit pattern np.trapz returns; the Python
// caller checks that against NumPy on its first evaluation and keeps the NumPy expressions if they ever differ.
#include <cmath>
#include <cstdint>

#include "../../include/gpk.h"

namespace {

double pairwise_sum(const double* a, long long n) {
    if (n < 8) {
        double res = 0.0;
        for (long long i = 0; i < n; ++i) res += a[i];
        return res;
    }
    if (n <= 128) {
        double r[8];
        for (int k = 0; k < 8; ++k) r[k] = a[k];
        long long i;
        for (i = 8; i < n - (n % 8); i += 8)
            for (int k = 0; k < 8; ++k) r[k] += a[i + k];
        double res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
        for (; i < n; ++i) res += a[i];
        return res;
    }
    long long n2 = n / 2;
    n2 -= n2 % 8;
    return pairwise_sum(a, n2) + pairwise_sum(a + n2, n - n2);
}

}  // namespace

extern "C" int gpk_finish_wasserstein(const int* counts, long long P, int S, const double* measured_of, int n_points,
                                      const double* expected, const double* d_sigma, double* scratch, double* out) {
    if (P < 0 || S < 2 || n_points < 1 || !counts || !measured_of || !expected || !d_sigma || !scratch || !out) return 2;
    double* t = scratch;              // S - 1 terms (scratch holds 2 S doubles for older callers; S - 1 are used)
    for (long long p = 0; p < P; ++p) {
        const int* c = counts + p * (long long)S;
        // one pass: y_s = |measured - expected| and t_s = d_s * (y_{s+1} + y_s) / 2.0 (NumPy's operation order; the division by
        // 2.0 is exact either way); a count outside 0..n_points is remembered and reported after the loop
        unsigned bad = (unsigned)c[0] > (unsigned)n_points;
        double y_prev = std::fabs(measured_of[bad ? 0 : c[0]] - expected[0]);
        for (int s = 1; s < S; ++s) {
            const unsigned k = (unsigned)c[s];
            const unsigned oob = k > (unsigned)n_points;
            bad |= oob;
            const double y_cur = std::fabs(measured_of[oob ? 0u : k] - expected[s]);
            t[s - 1] = d_sigma[s - 1] * (y_cur + y_prev) / 2.0;
            y_prev = y_cur;
        }
        if (bad) return 3;
        out[p] = pairwise_sum(t, S - 1);
    }
    return 0;
}
