// Text side of the result tables: float64 table -> CSV bytes exactly as the reference writes them with
// pandas.DataFrame.to_csv(index=False) (GP_fit.py:63,81,104; calc_prob_surf.py:84).  pandas formats a float64 column
// with numpy's astype(str): the shortest digit string that reads back to the same double, positional for
// 1e-4 <= |x| < 1e16 and scientific otherwise ("1e-05", "1.5e+16": at least two exponent digits), "inf" / "-inf" as is,
// NaN as the empty field, "-0.0" keeps its sign, whole numbers end in ".0".  std::to_chars supplies the shortest digits;
// the layout rules are applied here.  Rows are split over host threads; every thread formats into its worst-case slot
// of the output buffer and the slots are closed up afterwards.
#include <algorithm>
#include <charconv>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

#include "../../include/gpk.h"

namespace {

constexpr int kMaxField = 26;      // "-2.2250738585072014e-308" is 24 characters

// writes one value, returns the end pointer
inline char* format_value(double v, char* p) {
    if (std::isnan(v)) return p;                                  // na_rep = ""
    if (std::isinf(v)) {
        if (v < 0) *p++ = '-';
        std::memcpy(p, "inf", 3);
        return p + 3;
    }
    if (std::signbit(v)) { *p++ = '-'; v = -v; }
    if (v == 0.0) { std::memcpy(p, "0.0", 3); return p + 3; }
    // shortest round-trip digits in scientific form: d[.ddd]e[+-]XX
    char buf[40];
    auto res = std::to_chars(buf, buf + sizeof(buf), v, std::chars_format::scientific);
    char* e = buf;
    while (*e != 'e') ++e;
    int exp10 = 0;
    {
        const char* q = e + 1;
        const bool neg = (*q == '-');
        if (*q == '-' || *q == '+') ++q;
        for (; q < res.ptr; ++q) exp10 = exp10 * 10 + (*q - '0');
        if (neg) exp10 = -exp10;
    }
    char digits[24];
    int nd = 0;
    for (const char* q = buf; q < e; ++q)
        if (*q != '.') digits[nd++] = *q;
    if (exp10 >= -4 && exp10 < 16) {                              // positional
        if (exp10 >= 0) {
            const int int_digits = exp10 + 1;
            for (int i = 0; i < int_digits; ++i) *p++ = (i < nd) ? digits[i] : '0';
            *p++ = '.';
            if (nd > int_digits) {
                std::memcpy(p, digits + int_digits, nd - int_digits);
                p += nd - int_digits;
            } else {
                *p++ = '0';
            }
        } else {
            *p++ = '0';
            *p++ = '.';
            for (int i = 0; i < -exp10 - 1; ++i) *p++ = '0';
            std::memcpy(p, digits, nd);
            p += nd;
        }
        return p;
    }
    *p++ = digits[0];                                             // scientific: no ".0" for a single digit
    if (nd > 1) {
        *p++ = '.';
        std::memcpy(p, digits + 1, nd - 1);
        p += nd - 1;
    }
    *p++ = 'e';
    *p++ = (exp10 < 0) ? '-' : '+';
    int a = std::abs(exp10);
    if (a >= 100) { *p++ = char('0' + a / 100); a %= 100; *p++ = char('0' + a / 10); *p++ = char('0' + a % 10); }
    else { *p++ = char('0' + a / 10); *p++ = char('0' + a % 10); }
    return p;
}

}  // namespace

extern "C" {

size_t gpk_csv_bound(long long rows, int cols) {
    if (rows <= 0 || cols <= 0) return 0;
    return (size_t)rows * (size_t)cols * (kMaxField + 1);
}

long long gpk_csv_format_f64(const double* table, long long rows, int cols, long long ld, char* out, size_t out_cap,
                             int threads) {
    if (rows < 0 || cols <= 0 || ld < cols || (rows > 0 && (!table || !out))) return -1;
    if (out_cap < gpk_csv_bound(rows, cols)) return -1;
    if (rows == 0) return 0;
    int nt = threads > 0 ? threads : (int)std::thread::hardware_concurrency();
    nt = (int)std::max<long long>(1, std::min<long long>(nt, (rows + 4095) / 4096));
    const size_t row_slot = (size_t)cols * (kMaxField + 1);
    std::vector<long long> first(nt + 1);
    std::vector<size_t> len(nt, 0);
    for (int t = 0; t <= nt; ++t) first[t] = rows * t / nt;
    auto work = [&](int t) {
        char* p = out + (size_t)first[t] * row_slot;
        char* const p0 = p;
        for (long long r = first[t]; r < first[t + 1]; ++r) {
            const double* row = table + r * ld;
            for (int c = 0; c < cols; ++c) {
                p = format_value(row[c], p);
                *p++ = (c + 1 < cols) ? ',' : '\n';
            }
        }
        len[t] = (size_t)(p - p0);
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < nt; ++t) pool.emplace_back(work, t);
    work(0);
    for (auto& th : pool) th.join();
    size_t total = len[0];
    for (int t = 1; t < nt; ++t) {                                // close the gaps between the threads' slots
        std::memmove(out + total, out + (size_t)first[t] * row_slot, len[t]);
        total += len[t];
    }
    return (long long)total;
}

size_t gpk_csv_surface_bound(long long rows, int ndims, int points) {
    if (rows <= 0 || ndims < 0 || points <= 0) return 0;
    return (size_t)rows * (size_t)points * (size_t)(ndims + 2) * (kMaxField + 1);
}

// Long-format surface block (calc_prob_surf.py:80-81: every merged row emits `points` output rows
// [coords..., quantity, prob]) straight from the reduction's outputs: the text of the coordinates is formatted once
// per merged row and copied in front of each of its output rows, only quantity and prob are formatted per output row.
// Same bytes as gpk_csv_format_f64 on the expanded (rows * points, ndims + 2) table.
long long gpk_csv_format_surface(const double* coords, const double* y, const double* p, long long rows, int ndims,
                                 int points, char* out, size_t out_cap, int threads) {
    if (rows < 0 || ndims < 0 || points <= 0 || (rows > 0 && (!y || !p || !out || (ndims > 0 && !coords)))) return -1;
    if (out_cap < gpk_csv_surface_bound(rows, ndims, points)) return -1;
    if (rows == 0) return 0;
    int nt = threads > 0 ? threads : (int)std::thread::hardware_concurrency();
    nt = (int)std::max<long long>(1, std::min<long long>(nt, (rows + 63) / 64));
    const size_t row_slot = (size_t)points * (size_t)(ndims + 2) * (kMaxField + 1);
    std::vector<long long> first(nt + 1);
    std::vector<size_t> len(nt, 0);
    for (int t = 0; t <= nt; ++t) first[t] = rows * t / nt;
    auto work = [&](int t) {
        char* q = out + (size_t)first[t] * row_slot;
        char* const q0 = q;
        std::vector<char> prefix((size_t)ndims * (kMaxField + 1) + 1);
        for (long long r = first[t]; r < first[t + 1]; ++r) {
            char* e = prefix.data();
            for (int c = 0; c < ndims; ++c) {
                e = format_value(coords[r * ndims + c], e);
                *e++ = ',';
            }
            const size_t plen = (size_t)(e - prefix.data());
            const double* yr = y + r * (long long)points;
            const double* pr = p + r * (long long)points;
            for (int j = 0; j < points; ++j) {
                std::memcpy(q, prefix.data(), plen);
                q += plen;
                q = format_value(yr[j], q);
                *q++ = ',';
                q = format_value(pr[j], q);
                *q++ = '\n';
            }
        }
        len[t] = (size_t)(q - q0);
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < nt; ++t) pool.emplace_back(work, t);
    work(0);
    for (auto& th : pool) th.join();
    size_t total = len[0];
    for (int t = 1; t < nt; ++t) {
        std::memmove(out + total, out + (size_t)first[t] * row_slot, len[t]);
        total += len[t];
    }
    return (long long)total;
}

}  // extern "C"
