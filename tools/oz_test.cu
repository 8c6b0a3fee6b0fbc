// Standalone check + timing of the emulated-FP64 (int8 tcgen05) primitives of libgpk, without torch.
//   build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tools/oz_test tools/oz_test.cu -Lgaussian-process_b200 -lgpk -Xlinker -rpath='$ORIGIN/../gaussian-process_b200'
//   run  : tools/oz_test [perf]
// 1. slices random FP64 operands on the GPU and compares the digit records byte by byte with a host slicer;
// 2. runs the digit GEMM, compares the raw int32 accumulators with host integer sums and the FP64 result with a
//    long-double reference; tries alternative shared-memory descriptors when the default one mismatches;
// 3. (perf) times a large batch of tiles.
#include <cuda_runtime.h>
#include <unistd.h>

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <vector>

#include "../include/gpk.h"

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(2); } } while (0)
#define GK(x) do { int s_ = (x); if (s_ != 0) { printf("gpk error %d: %s at %s:%d\n", s_, gpk_last_error_string(), __FILE__, __LINE__); exit(3); } } while (0)

static const int NS = 7, DIGIT = 4096, CHUNK = NS * DIGIT;
static const long long BIAS = 0x0000808080808080LL;

// host restatement of the record layout: digit s of element (r, k) of a rows x K matrix
static size_t rec_offset(int r, int k, int s, int kch) {
    const int rb = r / 128, rr = r % 128, kc = k / 32, kk = k % 32;
    return ((size_t)rb * kch + kc) * CHUNK + (size_t)s * DIGIT + (rr / 8) * 256 + (kk / 16) * 128 + (rr % 8) * 16 + (kk % 16);
}
static void host_digits(double x, double rowinv, int8_t* d) {
    long long v = llrint(x * rowinv * 0x1p55);
    v = (v + BIAS) ^ BIAS;
    for (int s = 0; s < NS; ++s) d[s] = (int8_t)((v >> (8 * (6 - s))) & 0xFF);
}

struct Problem {
    int M, N, K;
    std::vector<double> A, B, sa, sb, ia, ib;
};

static Problem make_problem(int M, int N, int K, unsigned seed) {
    Problem p{M, N, K};
    std::mt19937_64 rng(seed);
    std::uniform_real_distribution<double> U(-1.0, 1.0);
    p.A.resize((size_t)M * K); p.B.resize((size_t)N * K);
    p.sa.resize(M); p.sb.resize(N); p.ia.resize(M); p.ib.resize(N);
    auto fill = [&](std::vector<double>& X, std::vector<double>& sg, std::vector<double>& inv, int R) {
        for (int r = 0; r < R; ++r) {
            const double amp = std::ldexp(1.0, (int)(rng() % 9) - 4) * (1.0 + 0.3 * U(rng));
            double mx = 0;
            for (int k = 0; k < K; ++k) { double v = amp * U(rng) * ((k % 7 == 0) ? 1e-6 : 1.0); X[(size_t)r * K + k] = v; mx = std::fmax(mx, std::fabs(v)); }
            sg[r] = std::ldexp(1.0, std::ilogb(mx) + 1);
            if (mx > 0.98 * sg[r]) sg[r] *= 2;
            inv[r] = 1.0 / sg[r];
        }
    };
    fill(p.A, p.sa, p.ia, M);
    fill(p.B, p.sb, p.ib, N);
    return p;
}

int main(int argc, char** argv) {
    alarm(170);
    const bool perf = argc > 1 && !strcmp(argv[1], "perf");
    printf("libgpk version %d\n", gpk_version());
    int fails = 0;
    {
        const int M = 256, N = 128, K = 256, kch = K / 32;
        Problem p = make_problem(M, N, K, 1);
        double *dA, *dB, *dsa, *dsb, *dia, *dib, *dC;
        uint8_t *sA, *sB;
        int* dbg;
        const size_t bA = gpk_oz_slice_bytes(M, K), bB = gpk_oz_slice_bytes(N, K);
        const int tiles = (M / 128) * (N / 64);
        CK(cudaMalloc(&dA, p.A.size() * 8)); CK(cudaMalloc(&dB, p.B.size() * 8));
        CK(cudaMalloc(&dsa, M * 8)); CK(cudaMalloc(&dsb, N * 8)); CK(cudaMalloc(&dia, M * 8)); CK(cudaMalloc(&dib, N * 8));
        CK(cudaMalloc(&dC, (size_t)M * N * 8)); CK(cudaMalloc(&sA, bA)); CK(cudaMalloc(&sB, bB));
        CK(cudaMalloc(&dbg, (size_t)tiles * NS * 128 * 64 * 4));
        CK(cudaMemcpy(dA, p.A.data(), p.A.size() * 8, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(dB, p.B.data(), p.B.size() * 8, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(dsa, p.sa.data(), M * 8, cudaMemcpyHostToDevice)); CK(cudaMemcpy(dsb, p.sb.data(), N * 8, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(dia, p.ia.data(), M * 8, cudaMemcpyHostToDevice)); CK(cudaMemcpy(dib, p.ib.data(), N * 8, cudaMemcpyHostToDevice));
        GK(gpk_oz_slice(dA, K, M, K, dia, sA, nullptr));
        GK(gpk_oz_slice(dB, K, N, K, dib, sB, nullptr));
        CK(cudaDeviceSynchronize());
        std::vector<uint8_t> hA(bA), hB(bB);
        CK(cudaMemcpy(hA.data(), sA, bA, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(hB.data(), sB, bB, cudaMemcpyDeviceToHost));
        // 1. slicer
        std::vector<int8_t> DA((size_t)M * K * NS), DB((size_t)N * K * NS);
        long long bad = 0;
        for (int r = 0; r < M; ++r) for (int k = 0; k < K; ++k) {
            int8_t* d = &DA[((size_t)r * K + k) * NS];
            host_digits(p.A[(size_t)r * K + k], p.ia[r], d);
            for (int s = 0; s < NS; ++s) if ((int8_t)hA[rec_offset(r, k, s, kch)] != d[s]) { if (bad < 5) printf("  slice A mismatch r=%d k=%d s=%d gpu=%d host=%d\n", r, k, s, (int)(int8_t)hA[rec_offset(r, k, s, kch)], (int)d[s]); ++bad; }
        }
        for (int r = 0; r < N; ++r) for (int k = 0; k < K; ++k) {
            int8_t* d = &DB[((size_t)r * K + k) * NS];
            host_digits(p.B[(size_t)r * K + k], p.ib[r], d);
            for (int s = 0; s < NS; ++s) if ((int8_t)hB[rec_offset(r, k, s, kch)] != d[s]) ++bad;
        }
        printf("[slicer] digit bytes differing from the host slicer: %lld\n", bad);
        fails += bad != 0;
        // digit reconstruction error on the host
        double maxrec = 0;
        for (int r = 0; r < M; ++r) for (int k = 0; k < K; ++k) {
            long double u = 0; const int8_t* d = &DA[((size_t)r * K + k) * NS];
            for (int s = 0; s < NS; ++s) u += (long double)d[s] * ldexpl(1.0L, -7 - 8 * s);
            maxrec = std::fmax(maxrec, (double)fabsl(u * p.sa[r] - p.A[(size_t)r * K + k]) / p.sa[r]);
        }
        printf("[slicer] max |reconstruction - x| / sigma = %.3e (expect <= 2^-56 = 1.4e-17)\n", maxrec);
        // 2. GEMM: groups g = s + t < NG; host integer sums for all 13 groups, FP64 reference in long double
        const int NGMAX = 8;
        std::vector<long long> acc((size_t)M * N * NGMAX, 0);
        std::vector<double> Cref((size_t)M * N);
        for (int m = 0; m < M; ++m) for (int n = 0; n < N; ++n) {
            long double ref = 0;
            long long* a = &acc[((size_t)m * N + n) * NGMAX];
            for (int k = 0; k < K; ++k) {
                ref += (long double)p.A[(size_t)m * K + k] * (long double)p.B[(size_t)n * K + k];
                const int8_t* da = &DA[((size_t)m * K + k) * NS]; const int8_t* db = &DB[((size_t)n * K + k) * NS];
                for (int s = 0; s < NS; ++s) for (int t = 0; t < NS && s + t < NGMAX; ++t) a[s + t] += (long long)da[s] * db[t];
            }
            Cref[(size_t)m * N + n] = (double)ref;
        }
        cudaFree(dbg);
        CK(cudaMalloc(&dbg, (size_t)tiles * NGMAX * 128 * 64 * 4));
        for (int NG : {7, 8}) {
            CK(cudaMemset(dC, 0xFF, (size_t)M * N * 8)); CK(cudaMemset(dbg, 0xFF, (size_t)tiles * NGMAX * 128 * 64 * 4));
            GK(gpk_oz_gemm_tn(sA, sB, M, N, K, dsa, dsb, dC, N, dbg, NG, nullptr));
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("[gemm NG=%d] kernel failed: %s\n", NG, cudaGetErrorString(e)); fails++; break; }
            std::vector<int> hd((size_t)tiles * NG * 128 * 64);
            std::vector<double> hC((size_t)M * N);
            CK(cudaMemcpy(hd.data(), dbg, hd.size() * 4, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(hC.data(), dC, hC.size() * 8, cudaMemcpyDeviceToHost));
            long long ibad = 0; double maxerr = 0;
            for (int m = 0; m < M; ++m) for (int n = 0; n < N; ++n) {
                const int mb = m / 128, nb = n / 64, tile = mb * (N / 64) + nb;
                for (int g = 0; g < NG; ++g) {
                    const int got = hd[(((size_t)tile * NG + g) * 128 + (m % 128)) * 64 + (n % 64)];
                    if ((long long)got != acc[((size_t)m * N + n) * NGMAX + g]) { if (ibad < 6) printf("  acc mismatch m=%d n=%d g=%d gpu=%d host=%lld\n", m, n, g, got, acc[((size_t)m * N + n) * NGMAX + g]); ++ibad; }
                }
                const double scale = p.sa[m] * p.sb[n];
                maxerr = std::fmax(maxerr, std::fabs(hC[(size_t)m * N + n] - Cref[(size_t)m * N + n]) / scale);
            }
            printf("[gemm NG=%d] int32 accumulators differing: %lld of %d; max |C - ref| / (sigma_a sigma_b) = %.3e\n", NG, ibad, M * N * NG, maxerr);
            fails += !(ibad == 0 && maxerr < 1e-13);
        }
    }
    if (perf && fails == 0) {
        // many tiles, K = 2048: 2 M N K FP64-equivalent flop, 28 int8 MACs each
        const int M = 16384, N = 4096, K = 2048;
        uint8_t *sA, *sB; double* dC;
        const size_t bA = gpk_oz_slice_bytes(M, K), bB = gpk_oz_slice_bytes(N, K);
        CK(cudaMalloc(&sA, bA)); CK(cudaMalloc(&sB, bB)); CK(cudaMalloc(&dC, (size_t)M * N * 8));
        CK(cudaMemset(sA, 0x11, bA)); CK(cudaMemset(sB, 0x07, bB));
        cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
        const int reps = 5;
        float ms;
        for (int ng : {7, 8}) {
            for (int it = 0; it < 2; ++it) GK(gpk_oz_gemm_tn(sA, sB, M, N, K, nullptr, nullptr, dC, N, nullptr, ng, nullptr));
            CK(cudaDeviceSynchronize());
            CK(cudaEventRecord(e0));
            for (int it = 0; it < reps; ++it) GK(gpk_oz_gemm_tn(sA, sB, M, N, K, nullptr, nullptr, dC, N, nullptr, ng, nullptr));
            CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
            CK(cudaEventElapsedTime(&ms, e0, e1)); ms /= reps;
            const double flop = 2.0 * M * N * K;
            const int prods = ng == 7 ? 28 : 34;
            printf("[perf] groups %d (%d digit products) M=%d N=%d K=%d: %.3f ms per GEMM, %.1f TFLOP/s FP64-equivalent, %.0f int8 TOPS\n", ng, prods, M, N, K, ms, flop / ms * 1e-9, prods * flop / ms * 1e-9);
        }
        // K sweep (per-tile overheads)
        for (int k2 : {256, 1024, 4096}) {
            if ((size_t)gpk_oz_slice_bytes(M, k2) > bA * 2) continue;
            uint8_t *tA, *tB; CK(cudaMalloc(&tA, gpk_oz_slice_bytes(M, k2))); CK(cudaMalloc(&tB, gpk_oz_slice_bytes(N, k2)));
            CK(cudaMemset(tA, 0x11, gpk_oz_slice_bytes(M, k2))); CK(cudaMemset(tB, 0x07, gpk_oz_slice_bytes(N, k2)));
            GK(gpk_oz_gemm_tn(tA, tB, M, N, k2, nullptr, nullptr, dC, N, nullptr, 8, nullptr));
            CK(cudaDeviceSynchronize());
            CK(cudaEventRecord(e0));
            for (int it = 0; it < reps; ++it) GK(gpk_oz_gemm_tn(tA, tB, M, N, k2, nullptr, nullptr, dC, N, nullptr, 8, nullptr));
            CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
            CK(cudaEventElapsedTime(&ms, e0, e1)); ms /= reps;
            printf("[perf] K=%d: %.3f ms, %.1f TFLOP/s FP64-equivalent\n", k2, ms, 2.0 * M * N * k2 / ms * 1e-9);
            cudaFree(tA); cudaFree(tB);
        }
    }
    printf(fails ? "OZ TEST FAILED (%d)\n" : "OZ TEST OK\n", fails);
    return fails ? 1 : 0;
}
