// Data-dependent power of tcgen05.mma.kind::i8 (B200): sustained int8 throughput of an SMEM-resident MMA loop
// (no global traffic, one CTA per SM, N = 256) as a function of the operand bit patterns and of the signed /
// unsigned operand format.  Each configuration runs ~3 s so that the board reaches its power cap; clock and power
// are read from nvidia-smi in the middle of the run.  Question answered: would another digit representation of the
// emulated-FP64 GEMM (non-negative digits, 7-bit digits, unsigned format) draw less power per multiply-add?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tools/umma_power tools/umma_power.cu
#include <cuda_runtime.h>
#include <unistd.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(2); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    const long long t0 = clock64();
    while (!mbar_try_wait(bar, parity)) if (clock64() - t0 > 4000000000LL) __trap();
}
__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n}\n" ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
}

constexpr int N = 256;
constexpr int SMEM_BYTES = 160 * 1024;       // A: 8 slices x 4 KB at 0; B: 4 slices x 8 KB at 32 KB

// fmt: 1 = signed 8 bit operands, 0 = unsigned.  reset_every: batches between accumulator resets (keeps int32 sums bounded
// the way a K = 4096 contraction does: 128 MMAs of K = 32 per accumulator)
__global__ void __launch_bounds__(128, 1) k_power(const uint8_t* __restrict__ pattern, int batches, uint32_t fmt, long long* clocks_out) {
    constexpr int per_batch = 28;
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ uint64_t bar[4];
    __shared__ uint32_t tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < 64 * 1024 / 16; i += 128) reinterpret_cast<uint4*>(smem)[i] = reinterpret_cast<const uint4*>(pattern)[i];
    if (tid == 0) {
        for (int i = 0; i < 4; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" ::"r"(smem_u32(&bar[i])));
        asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(smem_u32(&tmem_slot)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
    const uint32_t tmem = tmem_slot;
    if (tid == 0) {
        const uint32_t idesc = (2u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(N >> 3) << 17) | ((128u >> 4) << 24);
        const uint64_t sdesc = ((uint64_t)(128 >> 4) << 16) | ((uint64_t)(256 >> 4) << 32) | (1ull << 46);
        const uint32_t a0 = smem_u32(smem), b0 = a0 + 32 * 1024;
        constexpr int nacc = 512 / N;
        const uint64_t da = sdesc | ((a0 >> 4) & 0x3FFF), db = sdesc | ((b0 >> 4) & 0x3FFF);
        const long long t0 = clock64();
        for (int it = 0; it < batches; ++it) {
            const bool reset = (it % 9) == 0;                     // 9 x 28 / 2 = 126 accumulations per accumulator
#pragma unroll
            for (int j = 0; j < per_batch; ++j) {
                const uint32_t ao = (j & 7) * (4096 >> 4);
                const uint32_t bo = (j % 4) * ((N * 32) >> 4);
                umma_i8(tmem + (j % nacc) * N, da + ao, db + bo, idesc, (reset && j < nacc) ? 0u : 1u);
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(smem_u32(&bar[it & 3])) : "memory");
            if (it >= 2) mbar_wait(smem_u32(&bar[(it - 2) & 3]), ((it - 2) >> 2) & 1);
        }
        for (int it = (batches >= 2 ? batches - 2 : 0); it < batches; ++it) mbar_wait(smem_u32(&bar[it & 3]), (it >> 2) & 1);
        clocks_out[blockIdx.x] = clock64() - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tmem), "r"(512u) : "memory");
}

struct Pattern { const char* name; int lo_a, hi_a, lo_b, hi_b; uint32_t fmt; int zero_pct; };

int main(int argc, char** argv) {
    alarm(280);
    const double seconds = argc > 1 ? atof(argv[1]) : 3.0;
    CK(cudaFuncSetAttribute(k_power, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    long long* d_clk; uint8_t* d_pat;
    CK(cudaMalloc(&d_clk, 148 * 8)); CK(cudaMalloc(&d_pat, 64 * 1024));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    const Pattern pats[] = {
        {"zeros                      ", 0, 0, 0, 0, 1, 0},
        {"constant 0x11 x 0x07       ", 0x11, 0x11, 0x07, 0x07, 1, 0},
        {"s8 uniform [-128,127]      ", -128, 127, -128, 127, 1, 0},
        {"s8 uniform [0,127]         ", 0, 127, 0, 127, 1, 0},
        {"u8 uniform [0,255] (u fmt) ", 0, 255, 0, 255, 0, 0},
        {"u8 uniform [0,127] (u fmt) ", 0, 127, 0, 127, 0, 0},
        {"s8 uniform [-64,63]        ", -64, 63, -64, 63, 1, 0},
        {"s8 uniform [0,63]          ", 0, 63, 0, 63, 1, 0},
        {"s8 uniform [-16,15]        ", -16, 15, -16, 15, 1, 0},
        {"A s8 full, B [0,127]       ", -128, 127, 0, 127, 1, 0},
        {"s8 full, 50% zero elements ", -128, 127, -128, 127, 1, 50},
        {"s8 full, 2 of 7 A slices 0 ", -128, 127, -128, 127, 1, -2},
    };
    const int per_batch = 28;
    const double clk_per_batch = per_batch * 128.0, hz = 1.5e9;
    const int batches = (int)(seconds * hz / clk_per_batch);
    std::vector<uint8_t> h(64 * 1024);
    for (const Pattern& p : pats) {
        std::mt19937 rng(12345);
        for (size_t i = 0; i < h.size(); ++i) {
            const bool isA = i < 32 * 1024;
            const int lo = isA ? p.lo_a : p.lo_b, hi = isA ? p.hi_a : p.hi_b;
            int v = lo + (int)(rng() % (unsigned)(hi - lo + 1));
            if (p.zero_pct > 0 && (int)(rng() % 100) < p.zero_pct) v = 0;
            if (p.zero_pct == -2 && isA && (i / 4096) < 2) v = 0;
            h[i] = (uint8_t)v;
        }
        CK(cudaMemcpy(d_pat, h.data(), h.size(), cudaMemcpyHostToDevice));
        k_power<<<148, 128, SMEM_BYTES>>>(d_pat, 2000, p.fmt, d_clk);
        CK(cudaDeviceSynchronize());
        CK(cudaEventRecord(e0));
        k_power<<<148, 128, SMEM_BYTES>>>(d_pat, batches, p.fmt, d_clk);
        CK(cudaEventRecord(e1));
        usleep((useconds_t)(seconds * 0.45e6));
        char smi[256] = "";
        if (FILE* f = popen("nvidia-smi --query-gpu=clocks.sm,power.draw --format=csv,noheader,nounits -i 0 2>/dev/null", "r")) {
            if (!fgets(smi, sizeof smi, f)) smi[0] = 0;
            pclose(f);
            if (char* nl = strchr(smi, '\n')) *nl = 0;
        }
        CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        long long hc[148]; CK(cudaMemcpy(hc, d_clk, sizeof hc, cudaMemcpyDeviceToHost));
        double avg = 0; for (int i = 0; i < 148; ++i) avg += (double)hc[i]; avg /= 148;
        const double n_mma = (double)batches * per_batch, macs = n_mma * 128.0 * N * 32.0 * 148;
        printf("%s: %7.0f int8 TOPS over %.2f s, %.1f clk/MMA, avg SM clock %.0f MHz, nvidia-smi mid-run [MHz, W]: %s\n", p.name,
               2 * macs / ms * 1e-9, ms * 1e-3, avg / n_mma, avg / ms * 1e-3, smi);
        fflush(stdout);
        usleep(300000);
    }
    return 0;
}
