// tcgen05.mma.kind::i8 issue-rate probe (B200): clocks per MMA as a function of the instruction N, with the operands
// resident in shared memory (no global traffic), one CTA per SM.  Answers: is the 128 x 64 x 32 shape of the digit
// GEMM bound by the tensor pipe or by the shared-memory operand reads?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tools/umma_probe tools/umma_probe.cu
#include <cuda_runtime.h>
#include <unistd.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>

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

// mode 0: every MMA reads a different A slice and B slice (like the digit GEMM); mode 1: same A for all MMAs of a batch
template <int N, int MODE>
__global__ void __launch_bounds__(128, 1) k_probe(int batches, long long* clocks_out) {
    constexpr int per_batch = 28;
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ uint64_t bar[4];
    __shared__ uint32_t tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < 160 * 1024 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0x01010101u * (i & 3);
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
        const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((128u >> 4) << 24);
        // K-major no-swizzle, slice-major layout: [slice][row group][2][8 x 16 B]: LBO = 128, SBO = 256
        const uint64_t sdesc = ((uint64_t)(128 >> 4) << 16) | ((uint64_t)(256 >> 4) << 32) | (1ull << 46);
        const uint32_t a0 = smem_u32(smem), b0 = a0 + 32 * 1024;     // A: 8 slices x 4 KB; B: up to 128 KB
        constexpr int nacc = 512 / N;
        const uint64_t da = sdesc | ((a0 >> 4) & 0x3FFF), db = sdesc | ((b0 >> 4) & 0x3FFF);
        const long long t0 = clock64();
        for (int it = 0; it < batches; ++it) {
#pragma unroll
            for (int j = 0; j < per_batch; ++j) {
                const uint32_t ao = (MODE == 1) ? 0 : (j & 7) * (4096 >> 4);
                const uint32_t bo = (j % 4) * ((N * 32) >> 4);
                umma_i8(tmem + (j % nacc) * N, da + ao, db + bo, idesc, (it > 0 || j >= nacc) ? 1u : 0u);
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(smem_u32(&bar[it & 3])) : "memory");
            if (it >= 2) mbar_wait(smem_u32(&bar[(it - 2) & 3]), ((it - 2) >> 2) & 1);   // at most 3 batches in flight
        }
        for (int it = (batches >= 2 ? batches - 2 : 0); it < batches; ++it) mbar_wait(smem_u32(&bar[it & 3]), (it >> 2) & 1);
        clocks_out[blockIdx.x] = clock64() - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tmem), "r"(512u) : "memory");
}

int main() {
    alarm(120);
    const int smem_bytes = 160 * 1024;
    long long* d_clk;
    CK(cudaMalloc(&d_clk, 148 * 8));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    const int batches = 400, per_batch = 28;
    typedef void (*kern_t)(int, long long*);
    kern_t kerns[2][3] = {{k_probe<64, 0>, k_probe<128, 0>, k_probe<256, 0>}, {k_probe<64, 1>, k_probe<128, 1>, k_probe<256, 1>}};
    for (int m = 0; m < 2; ++m) for (int n = 0; n < 3; ++n) CK(cudaFuncSetAttribute(kerns[m][n], cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    for (int grid : {1, 148}) {
        for (int mode = 0; mode < 2; ++mode) {
            for (int ni = 0; ni < 3; ++ni) {
                const int N = 64 << ni;
                kerns[mode][ni]<<<grid, 128, smem_bytes>>>(20, d_clk);      // warm-up
                CK(cudaDeviceSynchronize());
                CK(cudaEventRecord(e0));
                kerns[mode][ni]<<<grid, 128, smem_bytes>>>(batches, d_clk);
                CK(cudaEventRecord(e1));
                CK(cudaEventSynchronize(e1));
                float ms;
                CK(cudaEventElapsedTime(&ms, e0, e1));
                long long h[148];
                CK(cudaMemcpy(h, d_clk, grid * 8, cudaMemcpyDeviceToHost));
                double avg = 0;
                for (int i = 0; i < grid; ++i) avg += (double)h[i];
                avg /= grid;
                const double n_mma = (double)batches * per_batch;
                const double macs = n_mma * 128.0 * N * 32.0 * grid;
                printf("grid %3d mode %d N %3d: %.1f clk / MMA (ideal %d), kernel %.3f ms, %.0f int8 TOPS, SM clock ~%.0f MHz\n", grid, mode, N,
                       avg / n_mma, N / 2, ms, 2 * macs / ms * 1e-9, avg / ms * 1e-3);
            }
        }
    }
    return 0;
}
