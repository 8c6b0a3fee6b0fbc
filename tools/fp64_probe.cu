// FP64 issue-rate probe for B200 (sm_100a): DFMA vs DMMA shapes. Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3
#include <cstdio>
#include <cuda_runtime.h>

#define CK(x) do{cudaError_t e_=(x); if(e_!=cudaSuccess){printf("CUDA error %s at %d\n", cudaGetErrorString(e_), __LINE__); return 1;}}while(0)

template<int ILP>
__global__ void k_dfma(double* out, int iters, double a, double b) {
    double acc[ILP];
#pragma unroll
    for (int i = 0; i < ILP; i++) acc[i] = threadIdx.x + i;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < ILP; i++) acc[i] = fma(acc[i], a, b);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template<int ILP>
__global__ void k_dmma884(double* out, int iters, double a, double b) {
    double c0[ILP], c1[ILP];
#pragma unroll
    for (int i = 0; i < ILP; i++) { c0[i] = threadIdx.x + i; c1[i] = i; }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < ILP; i++) {
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                         : "+d"(c0[i]), "+d"(c1[i]) : "d"(a), "d"(b));
        }
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) s += c0[i] + c1[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template<int ILP>
__global__ void k_dmma16816(double* out, int iters, double a, double b) {
    double c[ILP][4];
#pragma unroll
    for (int i = 0; i < ILP; i++) { c[i][0] = threadIdx.x + i; c[i][1] = i; c[i][2] = 1; c[i][3] = 2; }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < ILP; i++) {
            asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};\n"
                         : "+d"(c[i][0]), "+d"(c[i][1]), "+d"(c[i][2]), "+d"(c[i][3])
                         : "d"(a), "d"(b), "d"(a), "d"(b), "d"(a), "d"(b), "d"(a), "d"(b), "d"(b), "d"(a), "d"(b), "d"(a));
        }
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template<int ILP>
__global__ void k_dmma1688(double* out, int iters, double a, double b) {
    double c[ILP][4];
#pragma unroll
    for (int i = 0; i < ILP; i++) { c[i][0] = threadIdx.x + i; c[i][1] = i; c[i][2] = 1; c[i][3] = 2; }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < ILP; i++) {
            asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                         : "+d"(c[i][0]), "+d"(c[i][1]), "+d"(c[i][2]), "+d"(c[i][3])
                         : "d"(a), "d"(b), "d"(a), "d"(b), "d"(b), "d"(a));
        }
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template<typename F>
float timeit(F f) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    f(); cudaDeviceSynchronize();
    cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); return ms;
}

int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    printf("device %s sms=%d clock=%d kHz\n", p.name, p.multiProcessorCount, p.clockRate);
    int sms = p.multiProcessorCount;
    double* out; CK(cudaMalloc(&out, sizeof(double) * sms * 8 * 1024));
    const int iters = 20000;
    for (int bps = 1; bps <= 2; bps++)
    for (int threads = 128; threads <= 1024; threads *= 2) {
        int grid = sms * bps;
        double warps = (double)grid * threads / 32;
        {
            float ms = timeit([&]{ k_dfma<8><<<grid, threads>>>(out, iters, 1.0000001, 1e-9); });
            double fl = warps * 32 * 8.0 * iters * 2;
            printf("dfma   ilp8  bps=%d thr=%4d: %8.3f ms  %7.2f TFLOP/s\n", bps, threads, ms, fl / ms * 1e-9);
        }
        {
            float ms = timeit([&]{ k_dmma884<8><<<grid, threads>>>(out, iters, 1.0000001, 1e-9); });
            double fl = warps * 8.0 * iters * (8 * 8 * 4 * 2);
            printf("dmma884 ilp8 bps=%d thr=%4d: %8.3f ms  %7.2f TFLOP/s\n", bps, threads, ms, fl / ms * 1e-9);
        }
        {
            float ms = timeit([&]{ k_dmma1688<4><<<grid, threads>>>(out, iters, 1.0000001, 1e-9); });
            double fl = warps * 4.0 * iters * (16 * 8 * 8 * 2);
            printf("dmma1688 ilp4 bps=%d thr=%4d: %8.3f ms  %7.2f TFLOP/s\n", bps, threads, ms, fl / ms * 1e-9);
        }
        {
            float ms = timeit([&]{ k_dmma16816<4><<<grid, threads>>>(out, iters, 1.0000001, 1e-9); });
            double fl = warps * 4.0 * iters * (16 * 8 * 16 * 2);
            printf("dmma16816 ilp4 bps=%d thr=%4d: %8.3f ms  %7.2f TFLOP/s\n", bps, threads, ms, fl / ms * 1e-9);
        }
    }
    // dependent-chain latency (1 warp)
    {
        float ms = timeit([&]{ k_dfma<1><<<1, 32>>>(out, iters, 1.0000001, 1e-9); });
        printf("dfma latency   ~ %.1f ns/op\n", ms * 1e6 / iters);
        ms = timeit([&]{ k_dmma884<1><<<1, 32>>>(out, iters, 1.0000001, 1e-9); });
        printf("dmma884 latency ~ %.1f ns/op\n", ms * 1e6 / iters);
        ms = timeit([&]{ k_dmma16816<1><<<1, 32>>>(out, iters, 1.0000001, 1e-9); });
        printf("dmma16816 latency ~ %.1f ns/op\n", ms * 1e6 / iters);
    }
    return 0;
}
