// dmma.cu -- FP64 rate on B200: plain DFMA chains against mma.sync m8n8k4 / m16n8k8 .f64 (DMMA).
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o dmma dmma.cu && ./dmma
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void __launch_bounds__(256) k(double* out, double x, int iters) {
    double acc[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = threadIdx.x * 1e-9 + i;
    double a = x + threadIdx.x * 1e-12, b = x * 0.5;
    for (int it = 0; it < iters; ++it) {
        if (MODE == 0) {            // 16 independent DFMA per thread
#pragma unroll
            for (int i = 0; i < 16; ++i) acc[i] = fma(a, b, acc[i]);
        } else if (MODE == 1) {     // 8 independent m8n8k4 (2 accumulators each): 256 FMA per warp instruction
#pragma unroll
            for (int i = 0; i < 8; ++i)
                asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                             : "+d"(acc[2 * i]), "+d"(acc[2 * i + 1]) : "d"(a), "d"(b));
        } else {                    // 4 independent m16n8k8 (4 accumulators, 4 A, 2 B): 1024 FMA per warp instruction
#pragma unroll
            for (int i = 0; i < 4; ++i)
                asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+d"(acc[4 * i]), "+d"(acc[4 * i + 1]), "+d"(acc[4 * i + 2]), "+d"(acc[4 * i + 3])
                             : "d"(a), "d"(b), "d"(a), "d"(b), "d"(b), "d"(a));
        }
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE>
void run(const char* name, double fma_per_thread_iter, int sms) {
    double* out;
    const int blocks = sms * 4, iters = 20000;
    cudaMalloc(&out, blocks * 256 * sizeof(double));
    k<MODE><<<blocks, 256>>>(out, 1.0000001, 10);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k<MODE><<<blocks, 256>>>(out, 1.0000001, iters);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double fma = (double)blocks * 256 * iters * fma_per_thread_iter;
    printf("%-28s %8.3f ms  %7.2f TFMA/s chip = %6.2f TFLOP/s  (%s)\n", name, ms, fma / ms * 1e-9, 2 * fma / ms * 1e-9,
           cudaGetErrorString(cudaGetLastError()));
    cudaFree(out);
}

int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    printf("%s, %d SMs\n", p.name, p.multiProcessorCount);
    run<0>("DFMA x16", 16, p.multiProcessorCount);
    run<1>("DMMA m8n8k4 x8", 8 * 256 / 32.0, p.multiProcessorCount);
    run<2>("DMMA m16n8k8 x4", 4 * 1024 / 32.0, p.multiProcessorCount);
    return 0;
}
