// fp64bfly.cu -- can the FP64 pipe carry NTT butterflies for moduli below 2^50?  ("a narrower
// quotient path could move the integer ceiling".)  Register-only butterfly loops, time-based rates per SM:
//   int   : Cooley-Tukey butterfly with the approximate Shoup quotient (what ntt_flat.cuh runs, fmaheavy pipe)
//   fp    : the same butterfly on doubles: t = y*w - rint(y*w/q)*q exactly with 2 products + 3 FMAs (6 FP64-pipe ops),
//           x +- t (2 ops), x folded every 4th stage (3 ops)
//   mixed : both kinds of chains in one thread (the two pipes issue side by side)
// Build + run:  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64bfly fp64bfly.cu && ./fp64bfly
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../matrix_fhe_gpu_b200/csrc/modmath.cuh"
using mfhe::u64; using mfhe::u32; using mfhe::csub;

constexpr int ITERS = 4096;
constexpr double kMagic = 6755399441055744.0;
__device__ __forceinline__ double fp_mulmod(double y, double w, double wq, double q) {
    const double h = __dmul_rn(y, w), l = __fma_rn(y, w, -h);
    const double k = __dadd_rn(__fma_rn(y, wq, kMagic), -kMagic);
    return __dadd_rn(__fma_rn(-k, q, h), l);
}
__device__ __forceinline__ double fp_fold(double x, double q, double qinv) {
    const double k = __dadd_rn(__fma_rn(x, qinv, kMagic), -kMagic);
    return __fma_rn(-k, q, x);
}

template <int NI, int NF>
__global__ void __launch_bounds__(256) k(u64* out, u64 q, u64 w, u64 ws, double qd, double wd, double wqd) {
    u64 x[NI ? NI : 1], y[NI ? NI : 1];
    double a[NF ? NF : 1], b[NF ? NF : 1];
#pragma unroll
    for (int i = 0; i < NI; ++i) { x[i] = ((threadIdx.x + 1) * 0x9E3779B97F4A7C15ull + i + blockIdx.x) % q; y[i] = (x[i] * 3 + 1) % q; }
#pragma unroll
    for (int i = 0; i < NF; ++i) { a[i] = (double)(((threadIdx.x + 7) * 0x9E3779B97F4A7C15ull + i + blockIdx.x) % q); b[i] = a[i] * 0.5 + 1.0; b[i] = floor(b[i]); }
    const u64 q2 = 2 * q;
    const double qinv = 1.0 / qd;
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < (NI > NF ? NI : NF); ++i) {
            if (i < NI) {
                mfhe::tw_t tw = {w, ws};
                u64 t = mfhe::mul_shoup_lazy4(y[i], tw, 0 - q);
                u64 xx = x[i];
                y[i] = xx + 2 * q2 - t; x[i] = xx + t;
                if ((it & 1) == 1) x[i] = csub(x[i], 8 * q);
            }
            if (i < NF) {
                const double t = fp_mulmod(b[i], wd, wqd, qd);
                const double xx = a[i];
                b[i] = __dadd_rn(xx, -t); a[i] = __dadd_rn(xx, t);
                if ((it & 3) == 3) a[i] = fp_fold(a[i], qd, qinv);
            }
        }
    }
    u64 acc = 0;
#pragma unroll
    for (int i = 0; i < NI; ++i) acc += x[i] + y[i];
#pragma unroll
    for (int i = 0; i < NF; ++i) acc += (u64)(long long)a[i] + (u64)(long long)b[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int NI, int NF>
void run(u64* out, int sms, int bps, const char* name) {
    const u64 q = 1125899903827969ull, w = 719028594519ull;       // 50-bit NTT prime
    const u64 ws = (u64)(((unsigned __int128)w << 64) / q);
    const double qd = (double)q, wd = (double)w, wqd = wd / qd;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<NI, NF><<<sms * bps, 256>>>(out, q, w, ws, qd, wd, wqd);
    cudaEventRecord(e0);
    for (int r = 0; r < 5; ++r) k<NI, NF><<<sms * bps, 256>>>(out, q, w, ws, qd, wd, wqd);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= 5;
    const double bf = (double)sms * bps * 256 * ITERS * (NI + NF);
    printf("%-34s %8.3f ms  %7.2f G butterflies/s/SM  %6.2f T/s chip   (int %d : fp %d chains per thread)\n", name, ms,
           bf / (ms * 1e-3) / sms / 1e9, bf / (ms * 1e-3) / 1e12, NI, NF);
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount;
    printf("%s, %d SMs\n", p.name, sms);
    u64* out; cudaMalloc(&out, (size_t)sms * 8 * 256 * 8);
    for (int bps : {4, 8}) {
        printf("-- %d blocks of 256 threads per SM\n", bps);
        run<8, 0>(out, sms, bps, "int only (8 chains)");
        run<0, 8>(out, sms, bps, "fp64 only (8 chains)");
        run<0, 4>(out, sms, bps, "fp64 only (4 chains)");
        run<4, 4>(out, sms, bps, "mixed 1:1 (4 + 4)");
        run<3, 6>(out, sms, bps, "mixed 1:2 (3 + 6)");
        run<2, 6>(out, sms, bps, "mixed 1:3 (2 + 6)");
        run<4, 2>(out, sms, bps, "mixed 2:1 (4 + 2)");
    }
    printf("err=%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
