// intpipe.cu -- issue-rate micro-benchmark for the integer work of a 64-bit NTT butterfly on B200.
// "the first artefact to produce on the box".  Build + run:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o intpipe intpipe.cu && ./intpipe
// Each kernel keeps ILP independent dependency chains per thread in registers; results are
// folded into one store so nothing is optimised away.  Reported: lane-ops per clock per SM
// (using the SM clock measured inside the kernel with clock64 / globaltimer).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef unsigned long long u64;
typedef unsigned int u32;

constexpr int ILP = 8;
constexpr int ITERS = 4096;

__device__ __forceinline__ u64 csub(u64 x, u64 m) { u64 t = x - m; return t < x ? t : x; }

enum Op { IMAD32, IMADWIDE, MULHI64, MULLO64, SHOUP_LAZY, CT_BFLY, CT_BFLY_LAZY4, DFMA, DFMA_PLUS_IMAD, IADD64, LOP3, GS_BFLY,
          SOLINAS60, NOPS };
const char* names[] = {"mad.lo.u32", "mad.wide.u32", "mul.hi.u64", "mul.lo.u64", "shoup_lazy modmul", "CT butterfly (csub each)",
                       "CT butterfly (lazy, no csub)", "fma.f64", "fma.f64 + mad.lo.u32 (pairs)", "add.u64", "lop3", "GS butterfly",
                       "solinas 2^60-d modmul"};

template <int OP>
__global__ void __launch_bounds__(256) k(u64* out, u64 q, u64 w, u64 ws, long long* clk) {
    u64 x[ILP], y[ILP];
    double d[ILP];
    u32 a[ILP];
#pragma unroll
    for (int i = 0; i < ILP; ++i) {
        x[i] = (threadIdx.x + 1) * 0x9E3779B97F4A7C15ull + i + blockIdx.x;
        y[i] = x[i] * 3 + 1; x[i] %= q; y[i] %= q;
        d[i] = (double)x[i] * 1e-18; a[i] = (u32)x[i];
    }
    const u64 q2 = 2 * q;
    const u32 dd = (u32)(0 - (u32)q);   // for q = 2^60 - d, d < 2^32
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) {
            if (OP == IMAD32) a[i] = a[i] * (u32)w + (u32)ws;
            if (OP == IMADWIDE) x[i] = (u64)(u32)x[i] * (u32)w + x[i];
            if (OP == MULHI64) x[i] = __umul64hi(x[i], ws) + w;
            if (OP == MULLO64) x[i] = x[i] * w + ws;
            if (OP == SHOUP_LAZY) { u64 hi = __umul64hi(x[i], ws); x[i] = x[i] * w - hi * q; }
            if (OP == CT_BFLY) {
                u64 xx = csub(x[i], q2);
                u64 hi = __umul64hi(y[i], ws); u64 t = y[i] * w - hi * q;
                y[i] = xx + q2 - t; x[i] = xx + t;
            }
            if (OP == CT_BFLY_LAZY4) {
                u64 hi = __umul64hi(y[i], ws); u64 t = y[i] * w - hi * q;
                u64 xx = x[i];
                y[i] = xx + q2 - t; x[i] = xx + t;
                if ((it & 3) == 3) x[i] = csub(x[i], 8 * q);
            }
            if (OP == GS_BFLY) {
                u64 t = x[i] + q2 - y[i];
                x[i] = csub(x[i] + y[i], q2);
                u64 hi = __umul64hi(t, ws); y[i] = t * w - hi * q;
            }
            if (OP == DFMA) d[i] = fma(d[i], 1.0000001, 1e-9);
            if (OP == DFMA_PLUS_IMAD) { d[i] = fma(d[i], 1.0000001, 1e-9); a[i] = a[i] * (u32)w + (u32)ws; }
            if (OP == IADD64) x[i] = x[i] + y[i];
            if (OP == LOP3) a[i] = (a[i] ^ (u32)w) & ((u32)ws | a[(i + 1) % ILP]);
            if (OP == SOLINAS60) {
                // x*w mod (2^60 - dd): 128-bit product, fold twice with 2^60 = dd
                u64 lo = x[i] * w, hi = __umul64hi(x[i], w);
                u64 ph = (hi << 4) | (lo >> 60), pl = lo & ((1ull << 60) - 1);    // ph < 2^60
                u64 m_lo = (u64)(u32)ph * dd, m_hi = (ph >> 32) * dd;              // ph*dd = m_lo + m_hi<<32
                u64 s = pl + m_lo; u64 c = s < pl;
                u64 hi2 = (m_hi >> 28) + (s >> 60) + (c << 4);                      // approx fold
                u64 lo2 = (s & ((1ull << 60) - 1)) + ((m_hi << 32) & ((1ull << 60) - 1));
                x[i] = lo2 + hi2 * dd;
            }
        }
    }
    long long t1 = clock64();
    u64 acc = 0;
#pragma unroll
    for (int i = 0; i < ILP; ++i) acc += x[i] + y[i] + a[i] + (u64)d[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0 && blockIdx.x == 0) clk[0] = t1 - t0;
}

template <int OP>
void run(u64* out, long long* clk, int sms, int bps) {
    const u64 q = 1152921504606584833ull, w = 719028594519ull;
    const u64 ws = (u64)(((unsigned __int128)w << 64) / q);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<OP><<<sms * bps, 256>>>(out, q, w, ws, clk);
    cudaEventRecord(e0);
    k<OP><<<sms * bps, 256>>>(out, q, w, ws, clk);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long c; cudaMemcpy(&c, clk, 8, cudaMemcpyDeviceToHost);
    // one block's clock count ~ kernel duration / waves; derive effective MHz from it
    double lane_ops = (double)sms * bps * 256 * ITERS * ILP;
    double per_s = lane_ops / (ms * 1e-3);
    // cycles: the measured block ran concurrently with (bps-1) others on its SM
    double per_clk_sm = (double)bps * 256 * ITERS * ILP / (double)c;
    printf("%-32s %8.3f ms  %8.2f Gop/s/SM  %7.2f lane-ops/clk/SM (block clk %lld)  %7.1f Top/s chip\n", names[OP], ms,
           per_s / sms / 1e9, per_clk_sm, c, per_s / 1e12);
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int sms = p.multiProcessorCount, bps = 8;     // 8 blocks x 256 thr = 2048 threads/SM, one wave
    printf("%s, %d SMs, clockRate %d kHz\n", p.name, sms, p.clockRate);
    u64* out; long long* clk;
    cudaMalloc(&out, (size_t)sms * bps * 256 * 8); cudaMalloc(&clk, 8);
    run<IMAD32>(out, clk, sms, bps);
    run<IMADWIDE>(out, clk, sms, bps);
    run<MULHI64>(out, clk, sms, bps);
    run<MULLO64>(out, clk, sms, bps);
    run<SHOUP_LAZY>(out, clk, sms, bps);
    run<CT_BFLY>(out, clk, sms, bps);
    run<CT_BFLY_LAZY4>(out, clk, sms, bps);
    run<GS_BFLY>(out, clk, sms, bps);
    run<SOLINAS60>(out, clk, sms, bps);
    run<DFMA>(out, clk, sms, bps);
    run<DFMA_PLUS_IMAD>(out, clk, sms, bps);
    run<IADD64>(out, clk, sms, bps);
    run<LOP3>(out, clk, sms, bps);
    printf("err=%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
