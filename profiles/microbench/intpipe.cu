// intpipe.cu -- issue-rate micro-benchmark for the integer work of a 64-bit NTT butterfly on B200.
// "the first artefact to produce on the box".  Build + run:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -DMFHE_MODMUL=0 -o intpipe intpipe.cu && ./intpipe
// (-DMFHE_MODMUL=0: the FP64-assisted quotient variants measured here need the prepared-twiddle form of modmath.cuh)
// Each kernel keeps ILP independent dependency chains per thread in registers; results are
// folded into one store so nothing is optimised away.  Reported: lane-ops per clock per SM
// (using the SM clock measured inside the kernel with clock64 / globaltimer).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../matrix_fhe_gpu_b200/csrc/modmath.cuh"
using mfhe::u64; using mfhe::u32;

constexpr int ILP = 8;
constexpr int ITERS = 4096;

using mfhe::csub;


__device__ __forceinline__ u64 mul_fpq_cvt(u64 x, const mfhe::twf_t& w, u64 nq, u64 cneg) {
    const u32 x0 = (u32)x, x1 = (u32)(x >> 32);
    const double c = 29014219670751100192948224.0 - 4294967296.0;
    const double d = fma(__uint2double_rn(x1), w.s0d, fma(__uint2double_rn(x0), w.s1d, c));
    const u32 d0 = (u32)__double2loint(d), d1 = (u32)__double2hiint(d);
    const u32 w0 = (u32)w.w, w1 = (u32)(w.w >> 32), n0 = (u32)nq, n1 = (u32)(nq >> 32);
    const u32 c0 = (u32)cneg, c1 = (u32)(cneg >> 32);
    u32 ql, qh, t0, t1;
    asm("mad.lo.cc.u32 %0, %2, %3, %4;\n\tmadc.hi.u32 %1, %2, %3, %5;" : "=r"(ql), "=r"(qh) : "r"(x1), "r"(w.s1), "r"(d0), "r"(d1));
    asm("mad.lo.cc.u32 %0, %2, %3, %10;\n\tmadc.hi.u32 %1, %2, %3, %11;\n\tmad.lo.u32 %1, %2, %4, %1;\n\tmad.lo.u32 %1, %5, %3, %1;\n\t"
        "mad.lo.cc.u32 %0, %6, %7, %0;\n\tmadc.hi.u32 %1, %6, %7, %1;\n\tmad.lo.u32 %1, %6, %8, %1;\n\tmad.lo.u32 %1, %9, %7, %1;"
        : "=&r"(t0), "=&r"(t1) : "r"(x0), "r"(w0), "r"(w1), "r"(x1), "r"(ql), "r"(n0), "r"(n1), "r"(qh), "r"(c0), "r"(c1));
    return ((u64)t1 << 32) | t0;
}
__device__ __forceinline__ u64 mul_fpq_prmt(u64 x, const mfhe::twf_t& w, u64 nq, u64 cneg, u32 magic) {
    const u32 x0 = (u32)x, x1 = (u32)(x >> 32);
    u32 a0, a1;
    asm("prmt.b32 %0, %1, %2, 0x3210;" : "=r"(a0) : "r"(x0), "r"(magic));
    asm("prmt.b32 %0, %1, %2, 0x3210;" : "=r"(a1) : "r"(x1), "r"(magic));
    const double m0 = __hiloint2double(magic, (int)a0), m1 = __hiloint2double(magic, (int)a1);
    const double d = fma(m1, w.s0d, fma(m0, w.s1d, w.k));
    const u32 d0 = (u32)__double2loint(d), d1 = (u32)__double2hiint(d);
    const u32 w0 = (u32)w.w, w1 = (u32)(w.w >> 32), n0 = (u32)nq, n1 = (u32)(nq >> 32);
    const u32 c0 = (u32)cneg, c1 = (u32)(cneg >> 32);
    u32 ql, qh, t0, t1;
    asm("mad.lo.cc.u32 %0, %2, %3, %4;\n\tmadc.hi.u32 %1, %2, %3, %5;" : "=r"(ql), "=r"(qh) : "r"(x1), "r"(w.s1), "r"(d0), "r"(d1));
    asm("mad.lo.cc.u32 %0, %2, %3, %10;\n\tmadc.hi.u32 %1, %2, %3, %11;\n\tmad.lo.u32 %1, %2, %4, %1;\n\tmad.lo.u32 %1, %5, %3, %1;\n\t"
        "mad.lo.cc.u32 %0, %6, %7, %0;\n\tmadc.hi.u32 %1, %6, %7, %1;\n\tmad.lo.u32 %1, %6, %8, %1;\n\tmad.lo.u32 %1, %9, %7, %1;"
        : "=&r"(t0), "=&r"(t1) : "r"(x0), "r"(w0), "r"(w1), "r"(x1), "r"(ql), "r"(n0), "r"(n1), "r"(qh), "r"(c0), "r"(c1));
    return ((u64)t1 << 32) | t0;
}
__device__ __forceinline__ u64 csub_pred(u64 x, u64 m) {
    asm("{\n\t.reg .pred p;\n\tsetp.ge.u64 p, %0, %1;\n\t@p sub.u64 %0, %0, %1;\n\t}" : "+l"(x) : "l"(m));
    return x;
}
__device__ __forceinline__ u64 csub_sign(u64 x, u64 m) {   // subtract m when bit 63 of x is set
    asm("{\n\t.reg .pred p;\n\t.reg .b32 lo, hi;\n\tmov.b64 {lo, hi}, %0;\n\tsetp.lt.s32 p, hi, 0;\n\t@p sub.u64 %0, %0, %1;\n\t}" : "+l"(x) : "l"(m));
    return x;
}

enum Op { IMAD32, IMADWIDE, MULHI64, MULLO64, SHOUP_LAZY, CT_BFLY, CT_BFLY_LAZY4, DFMA, DFMA_PLUS_IMAD, IADD64, LOP3, GS_BFLY,
          SOLINAS60, CT_APPROX, GS_APPROX, IMADHI32, CT_FPQ, GS_FPQ, DFMA_PLUS_WIDE, CT_FPQ_CVT, CT_FPQ_PRMT, CSUB_SEL, CSUB_PRED, CSUB_SIGN, NOPS };
const char* names[] = {"mad.lo.u32", "mad.wide.u32", "mul.hi.u64", "mul.lo.u64", "shoup_lazy modmul", "CT butterfly (csub each)",
                       "CT butterfly (lazy, no csub)", "fma.f64", "fma.f64 + mad.lo.u32 (pairs)", "add.u64", "lop3", "GS butterfly",
                       "solinas 2^60-d modmul", "CT bfly approx-quotient (product)", "GS bfly approx-quotient (product)", "mul.hi.u32", "CT bfly FP64-assisted quotient", "GS bfly FP64-assisted quotient", "fma.f64 + mad.wide.u32 (pairs)", "CT bfly FPQ, I2F conversions", "CT bfly FPQ, PRMT-formed doubles", "csub select form", "csub predicated form", "csub sign-bit predicated"};

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
    const mfhe::twf_t twf = mfhe::fpq_prep(mfhe::tw_t{w, ws});
    const u64 cneg = mfhe::fpq_cneg(q);
    const u32 magic = 0x43300000u + (u32)(clk[1] & 0);
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
            if (OP == CT_APPROX) {
                mfhe::tw_t tw = {w, ws};
                u64 t = mfhe::mul_shoup_lazy4(y[i], tw, 0 - q);
                u64 xx = x[i];
                y[i] = xx + 2 * q2 - t; x[i] = xx + t;
                if ((it & 1) == 1) x[i] = csub(x[i], 8 * q);
            }
            if (OP == GS_APPROX) {
                mfhe::tw_t tw = {w, ws};
                u64 t = x[i] + 2 * q2 - y[i];
                x[i] = csub(x[i] + y[i], 2 * q2);
                y[i] = mfhe::mul_shoup_lazy4(t, tw, 0 - q);
            }
            if (OP == IMADHI32) a[i] = __umulhi(a[i], (u32)w) + (u32)ws;
            if (OP == DFMA_PLUS_WIDE) { d[i] = fma(d[i], 1.0000001, 1e-9); x[i] = (u64)(u32)x[i] * (u32)w + x[i]; }
            if (OP == CT_FPQ) {
                u64 t = mfhe::mul_fpq(y[i], twf, 0 - q, cneg);
                u64 xx = x[i];
                y[i] = xx + 2 * q2 - t; x[i] = xx + t;
                if ((it & 1) == 1) x[i] = csub(x[i], 8 * q);
            }
            if (OP == GS_FPQ) {
                u64 t = x[i] + 2 * q2 - y[i];
                x[i] = csub(x[i] + y[i], 2 * q2);
                y[i] = mfhe::mul_fpq(t, twf, 0 - q, cneg);
            }
            if (OP == CT_FPQ_CVT) {
                u64 t = mul_fpq_cvt(y[i], twf, 0 - q, cneg);
                u64 xx = x[i];
                y[i] = xx + 2 * q2 - t; x[i] = xx + t;
                if ((it & 1) == 1) x[i] = csub(x[i], 8 * q);
            }
            if (OP == CT_FPQ_PRMT) {
                u64 t = mul_fpq_prmt(y[i], twf, 0 - q, cneg, magic);
                u64 xx = x[i];
                y[i] = xx + 2 * q2 - t; x[i] = xx + t;
                if ((it & 1) == 1) x[i] = csub(x[i], 8 * q);
            }
            if (OP == CSUB_SEL) { x[i] = csub(x[i] + y[i], q2); }
            if (OP == CSUB_PRED) { x[i] = csub_pred(x[i] + y[i], q2); }
            if (OP == CSUB_SIGN) { x[i] = csub_sign(x[i] + y[i], 8 * q); }
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
    cudaMalloc(&out, (size_t)sms * bps * 256 * 8); cudaMalloc(&clk, 16); cudaMemset(clk, 0, 16);
    run<IMAD32>(out, clk, sms, bps);
    run<IMADWIDE>(out, clk, sms, bps);
    run<MULHI64>(out, clk, sms, bps);
    run<MULLO64>(out, clk, sms, bps);
    run<SHOUP_LAZY>(out, clk, sms, bps);
    run<CT_BFLY>(out, clk, sms, bps);
    run<CT_BFLY_LAZY4>(out, clk, sms, bps);
    run<GS_BFLY>(out, clk, sms, bps);
    run<SOLINAS60>(out, clk, sms, bps);
    run<CT_APPROX>(out, clk, sms, bps);
    run<GS_APPROX>(out, clk, sms, bps);
    run<IMADHI32>(out, clk, sms, bps);
    run<CT_FPQ>(out, clk, sms, bps);
    run<GS_FPQ>(out, clk, sms, bps);
    run<DFMA_PLUS_WIDE>(out, clk, sms, bps);
    run<CT_FPQ_CVT>(out, clk, sms, bps);
    run<CT_FPQ_PRMT>(out, clk, sms, bps);
    run<CSUB_SEL>(out, clk, sms, bps);
    run<CSUB_PRED>(out, clk, sms, bps);
    run<CSUB_SIGN>(out, clk, sms, bps);
    run<DFMA>(out, clk, sms, bps);
    run<DFMA_PLUS_IMAD>(out, clk, sms, bps);
    run<IADD64>(out, clk, sms, bps);
    run<LOP3>(out, clk, sms, bps);
    printf("err=%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
