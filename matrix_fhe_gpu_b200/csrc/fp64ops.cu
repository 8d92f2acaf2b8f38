// fp64ops.cu -- the floating-point half of encode / decode.
//
//   wdft_*        complex W-(I)DFT as one tiled complex GEMM over all n*n columns
//                 (replaces HE.cu:1116-1202 and batched_encoder.cu:104-123: one thread per output,
//                 table re-read from global memory per thread)
//   xy_sandwich   out = A * M * B for all lanes in ONE launch, M staged in shared memory
//                 (replaces the 512-iteration host loops of cudaMalloc + 2 launches of
//                 mat_mul_kernel_complex<<<(4,4),(16,16)>>> + cudaFree: encoder.cu:460-467, 492-501,
//                 batched_encoder.cu:192-196, HE.cu:1676-1679)
//   quantize      llround(z * delta) mod q_l (batched_encoder.cu:125-152, encoder.cu:36-50)
// fp64 results agree with the reference within summation-order round-off (tolerance-checked).
#include <cstdlib>
#include "ctx.h"

namespace mfhe {

constexpr int DTO = 64, DTC = 64, DTK = 8, DTHREADS = 256;

// IN: 0 interleaved complex, 1 two double planes, 2 two int64 planes.  OUT: 0 interleaved, 1 planes.
template <int IN, int OUT>
__global__ void __launch_bounds__(DTHREADS)
wdft_gemm_kernel(const double2* __restrict__ tab, const void* __restrict__ in0, const void* __restrict__ in1,
                 void* __restrict__ out0, void* __restrict__ out1, int phi, int cols) {
    __shared__ __align__(16) double2 sTab[2][DTK][DTO];
    __shared__ __align__(16) double2 sIn[2][DTK][DTC];
    const int t = threadIdx.x, tx = t & 15, ty = t >> 4;
    const int o0 = blockIdx.y * DTO, c0 = blockIdx.x * DTC;
    // staging: 8 rows x 64 entries = 512 entries per tile, 2 per thread
    const int lk = t >> 5, lg = (t & 31) << 1;
    double2 rt[2], ri[2];
    auto fetch = [&](int k0) {
        const int i = k0 + lk;
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            rt[j] = (i < phi && o0 + lg + j < phi) ? __ldg(tab + (size_t)i * phi + o0 + lg + j) : make_double2(0, 0);
            const int c = c0 + lg + j;
            if (i < phi && c < cols) {
                const size_t off = (size_t)i * cols + c;
                if (IN == 0) ri[j] = __ldg((const double2*)in0 + off);
                else if (IN == 1) ri[j] = make_double2(((const double*)in0)[off], ((const double*)in1)[off]);
                else ri[j] = make_double2((double)((const long long*)in0)[off], (double)((const long long*)in1)[off]);
            } else ri[j] = make_double2(0, 0);
        }
    };
    auto stash = [&](int buf) {
#pragma unroll
        for (int j = 0; j < 2; ++j) { sTab[buf][lk][lg + j] = rt[j]; sIn[buf][lk][lg + j] = ri[j]; }
    };
    double ar[4][4], ai[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) ar[i][j] = ai[i][j] = 0.0;
    fetch(0); stash(0);
    __syncthreads();
    const int ktiles = (phi + DTK - 1) / DTK;
    for (int kt = 0; kt < ktiles; ++kt) {
        const int buf = kt & 1;
        if (kt + 1 < ktiles) fetch((kt + 1) * DTK);
#pragma unroll
        for (int k = 0; k < DTK; ++k) {
            double2 a[4], b[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) { a[i] = sTab[buf][k][ty + 16 * i]; b[i] = sIn[buf][k][tx + 16 * i]; }   // lane-contiguous: no bank conflicts
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    ar[i][j] = fma(a[i].x, b[j].x, ar[i][j]); ar[i][j] = fma(-a[i].y, b[j].y, ar[i][j]);
                    ai[i][j] = fma(a[i].x, b[j].y, ai[i][j]); ai[i][j] = fma(a[i].y, b[j].x, ai[i][j]);
                }
        }
        if (kt + 1 < ktiles) stash(buf ^ 1);
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int o = o0 + ty + 16 * i;
        if (o >= phi) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int c = c0 + tx + 16 * j;
            if (c >= cols) continue;
            const size_t off = (size_t)o * cols + c;
            if (OUT == 0) ((double2*)out0)[off] = make_double2(ar[i][j], ai[i][j]);
            else { ((double*)out0)[off] = ar[i][j]; ((double*)out1)[off] = ai[i][j]; }
        }
    }
}

// ---- conjugate-pair forms: half the DFMAs of the complex GEMM ---------------------------------------------
// The evaluation points come in conjugate pairs (e, p - e), so the table rows of lanes w, w' are complex
// conjugates of each other (exactly for V, up to round-off for the Gauss-Jordan inverse, whose COLUMNS pair up).
// With t = c + i d the table entry of the pair's representative:
//   forward (pairs along the outputs):  E[w] = (Sca - Sdb) + i (Scb + Sda),  E[w'] = (Sca + Sdb) + i (Scb - Sda)
//       where Sca = sum c*a etc. over the inputs a + i b              -> 4 real FMAs per (pair, input), not 8;
//   inverse (pairs along the inputs):   out = sum_pairs c*(x+x') - d*(y-y') + i (c*(y+y') + d*(x-x'))
//       for the inputs x + i y (lane w) and x' + i y' (lane w')       -> 4 real FMAs per (output, pair).
template <int IN>
__device__ __forceinline__ double2 wdft_load(const void* in0, const void* in1, size_t off) {
    if (IN == 0) return __ldg((const double2*)in0 + off);
    if (IN == 1) return make_double2(((const double*)in0)[off], ((const double*)in1)[off]);
    return make_double2((double)((const long long*)in0)[off], (double)((const long long*)in1)[off]);
}
template <int OUT>
__device__ __forceinline__ void wdft_store(void* out0, void* out1, size_t off, double re, double im) {
    if (OUT == 0) ((double2*)out0)[off] = make_double2(re, im);
    else { ((double*)out0)[off] = re; ((double*)out1)[off] = im; }
}

constexpr int PTP = 32;      // forward: pairs per CTA tile (= 64 outputs)

// tabh[i][p] = tab[i][pairs[p].x], hp = phi / 2 pairs; CTA tile = 32 pairs x 64 columns, thread = 2 pairs x 4 columns
template <int IN, int OUT>
__global__ void __launch_bounds__(DTHREADS, 2)
wdft_opair_kernel(const double2* __restrict__ tabh, const int2* __restrict__ pairs, const void* __restrict__ in0,
                  const void* __restrict__ in1, void* __restrict__ out0, void* __restrict__ out1, int phi, int cols) {
    __shared__ __align__(16) double2 sTab[2][DTK][PTP];
    __shared__ __align__(16) double2 sIn[2][DTK][DTC];
    const int hp = phi >> 1;
    const int t = threadIdx.x, tx = t & 15, ty = t >> 4;
    const int p0 = blockIdx.y * PTP, c0 = blockIdx.x * DTC;
    const int lk = t >> 5, lp = t & 31, lg = (t & 31) << 1;
    double2 rt, ri[2];
    auto fetch = [&](int k0) {
        const int i = k0 + lk;
        rt = (i < phi && p0 + lp < hp) ? __ldg(tabh + (size_t)i * hp + p0 + lp) : make_double2(0, 0);
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int c = c0 + lg + j;
            ri[j] = (i < phi && c < cols) ? wdft_load<IN>(in0, in1, (size_t)i * cols + c) : make_double2(0, 0);
        }
    };
    auto stash = [&](int buf) {
        sTab[buf][lk][lp] = rt;
        *reinterpret_cast<double4*>(&sIn[buf][lk][lg]) = make_double4(ri[0].x, ri[0].y, ri[1].x, ri[1].y);
    };
    double ca[2][4], cb[2][4], da[2][4], db[2][4];
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) ca[i][j] = cb[i][j] = da[i][j] = db[i][j] = 0.0;
    fetch(0); stash(0);
    __syncthreads();
    const int ktiles = (phi + DTK - 1) / DTK;
    for (int kt = 0; kt < ktiles; ++kt) {
        const int buf = kt & 1;
        if (kt + 1 < ktiles) fetch((kt + 1) * DTK);
#pragma unroll
        for (int k = 0; k < DTK; ++k) {
            double2 a[2], b[4];
#pragma unroll
            for (int i = 0; i < 2; ++i) a[i] = sTab[buf][k][ty + 16 * i];
#pragma unroll
            for (int j = 0; j < 4; ++j) b[j] = sIn[buf][k][tx + 16 * j];
#pragma unroll
            for (int i = 0; i < 2; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    ca[i][j] = fma(a[i].x, b[j].x, ca[i][j]); cb[i][j] = fma(a[i].x, b[j].y, cb[i][j]);
                    da[i][j] = fma(a[i].y, b[j].x, da[i][j]); db[i][j] = fma(a[i].y, b[j].y, db[i][j]);
                }
        }
        if (kt + 1 < ktiles) stash(buf ^ 1);
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int p = p0 + ty + 16 * i;
        if (p >= hp) continue;
        const int2 ow = pairs[p];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int c = c0 + tx + 16 * j;
            if (c >= cols) continue;
            wdft_store<OUT>(out0, out1, (size_t)ow.x * cols + c, ca[i][j] - db[i][j], cb[i][j] + da[i][j]);
            wdft_store<OUT>(out0, out1, (size_t)ow.y * cols + c, ca[i][j] + db[i][j], cb[i][j] - da[i][j]);
        }
    }
}

// tabh[p][o] = tab[pairs[p].x][o]; CTA tile = 64 outputs x 64 columns, thread = 4 x 4, K runs over the hp pairs
// in steps of IKT = 4 (one table entry and one input pair per thread per step keeps the kernel at 128 registers,
// two CTAs per SM)
constexpr int IKT = 4;
template <int IN, int OUT>
__global__ void __launch_bounds__(DTHREADS, 2)
wdft_ipair_kernel(const double2* __restrict__ tabh, const int2* __restrict__ pairs, const void* __restrict__ in0,
                  const void* __restrict__ in1, void* __restrict__ out0, void* __restrict__ out1, int phi, int cols) {
    __shared__ __align__(16) double2 sTab[2][IKT][DTO];
    __shared__ __align__(16) double2 sS[2][IKT][DTC];      // (x + x', -(y - y'))
    __shared__ __align__(16) double2 sD[2][IKT][DTC];      // (y + y',   x - x')
    const int hp = phi >> 1;
    const int t = threadIdx.x, tx = t & 15, ty = t >> 4;
    const int o0 = blockIdx.y * DTO, c0 = blockIdx.x * DTC;
    const int lk = t >> 6, lg = t & 63;
    double2 rt, rs, rd;
    auto fetch = [&](int k0) {
        const int p = k0 + lk;
        rt = (p < hp && o0 + lg < phi) ? __ldg(tabh + (size_t)p * phi + o0 + lg) : make_double2(0, 0);
        const int c = c0 + lg;
        if (p < hp && c < cols) {
            const int2 ww = pairs[p];
            const double2 z = wdft_load<IN>(in0, in1, (size_t)ww.x * cols + c);
            const double2 w = wdft_load<IN>(in0, in1, (size_t)ww.y * cols + c);
            rs = make_double2(z.x + w.x, w.y - z.y);
            rd = make_double2(z.y + w.y, z.x - w.x);
        } else rs = rd = make_double2(0, 0);
    };
    auto stash = [&](int buf) { sTab[buf][lk][lg] = rt; sS[buf][lk][lg] = rs; sD[buf][lk][lg] = rd; };
    double ar[4][4], ai[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) ar[i][j] = ai[i][j] = 0.0;
    fetch(0); stash(0);
    __syncthreads();
    const int ktiles = (hp + IKT - 1) / IKT;
    for (int kt = 0; kt < ktiles; ++kt) {
        const int buf = kt & 1;
        if (kt + 1 < ktiles) fetch((kt + 1) * IKT);
#pragma unroll
        for (int k = 0; k < IKT; ++k) {
            double2 a[4], s[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) { a[i] = sTab[buf][k][ty + 16 * i]; s[i] = sS[buf][k][tx + 16 * i]; }
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) { ar[i][j] = fma(a[i].x, s[j].x, ar[i][j]); ar[i][j] = fma(a[i].y, s[j].y, ar[i][j]); }
#pragma unroll
            for (int i = 0; i < 4; ++i) s[i] = sD[buf][k][tx + 16 * i];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) { ai[i][j] = fma(a[i].x, s[j].x, ai[i][j]); ai[i][j] = fma(a[i].y, s[j].y, ai[i][j]); }
        }
        if (kt + 1 < ktiles) stash(buf ^ 1);
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int o = o0 + ty + 16 * i;
        if (o >= phi) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int c = c0 + tx + 16 * j;
            if (c < cols) wdft_store<OUT>(out0, out1, (size_t)o * cols + c, ar[i][j], ai[i][j]);
        }
    }
}

// ---- the same two forms on the FP64 tensor cores (mma.sync m8n8k4 .f64: 256 FMAs per warp instruction) ----------------
// Both conjugate-pair forms are REAL GEMMs.  Forward: rows (pair, c | d) x columns (col, a | b); the C fragment of a
// thread (row lane/4, two adjacent columns) is then (S.a, S.b) of one complex column, and with the c rows in m-tiles
// 0-1 and the d rows in m-tiles 2-3 of a warp the same thread also holds the matching d sums: the four sums of a pair
// meet in one thread.  Fragments (PTX ISA, mma.m8n8k4 .f64): A[row = lane/4][k = lane%4], B[k = lane%4][col = lane/4],
// C[row = lane/4][cols 2 (lane%4) + {0, 1}].  Shared rows are padded to stride % 32 == 8 doubles: a fragment load of a
// warp (4 k-rows x 8 consecutive doubles) then covers 32 distinct 8-byte slots, the 2-wavefront minimum.
__device__ __forceinline__ void dmma884(double (&c)[2], double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}
constexpr int MSP = PTP + 8, MSX = 2 * DTC + 8;      // padded row strides (doubles)

// CTA tile = 32 pairs x 64 complex columns, 8 warps = 2 (16 pairs) x 4 (16 columns)
template <int IN, int OUT>
__global__ void __launch_bounds__(DTHREADS, 2)
wdft_opair_dmma_kernel(const double2* __restrict__ tabh, const int2* __restrict__ pairs, const void* __restrict__ in0,
                       const void* __restrict__ in1, void* __restrict__ out0, void* __restrict__ out1, int phi, int cols) {
    __shared__ __align__(16) double sC[2][DTK][MSP];
    __shared__ __align__(16) double sD[2][DTK][MSP];
    __shared__ __align__(16) double sX[2][DTK][MSX];
    const int hp = phi >> 1;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5, wp = warp & 1, wc = warp >> 1;
    const int p0 = blockIdx.y * PTP, c0 = blockIdx.x * DTC;
    const int lk = t >> 5, lp = t & 31, lg = (t & 31) << 1;
    double2 rt, ri[2];
    auto fetch = [&](int k0) {
        const int i = k0 + lk;
        rt = (i < phi && p0 + lp < hp) ? __ldg(tabh + (size_t)i * hp + p0 + lp) : make_double2(0, 0);
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int c = c0 + lg + j;
            ri[j] = (i < phi && c < cols) ? wdft_load<IN>(in0, in1, (size_t)i * cols + c) : make_double2(0, 0);
        }
    };
    auto stash = [&](int buf) {
        sC[buf][lk][lp] = rt.x; sD[buf][lk][lp] = rt.y;
        *reinterpret_cast<double4*>(&sX[buf][lk][2 * lg]) = make_double4(ri[0].x, ri[0].y, ri[1].x, ri[1].y);
    };
    double aC[2][4][2], aD[2][4][2];
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) aC[i][j][0] = aC[i][j][1] = aD[i][j][0] = aD[i][j][1] = 0.0;
    fetch(0); stash(0);
    __syncthreads();
    const int ktiles = (phi + DTK - 1) / DTK;
    const int fk = lane & 3, fr = lane >> 2;
    for (int kt = 0; kt < ktiles; ++kt) {
        const int buf = kt & 1;
        if (kt + 1 < ktiles) fetch((kt + 1) * DTK);
#pragma unroll
        for (int kk = 0; kk < DTK / 4; ++kk) {
            double fc[2], fd[2], fb[4];
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                fc[i] = sC[buf][kk * 4 + fk][wp * 16 + i * 8 + fr];
                fd[i] = sD[buf][kk * 4 + fk][wp * 16 + i * 8 + fr];
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) fb[j] = sX[buf][kk * 4 + fk][wc * 32 + j * 8 + fr];
#pragma unroll
            for (int i = 0; i < 2; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) { dmma884(aC[i][j], fc[i], fb[j]); dmma884(aD[i][j], fd[i], fb[j]); }
        }
        if (kt + 1 < ktiles) stash(buf ^ 1);
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int p = p0 + wp * 16 + i * 8 + fr;
        if (p >= hp) continue;
        const int2 ow = pairs[p];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int c = c0 + wc * 16 + j * 4 + fk;
            if (c >= cols) continue;
            const double ca = aC[i][j][0], cb = aC[i][j][1], da = aD[i][j][0], db = aD[i][j][1];
            wdft_store<OUT>(out0, out1, (size_t)ow.x * cols + c, ca - db, cb + da);
            wdft_store<OUT>(out0, out1, (size_t)ow.y * cols + c, ca + db, cb - da);
        }
    }
}

// Inverse as a real GEMM: K runs over (pair, P | Q), B rows (pair, P) = (x + x', y + y'), (pair, Q) = (y' - y, x - x') per
// complex column, so C[o][(col, re | im)] is the result itself.  CTA tile = 64 outputs x 64 columns, 8 warps = 2 (32
// outputs) x 4 (16 columns), K tile = 4 pairs = 8 rows.
constexpr int MSO = DTO + 8;
template <int IN, int OUT>
__global__ void __launch_bounds__(DTHREADS, 2)
wdft_ipair_dmma_kernel(const double2* __restrict__ tabh, const int2* __restrict__ pairs, const void* __restrict__ in0,
                       const void* __restrict__ in1, void* __restrict__ out0, void* __restrict__ out1, int phi, int cols) {
    __shared__ __align__(16) double sA[2][DTK][MSO];
    __shared__ __align__(16) double sB[2][DTK][MSX];
    const int hp = phi >> 1;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5, wo = warp & 1, wc = warp >> 1;
    const int o0 = blockIdx.y * DTO, c0 = blockIdx.x * DTC;
    const int lpl = t >> 6, lg = t & 63;
    double2 rt, rs, rd;
    auto fetch = [&](int pk0) {
        const int p = pk0 + lpl;
        rt = (p < hp && o0 + lg < phi) ? __ldg(tabh + (size_t)p * phi + o0 + lg) : make_double2(0, 0);
        const int c = c0 + lg;
        if (p < hp && c < cols) {
            const int2 ww = pairs[p];
            const double2 z = wdft_load<IN>(in0, in1, (size_t)ww.x * cols + c);
            const double2 w = wdft_load<IN>(in0, in1, (size_t)ww.y * cols + c);
            rs = make_double2(z.x + w.x, z.y + w.y);
            rd = make_double2(w.y - z.y, z.x - w.x);
        } else rs = rd = make_double2(0, 0);
    };
    auto stash = [&](int buf) {
        sA[buf][2 * lpl][lg] = rt.x; sA[buf][2 * lpl + 1][lg] = rt.y;
        *reinterpret_cast<double2*>(&sB[buf][2 * lpl][2 * lg]) = rs;
        *reinterpret_cast<double2*>(&sB[buf][2 * lpl + 1][2 * lg]) = rd;
    };
    double acc[4][4][2];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
    fetch(0); stash(0);
    __syncthreads();
    constexpr int PK = DTK / 2;                        // pairs per K tile
    const int ktiles = (hp + PK - 1) / PK;
    const int fk = lane & 3, fr = lane >> 2;
    for (int kt = 0; kt < ktiles; ++kt) {
        const int buf = kt & 1;
        if (kt + 1 < ktiles) fetch((kt + 1) * PK);
#pragma unroll
        for (int kk = 0; kk < DTK / 4; ++kk) {
            double fa[4], fb[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) fa[i] = sA[buf][kk * 4 + fk][wo * 32 + i * 8 + fr];
#pragma unroll
            for (int j = 0; j < 4; ++j) fb[j] = sB[buf][kk * 4 + fk][wc * 32 + j * 8 + fr];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) dmma884(acc[i][j], fa[i], fb[j]);
        }
        if (kt + 1 < ktiles) stash(buf ^ 1);
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int o = o0 + wo * 32 + i * 8 + fr;
        if (o >= phi) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int c = c0 + wc * 16 + j * 4 + fk;
            if (c < cols) wdft_store<OUT>(out0, out1, (size_t)o * cols + c, acc[i][j][0], acc[i][j][1]);
        }
    }
}

template <int IN, int OUT>
static cudaError_t wdft_dispatch(const WdftTab& w, const void* in0, const void* in1, void* out0, void* out1, int phi,
                                 int cols, cudaStream_t st) {
    const int hp = phi / 2;
    if (w.mode == 1 && w.half && w.pairs) {
        dim3 grid((cols + DTC - 1) / DTC, (hp + PTP - 1) / PTP);
        static const bool dfma = std::getenv("MFHE_WDFT_DFMA") != nullptr;      // A/B switch: CUDA-core DFMA form
        if (dfma) wdft_opair_kernel<IN, OUT><<<grid, DTHREADS, 0, st>>>(w.half, w.pairs, in0, in1, out0, out1, phi, cols);
        else      wdft_opair_dmma_kernel<IN, OUT><<<grid, DTHREADS, 0, st>>>(w.half, w.pairs, in0, in1, out0, out1, phi, cols);
    } else if (w.mode == 2 && w.half && w.pairs) {
        dim3 grid((cols + DTC - 1) / DTC, (phi + DTO - 1) / DTO);
        static const bool dfma = std::getenv("MFHE_WDFT_DFMA") != nullptr;
        if (dfma) wdft_ipair_kernel<IN, OUT><<<grid, DTHREADS, 0, st>>>(w.half, w.pairs, in0, in1, out0, out1, phi, cols);
        else      wdft_ipair_dmma_kernel<IN, OUT><<<grid, DTHREADS, 0, st>>>(w.half, w.pairs, in0, in1, out0, out1, phi, cols);
    } else {
        dim3 grid((cols + DTC - 1) / DTC, (phi + DTO - 1) / DTO);
        wdft_gemm_kernel<IN, OUT><<<grid, DTHREADS, 0, st>>>(w.full, in0, in1, out0, out1, phi, cols);
    }
    return cudaGetLastError();
}
cudaError_t wdft_launch(const WdftTab& w, const double2* in, double2* out, int phi, int cols, cudaStream_t st) {
    return wdft_dispatch<0, 0>(w, in, nullptr, out, nullptr, phi, cols, st);
}
cudaError_t wdft_pair_launch(const WdftTab& w, const void* in_re, const void* in_im, double* out_re, double* out_im,
                             int phi, int cols, int int_input, cudaStream_t st) {
    return int_input ? wdft_dispatch<2, 1>(w, in_re, in_im, out_re, out_im, phi, cols, st)
                     : wdft_dispatch<1, 1>(w, in_re, in_im, out_re, out_im, phi, cols, st);
}

// llround(z * delta) of both parts of every complex value, once, as two int64 planes: the W-CRT loader then only
// reduces them modulo each limb (signed_mod) instead of converting the same double in every limb.
__global__ void llround_planes_kernel(const double2* __restrict__ in, long long* __restrict__ re, long long* __restrict__ im,
                                      size_t total, double delta) {
    const size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 2;
    if (i + 1 < total) {
        const double4 z = *reinterpret_cast<const double4*>(in + i);
        *reinterpret_cast<longlong2*>(re + i) = make_longlong2(llround(z.x * delta), llround(z.z * delta));
        *reinterpret_cast<longlong2*>(im + i) = make_longlong2(llround(z.y * delta), llround(z.w * delta));
    } else if (i < total) {
        const double2 z = in[i];
        re[i] = llround(z.x * delta); im[i] = llround(z.y * delta);
    }
}
cudaError_t llround_planes_launch(const double2* in, long long* re, long long* im, size_t total, double delta, cudaStream_t st) {
    llround_planes_kernel<<<(unsigned)((total / 2 + 256) / 256), 256, 0, st>>>(in, re, im, total, delta);
    return cudaGetLastError();
}

// ---- out = A * M * B per lane ---------------------------------------------------------------------
__device__ __forceinline__ double2 cfma(double2 a, double2 b, double2 c) {
    c.x = fma(a.x, b.x, c.x); c.x = fma(-a.y, b.y, c.x);
    c.y = fma(a.x, b.y, c.y); c.y = fma(a.y, b.x, c.y);
    return c;
}
// One CTA per lane, 256 threads as 16 x 16; thread (ty, tx) owns the R x R outputs (ty + 16 i, tx + 16 j), R = n / 16.
// T = A * M with M in shared memory and A streamed through the read-only path (it is the same n x n matrix for
// every lane, so it stays in L1/L2); out = T * B likewise.  R^2 complex FMAs per 2R loads.
template <int R>
__global__ void __launch_bounds__(256)
xy_sandwich_kernel(const double2* __restrict__ A, const double2* __restrict__ B, const double2* __restrict__ in,
                   double2* __restrict__ out) {
    constexpr int n = 16 * R;
    extern __shared__ __align__(16) double2 sm_xy[];
    double2* sM = sm_xy;                 // [n][n] the lane's matrix
    double2* sT = sm_xy + n * n;         // [n][n] A * M
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const double2* M = in + (size_t)blockIdx.x * n * n;
    for (int i = threadIdx.x; i < n * n; i += 256) sM[i] = M[i];
    __syncthreads();
    double2 acc[R][R];
#pragma unroll
    for (int i = 0; i < R; ++i)
#pragma unroll
        for (int j = 0; j < R; ++j) acc[i][j] = make_double2(0, 0);
#pragma unroll 4
    for (int k = 0; k < n; ++k) {
        double2 a[R], b[R];
#pragma unroll
        for (int i = 0; i < R; ++i) { a[i] = __ldg(A + (ty + 16 * i) * n + k); b[i] = sM[k * n + tx + 16 * i]; }
#pragma unroll
        for (int i = 0; i < R; ++i)
#pragma unroll
            for (int j = 0; j < R; ++j) acc[i][j] = cfma(a[i], b[j], acc[i][j]);
    }
#pragma unroll
    for (int i = 0; i < R; ++i)
#pragma unroll
        for (int j = 0; j < R; ++j) { sT[(ty + 16 * i) * n + tx + 16 * j] = acc[i][j]; acc[i][j] = make_double2(0, 0); }
    __syncthreads();
#pragma unroll 4
    for (int k = 0; k < n; ++k) {
        double2 a[R], b[R];
#pragma unroll
        for (int i = 0; i < R; ++i) { a[i] = sT[(ty + 16 * i) * n + k]; b[i] = __ldg(B + k * n + tx + 16 * i); }
#pragma unroll
        for (int i = 0; i < R; ++i)
#pragma unroll
            for (int j = 0; j < R; ++j) acc[i][j] = cfma(a[i], b[j], acc[i][j]);
    }
    double2* O = out + (size_t)blockIdx.x * n * n;
#pragma unroll
    for (int i = 0; i < R; ++i)
#pragma unroll
        for (int j = 0; j < R; ++j) O[(ty + 16 * i) * n + tx + 16 * j] = acc[i][j];
}
// n = 64 on the FP64 tensor cores.  A complex product C = A * M is the REAL GEMM
//     [C interleaved (n x 2n)] = [Ar | Ai] (n x 2n)  x  [ M interleaved ; M' ] (2n x 2n),   M'[k][(j, re | im)] = (-Mi, Mr)[k][j],
// and M' is just M read at the neighbouring part with a sign, so only M itself is staged.  One CTA per lane, 8 warps =
// 2 (32 rows) x 4 (32 real columns), 16 MMAs per 8 fragment loads; both products run through the same code, the first
// result goes back to shared memory as the (Tr | Ti) operand of the second.  Row strides: 132 doubles (== 4 mod 32) for
// the row-major A operand, 136 (== 8 mod 32) for the k-major B operand: conflict-free fragment loads.
constexpr int XSA = 132, XSM = 136;
__global__ void __launch_bounds__(256)
xy_sandwich_dmma64_kernel(const double2* __restrict__ A, const double2* __restrict__ B, const double2* __restrict__ in,
                          double2* __restrict__ out) {
    constexpr int n = 64;
    extern __shared__ __align__(16) double sm_xd[];
    double* sA = sm_xd;                    // [n][XSA]: (Ar | Ai), later (Tr | Ti)
    double* sM = sm_xd + n * XSA;          // [n][XSM]: M interleaved, later B
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5, wr = warp & 1, wc = warp >> 1;
    const int fk = lane & 3, fr = lane >> 2;
    const double2* M = in + (size_t)blockIdx.x * n * n;
    for (int e = t; e < n * n; e += 256) {
        const int r = e >> 6, c = e & 63;
        const double2 a = __ldg(A + e), m = M[e];
        sA[r * XSA + c] = a.x; sA[r * XSA + n + c] = a.y;
        *reinterpret_cast<double2*>(&sM[r * XSM + 2 * c]) = m;
    }
    __syncthreads();
    double acc[4][4][2];
    auto product = [&]() {
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
        const double* pa = sA + (wr * 32 + fr) * XSA + fk;
        const int col = wc * 32 + fr;
        const double sgn = (fr & 1) ? 1.0 : -1.0;                  // part 0 (re) of M' is -Mi
#pragma unroll 4
        for (int k4 = 0; k4 < n / 4; ++k4) {                       // K rows 0..n-1: (Ar, M)
            double fa[4], fb[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) fa[i] = pa[i * 8 * XSA + k4 * 4];
#pragma unroll
            for (int j = 0; j < 4; ++j) fb[j] = sM[(k4 * 4 + fk) * XSM + col + j * 8];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) dmma884(acc[i][j], fa[i], fb[j]);
        }
#pragma unroll 4
        for (int k4 = 0; k4 < n / 4; ++k4) {                       // K rows n..2n-1: (Ai, M')
            double fa[4], fb[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) fa[i] = pa[i * 8 * XSA + n + k4 * 4];
#pragma unroll
            for (int j = 0; j < 4; ++j) fb[j] = sgn * sM[(k4 * 4 + fk) * XSM + ((col + j * 8) ^ 1)];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) dmma884(acc[i][j], fa[i], fb[j]);
        }
    };
    product();                                                     // T = A * M
    __syncthreads();
    // C fragment: row wr*32 + i*8 + fr, complex column wc*16 + j*4 + fk, (re, im) = acc[i][j][0..1]
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int r = wr * 32 + i * 8 + fr, c = wc * 16 + j * 4 + fk;
            sA[r * XSA + c] = acc[i][j][0]; sA[r * XSA + n + c] = acc[i][j][1];
        }
    for (int e = t; e < n * n; e += 256) {
        const int r = e >> 6, c = e & 63;
        *reinterpret_cast<double2*>(&sM[r * XSM + 2 * c]) = __ldg(B + e);
    }
    __syncthreads();
    product();                                                     // out = T * B
    double2* O = out + (size_t)blockIdx.x * n * n;
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j)
            O[(wr * 32 + i * 8 + fr) * n + wc * 16 + j * 4 + fk] = make_double2(acc[i][j][0], acc[i][j][1]);
}
static cudaError_t xy_sandwich_dmma64_launch(const double2* A, const double2* B, const double2* in, double2* out, int lanes,
                                             cudaStream_t st) {
    constexpr size_t smem = (size_t)64 * (XSA + XSM) * sizeof(double);
    {   // set on every call: cheap, and right for any device / thread (no process-wide one-shot flag)
        cudaError_t e = cudaFuncSetAttribute(xy_sandwich_dmma64_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    xy_sandwich_dmma64_kernel<<<lanes, 256, smem, st>>>(A, B, in, out);
    return cudaGetLastError();
}

template <int R>
static cudaError_t xy_sandwich_launch_r(const double2* A, const double2* B, const double2* in, double2* out, int lanes,
                                        cudaStream_t st) {
    constexpr int n = 16 * R;
    constexpr size_t smem = 2 * (size_t)n * n * sizeof(double2);
    {   // set on every call: cheap, and right for any device / thread (no process-wide one-shot flag)
        cudaError_t e = cudaFuncSetAttribute(xy_sandwich_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    xy_sandwich_kernel<R><<<lanes, 256, smem, st>>>(A, B, in, out);
    return cudaGetLastError();
}
cudaError_t xy_sandwich_launch(const double2* A, const double2* B, const double2* in, double2* out, int n, int lanes,
                               cudaStream_t st) {
    switch (n) {
        case 16: return xy_sandwich_launch_r<1>(A, B, in, out, lanes, st);
        case 32: return xy_sandwich_launch_r<2>(A, B, in, out, lanes, st);
        case 64: {
            static const bool dfma = std::getenv("MFHE_XY_DFMA") != nullptr;      // A/B switch: CUDA-core version
            return dfma ? xy_sandwich_launch_r<4>(A, B, in, out, lanes, st) : xy_sandwich_dmma64_launch(A, B, in, out, lanes, st);
        }
        default: return cudaErrorInvalidValue;
    }
}

__global__ void matmul_complex_kernel(const double2* A, const double2* B, double2* C, int n) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n * n) return;
    const int r = idx / n, c = idx % n;
    double2 acc = make_double2(0, 0);
    for (int k = 0; k < n; ++k) acc = cfma(A[(size_t)r * n + k], B[(size_t)k * n + c], acc);
    C[idx] = acc;
}
cudaError_t matmul_complex_launch(const double2* A, const double2* B, double2* C, int n, cudaStream_t st) {
    matmul_complex_kernel<<<(n * n + 255) / 256, 256, 0, st>>>(A, B, C, n);
    return cudaGetLastError();
}

// ---- quantise ----------------------------------------------------------------------------------------
__global__ void quantize_kernel(const double2* __restrict__ in, u64* __restrict__ re, u64* __restrict__ im,
                                const u64* __restrict__ moduli, const ulonglong2* __restrict__ mu, int n2, int limbs,
                                size_t total, double delta) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const size_t lane = idx / n2, pos = idx % n2;
    const double2 z = in[idx];
    const long long xr = llround(z.x * delta), xi = llround(z.y * delta);
    for (int l = 0; l < limbs; ++l) {
        const u64 q = moduli[l], mh = mu[l].y;
        const size_t o = (lane * limbs + l) * n2 + pos;
        re[o] = signed_mod(xr, q, mh);
        im[o] = signed_mod(xi, q, mh);
    }
}
cudaError_t quantize_launch(const double2* in, u64* re, u64* im, const u64* moduli, const ulonglong2* mu, int n2,
                            int limbs, int lanes, double delta, cudaStream_t st) {
    const size_t total = (size_t)lanes * n2;
    quantize_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(in, re, im, moduli, mu, n2, limbs, total, delta);
    return cudaGetLastError();
}

}  // namespace mfhe
