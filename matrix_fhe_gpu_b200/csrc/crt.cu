// crt.cu -- exact CRT composition of RNS residues into a centred big integer and its fp64 image.
//
// Replaces src/core/encoder.cu:112-245 (dequantize_exact / crt_compose_centerlift(_big) kernels,
// launched once per lane from a 512-iteration host loop at HE.cu:1653-1668) and
// HE.cu:1007-1027 (compose_big_pair_to_complex_by_delta_kernel).  One launch covers all lanes;
// the big integer (W 64-bit words, 7 at L = 11, 14 at L = 24) lives in registers; mul_mod by the
// CRT inverse is a Barrett product instead of `unsigned __int128 %`.  Same integer, so the same
// (magnitude, sign) words and the same fp64 Horner value as the reference.
#include "ctx.h"

namespace mfhe {

template <int WC>
__device__ __forceinline__ void crt_compose(const CrtDev& c, const u64* __restrict__ res, size_t stride, int limbs,
                                            u64 (&mag)[WC], bool& neg) {
    u64 acc[WC];
#pragma unroll
    for (int i = 0; i < WC; ++i) acc[i] = 0;
    const int W = c.W;
    for (int k = 0; k < limbs; ++k) {
        const u64 q = c.moduli[k];
        const ulonglong2 mu = c.mu[k];
        const u64 t = mul_barrett(res[(size_t)k * stride], c.inv[k], q, mu.x, mu.y);
        const u64* M = c.M + (size_t)k * W;
        u64 mc = 0, ac = 0;                      // multiply carry, add carry
#pragma unroll
        for (int i = 0; i < WC; ++i) {
            if (i < W) {
                const u64 m = __ldg(M + i);
                const u64 lo = m * t, hi = __umul64hi(m, t);
                const u64 term = lo + mc;
                mc = hi + (term < lo);
                const u64 s = acc[i] + term, c1 = s < term;
                const u64 s2 = s + ac;
                ac = c1 | (s2 < s);
                acc[i] = s2;
            }
        }
        // acc >= Q ?  (acc < 2Q always)
        bool ge = true;
#pragma unroll
        for (int i = WC - 1; i >= 0; --i) {
            if (i < W) {
                const u64 qi = __ldg(c.Q + i);
                if (acc[i] != qi) { ge = acc[i] > qi; break; }
            }
        }
        if (ge) {
            u64 br = 0;
#pragma unroll
            for (int i = 0; i < WC; ++i) {
                if (i < W) {
                    const u64 qi = __ldg(c.Q + i);
                    const u64 d = acc[i] - qi, b1 = acc[i] < qi;
                    const u64 d2 = d - br;
                    br = b1 | (d < br);
                    acc[i] = d2;
                }
            }
        }
    }
    bool gt = false;                              // acc > Q/2 ?
#pragma unroll
    for (int i = WC - 1; i >= 0; --i) {
        if (i < W) {
            const u64 hi = __ldg(c.Qhalf + i);
            if (acc[i] != hi) { gt = acc[i] > hi; break; }
        }
    }
    neg = gt;
    if (gt) {
        u64 br = 0;
#pragma unroll
        for (int i = 0; i < WC; ++i) {
            if (i < W) {
                const u64 qi = __ldg(c.Q + i);
                const u64 d = qi - acc[i], b1 = qi < acc[i];
                const u64 d2 = d - br;
                br = b1 | (d < br);
                mag[i] = d2;
            } else mag[i] = 0;
        }
    } else {
#pragma unroll
        for (int i = 0; i < WC; ++i) mag[i] = i < W ? acc[i] : 0;
    }
}

// Fast exact path for values that fit two limbs (the decode case: |v| ~ delta * |message| << q0 q1).
// Candidate v' = centred Garner lift of (x0, x1) to (-q0 q1 / 2, q0 q1 / 2); it IS the full CRT value iff
// v' = x_j (mod q_j) for every remaining limb (both lie in (-Q/2, Q/2) and agree mod Q), which is checked
// exactly, one 64- or 128-bit Barrett per limb.  On a mismatch the caller runs the full composition, so
// results never differ from crt_compose; only the cost does (~350 instead of ~2000+ integer-pipe slots).
__device__ __forceinline__ bool crt_small(const CrtDev& c, const u64* __restrict__ res, size_t stride, int limbs,
                                          u64& mlo, u64& mhi, bool& neg) {
    const u64 q0 = c.moduli[0], q1 = c.moduli[1];
    const ulonglong2 mu1 = c.mu[1];
    const u64 x0 = res[0], x1 = res[stride];
    const u64 d = sub_mod(x1, reduce64(x0, q1, mu1.y), q1);
    const u64 t = mul_barrett(d, c.ginv, q1, mu1.x, mu1.y);            // (x1 - x0) / q0 mod q1
    u64 lo = q0 * t, hi = __umul64hi(q0, t);
    lo += x0; hi += lo < x0;                                            // v = x0 + q0 t  in [0, q0 q1)
    neg = hi > c.H2hi || (hi == c.H2hi && lo > c.H2lo);
    if (neg) { const u64 b = c.Q2lo < lo; lo = c.Q2lo - lo; hi = c.Q2hi - hi - b; }
    bool ok = true;
    for (int k = 2; k < limbs; ++k) {
        const u64 q = c.moduli[k];
        const ulonglong2 mu = c.mu[k];
        u64 m = hi ? reduce128(lo, hi, q, mu.x, mu.y) : reduce64(lo, q, mu.y);
        if (neg) m = neg_mod(m, q);
        ok &= m == res[(size_t)k * stride];
    }
    mlo = lo; mhi = hi;
    return ok;
}

template <int WC>
__device__ __forceinline__ double big_to_f64(const u64 (&a)[WC], int W) {
    const double two64 = 18446744073709551616.0;
    double v = 0.0;
#pragma unroll
    for (int i = WC - 1; i >= 0; --i)
        if (i < W) v = v * two64 + (double)a[i];     // Horner exactly as HE.cu:917-924
    return v;
}

template <int WC>
__global__ void crt_compose_big_kernel(CrtDev c, const u64* __restrict__ in, u64* __restrict__ mag_out,
                                       uint8_t* __restrict__ neg_out, int n2, int limbs, size_t total, int lane_limbs) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const size_t lane = idx / n2, pos = idx % n2;
    u64 mag[WC]; bool neg;
    crt_compose<WC>(c, in + lane * (size_t)lane_limbs * n2 + pos, (size_t)n2, limbs, mag, neg);
#pragma unroll
    for (int i = 0; i < WC; ++i)
        if (i < c.W) mag_out[idx * c.W + i] = mag[i];
    neg_out[idx] = neg ? 1 : 0;
}

template <int WC>
__global__ void crt_compose_i64_kernel(CrtDev c, const u64* __restrict__ in, long long* __restrict__ out, int n2, int limbs,
                                       size_t total, int saturate) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const size_t lane = idx / n2, pos = idx % n2;
    u64 mag[WC]; bool neg;
    crt_compose<WC>(c, in + lane * (size_t)limbs * n2 + pos, (size_t)n2, limbs, mag, neg);
    long long v;
    if (saturate) {                    // HE.cu:902-915 he_big_to_i64_checked
        bool over = mag[0] > (u64)0x7fffffffffffffffULL;
#pragma unroll
        for (int i = 1; i < WC; ++i) over |= (i < c.W) && mag[i] != 0;
        v = over ? (neg ? (long long)0x8000000000000000ULL : (long long)0x7fffffffffffffffULL)
                 : (neg ? -(long long)mag[0] : (long long)mag[0]);
    } else {                           // encoder.cu:186-188: low word only
        v = (long long)mag[0];
        v = neg ? -v : v;
    }
    out[idx] = v;
}

template <int WC>
__global__ void crt_to_complex_kernel(CrtDev c, const u64* __restrict__ in_re, const u64* __restrict__ in_im,
                                      double2* __restrict__ out, int n2, int limbs, size_t total, double delta) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const size_t lane = idx / n2, pos = idx % n2;
    const double two64 = 18446744073709551616.0;
    double v[2];
#pragma unroll 1
    for (int part = 0; part < 2; ++part) {
        const u64* res = (part ? in_im : in_re) + lane * (size_t)limbs * n2 + pos;
        u64 lo, hi; bool neg;
        double x;
        if (c.fast && limbs == c.L && crt_small(c, res, (size_t)n2, limbs, lo, hi, neg)) {
            // the Horner sum of big_to_f64 with every word above the second equal to zero
            x = c.W >= 2 ? (double)hi * two64 + (double)lo : (double)lo;
        } else {
            u64 mag[WC];
            crt_compose<WC>(c, res, (size_t)n2, limbs, mag, neg);
            x = big_to_f64<WC>(mag, c.W);
        }
        v[part] = (neg ? -x : x) / delta;
    }
    out[idx] = make_double2(v[0], v[1]);
}

#define MFHE_W_DISPATCH(W, CALL)                                   \
    if ((W) <= 4) { constexpr int WC = 4; CALL; }                  \
    else if ((W) <= 8) { constexpr int WC = 8; CALL; }             \
    else if ((W) <= 16) { constexpr int WC = 16; CALL; }           \
    else if ((W) <= 32) { constexpr int WC = 32; CALL; }           \
    else return cudaErrorInvalidValue;

cudaError_t crt_compose_big_launch(const CrtDev& c, const u64* in, u64* mag, uint8_t* neg, int n2, int limbs, int lanes,
                                   int lane_stride_limbs, cudaStream_t st) {
    const size_t total = (size_t)lanes * n2;
    const unsigned grid = (unsigned)((total + 127) / 128);
    MFHE_W_DISPATCH(c.W, (crt_compose_big_kernel<WC><<<grid, 128, 0, st>>>(c, in, mag, neg, n2, limbs, total, lane_stride_limbs)))
    return cudaGetLastError();
}
cudaError_t crt_compose_i64_launch(const CrtDev& c, const u64* in, long long* out, int n2, int limbs, int lanes,
                                   int saturate, cudaStream_t st) {
    const size_t total = (size_t)lanes * n2;
    const unsigned grid = (unsigned)((total + 127) / 128);
    MFHE_W_DISPATCH(c.W, (crt_compose_i64_kernel<WC><<<grid, 128, 0, st>>>(c, in, out, n2, limbs, total, saturate)))
    return cudaGetLastError();
}
cudaError_t crt_to_complex_launch(const CrtDev& c, const u64* in_re, const u64* in_im, double2* out, int n2, int limbs,
                                  int lanes, double delta, cudaStream_t st) {
    const size_t total = (size_t)lanes * n2;
    const unsigned grid = (unsigned)((total + 127) / 128);
    MFHE_W_DISPATCH(c.W, (crt_to_complex_kernel<WC><<<grid, 128, 0, st>>>(c, in_re, in_im, out, n2, limbs, total, delta)))
    return cudaGetLastError();
}

__global__ void big_pair_to_complex_kernel(const u64* __restrict__ re_mag, const uint8_t* __restrict__ re_neg,
                                           const u64* __restrict__ im_mag, const uint8_t* __restrict__ im_neg,
                                           double2* __restrict__ out, size_t total, int W, double delta) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const double two64 = 18446744073709551616.0;
    double vr = 0.0, vi = 0.0;
    for (int i = W - 1; i >= 0; --i) { vr = vr * two64 + (double)re_mag[idx * W + i]; vi = vi * two64 + (double)im_mag[idx * W + i]; }
    if (re_neg[idx]) vr = -vr;
    if (im_neg[idx]) vi = -vi;
    out[idx] = make_double2(vr / delta, vi / delta);
}
cudaError_t big_pair_to_complex_launch(const u64* re_mag, const uint8_t* re_neg, const u64* im_mag, const uint8_t* im_neg,
                                       double2* out, size_t total, int W, double delta, cudaStream_t st) {
    big_pair_to_complex_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(re_mag, re_neg, im_mag, im_neg, out, total, W, delta);
    return cudaGetLastError();
}

// HE.cu:1112-1113: residue mod q -> centred int64
__global__ void center_i64_kernel(const u64* __restrict__ in, long long* __restrict__ out, u64 q, size_t total) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const u64 a = in[idx];
    out[idx] = a > (q >> 1) ? (long long)a - (long long)q : (long long)a;
}
cudaError_t center_i64_launch(const u64* in, long long* out, u64 q, size_t total, cudaStream_t st) {
    center_i64_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(in, out, q, total);
    return cudaGetLastError();
}

}  // namespace mfhe
