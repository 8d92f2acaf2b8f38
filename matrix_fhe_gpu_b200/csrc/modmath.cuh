// modmath.cuh -- 64-bit modular arithmetic for the sm_100a kernels.
//
// Replaces the reference's device helpers:
//   include/core/common.cuh:12-39      add_mod / sub_mod / mul_mod (u128 %) / pow_mod
//   include/backend/phantom_math.cuh:9-28   compute_shoup / multiply_and_reduce_shoup
//   ph/include/uintmodmath.cuh:18-21,160-231  csub_q, lazy Shoup, Barrett
// All public results are canonical in [0, q); q < 2^60 (checked at plan creation)
// which leaves four spare bits for lazy accumulation up to 16q.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace mfhe {

typedef unsigned long long u64;
typedef unsigned int u32;

// {w, floor(w * 2^64 / q)} -- one 16-byte twiddle entry, one LDG.128.
typedef ulonglong2 tw_t;

#ifndef MFHE_CSUB_VARIANT
#define MFHE_CSUB_VARIANT 2
#endif
__device__ __forceinline__ u64 csub(u64 x, u64 m) {       // x in [0, 2m) -> [0, m)
#if MFHE_CSUB_VARIANT == 2
    // The borrow of the 64-bit subtraction IS the predicate: IADD3 + IADD3.X (carry out) + 2 SEL = 4 ALU slots instead of
    // the 6 of the compare form below (IADD3, IADD3.X, ISETP, ISETP.EX, SEL, SEL).  In the NTT row kernels a third of the
    // ALU-pipe instructions are conditional subtracts, and ptxas balances the pipes by instruction COUNT -- it moves
    // register copies and carry adds onto the fmaheavy pipe (IMAD.MOV, IMAD.X) when the ALU pipe is the fuller one, which
    // is the pipe that binds these kernels -- so every ALU slot saved here also takes work off the multiplier pipe
    // (cuobjdump, ntt_rows_fwd_kernel<16,2>: 1976 -> 1848 instructions, IMAD.MOV + IMAD.X 298 -> 203).
    const u32 xl = (u32)x, xh = (u32)(x >> 32), ml = (u32)m, mh = (u32)(m >> 32);
    u32 tl, th, b;
    asm("sub.cc.u32 %0, %3, %5;\n\tsubc.cc.u32 %1, %4, %6;\n\tsubc.u32 %2, 0, 0;"
        : "=r"(tl), "=r"(th), "=r"(b) : "r"(xl), "r"(xh), "r"(ml), "r"(mh));
    tl = b ? xl : tl; th = b ? xh : th;
    return ((u64)th << 32) | tl;
#else
    u64 t = x - m;
    return t < x ? t : x;                                 // unsigned min trick: wraps when x < m
#endif
}

// x * w mod q, lazily: any 64-bit x, result in [0, 2q).
__device__ __forceinline__ u64 mul_shoup_lazy(u64 x, tw_t w, u64 q) {
    u64 hi = __umul64hi(x, w.y);
    return x * w.x - hi * q;
}
__device__ __forceinline__ u64 mul_shoup(u64 x, tw_t w, u64 q) {
    return csub(mul_shoup_lazy(x, w, q), q);
}

// compile-time conditional-subtract chain: value known < B*q is brought below T*q.
template <int B, int T>
__device__ __forceinline__ void reduce_bound(u64& x, u64 q) {
    if constexpr (B > 8 && T <= 8) x = csub(x, 8 * q);
    if constexpr (B > 4 && T <= 4) x = csub(x, 4 * q);
    if constexpr (B > 2 && T <= 2) x = csub(x, 2 * q);
    if constexpr (B > 1 && T <= 1) x = csub(x, q);
}

// x * w mod q with an APPROXIMATE Shoup quotient: any 64-bit x, result in [0, 4q).
// Qhat = x1*s1 + hi32(x1*s0) + hi32(x0*s1) drops x0*s0 and the carries of the two cross terms, so it is
// within 2 of floor(x*w'/2^64); the remainder x*w - Qhat*q is formed mod 2^64 as x*w + Qhat*(-q) with the
// product-accumulate chain below: 3 IMAD.WIDE + 2 IMAD.HI + 4 IMAD and two carry adds, against
// 6 IMAD.WIDE + 4 IMAD + ~9 adds for the exact form (measured with cuobjdump, profiles/microbench).
__device__ __forceinline__ u64 mul_shoup_lazy4(u64 x, tw_t w, u64 nq) {
    const u32 x0 = (u32)x, x1 = (u32)(x >> 32), s0 = (u32)w.y, s1 = (u32)(w.y >> 32);
    const u32 w0 = (u32)w.x, w1 = (u32)(w.x >> 32), n0 = (u32)nq, n1 = (u32)(nq >> 32);
    u32 h0, h1, ql, qh, t0, t1;
    asm("mul.hi.u32 %0, %2, %3;\n\t"
        "mul.hi.u32 %1, %4, %5;"
        : "=r"(h0), "=r"(h1) : "r"(x1), "r"(s0), "r"(x0), "r"(s1));
    asm("mad.lo.cc.u32 %0, %2, %3, %4;\n\t"
        "madc.hi.u32 %1, %2, %3, 0;\n\t"
        "add.cc.u32 %0, %0, %5;\n\t"
        "addc.u32 %1, %1, 0;"
        : "=r"(ql), "=r"(qh) : "r"(x1), "r"(s1), "r"(h0), "r"(h1));
    asm("mul.lo.u32 %0, %2, %3;\n\t"
        "mul.hi.u32 %1, %2, %3;\n\t"
        "mad.lo.u32 %1, %2, %4, %1;\n\t"
        "mad.lo.u32 %1, %5, %3, %1;\n\t"
        "mad.lo.cc.u32 %0, %6, %7, %0;\n\t"
        "madc.hi.u32 %1, %6, %7, %1;\n\t"
        "mad.lo.u32 %1, %6, %8, %1;\n\t"
        "mad.lo.u32 %1, %9, %7, %1;"
        : "=&r"(t0), "=&r"(t1) : "r"(x0), "r"(w0), "r"(w1), "r"(x1), "r"(ql), "r"(n0), "r"(n1), "r"(qh));
    return ((u64)t1 << 32) | t0;
}

// ---- FP64-assisted approximate quotient -----------------------------------------------------------
// B200 measurements (profiles/microbench): IMAD.WIDE / IMAD.HI run at ~0.4x the rate of a 32-bit IMAD on
// the fmaheavy pipe, which is the pipe that binds the NTT kernels, while DFMA issues at the IMAD rate on
// its own pipe.  The two cross terms  (x1*s0 + x0*s1) >> 32  of the Shoup quotient only have to be right
// to within a couple of units, so they are moved to the FP64 pipe:
//   * x0, x1 enter as doubles without a conversion: the bit pattern {x_i, 0x43300000} IS 2^52 + x_i;
//   * the twiddle's Shoup word is converted once per twiddle (two I2F on the otherwise idle XU pipe) and
//     K = C - 2^52 (s0 + s1),  C = 1.5 * 2^84 - 2^32,  cancels the 2^52 offsets (exact: multiples of 2^32);
//   * D = fma(m1, s0, fma(m0, s1, K)) = 1.5 * 2^84 + (x1 s0 + x0 s1) - 2^32 + eps, |eps| <= 2^32 (two
//     roundings at ulp 2^32), so the LOW 52 bits of D's pattern hold c = ((x1 s0 + x0 s1) >> 32) - 1 +- 1
//     -- never above the true value -- and its pattern is kFpqOff + c;
//   * Q + kFpqOff = x1*s1 + D_bits is one IMAD.WIDE with D's registers as the addend, and the constant
//     kFpqOff * (-q) is cancelled by seeding the remainder chain with  cneg = -(kFpqOff * (-q)) mod 2^64.
// Q is in (rho - 4, rho] for rho = x*w/q, hence the product is in [0, 4q) exactly like mul_shoup_lazy4,
// at 3 IMAD.WIDE + 4 IMAD + 2 DFMA (+2 register moves) instead of 3 IMAD.WIDE + 2 IMAD.HI + 4 IMAD + 2 adds.
constexpr u64 kFpqOff = 0x4538000000000000ull;            // bit pattern of 1.5 * 2^84
#ifndef MFHE_FPQ_MODE
#define MFHE_FPQ_MODE 0      // how x0, x1 become doubles: 0 = {x_i, 0x43300000} register pairs, 1 = I2F (XU pipe), 2 = PRMT copies
#endif
#ifndef MFHE_MODMUL
#define MFHE_MODMUL 1      // butterfly product: 0 = FP64-assisted quotient, 1 = integer-only approximate quotient (mul_shoup_lazy4).
                           // In isolation 0 is 9-18 % faster (profiles/r1_04_intpipe.log); inside the NTT kernels its extra live
                           // registers (prepared twiddle = 9 words) cost occupancy or spills and it measured SLOWER
                           // (profiles/r1_06_ntt_ab.txt), so 1 is the default.
#endif
#if MFHE_MODMUL == 1
struct twf_t { tw_t t; };
__device__ __forceinline__ twf_t fpq_prep(tw_t t) { return twf_t{t}; }
#else
struct twf_t { u64 w; u32 s1; double s0d, s1d, k; };      // a twiddle prepared for mul_fpq
__device__ __forceinline__ twf_t fpq_prep(tw_t t) {
    twf_t r;
    r.w = t.x; r.s1 = (u32)(t.y >> 32);
    r.s0d = __uint2double_rn((u32)t.y);
    r.s1d = __uint2double_rn(r.s1);
    const double c = 29014219670751100192948224.0 - 4294967296.0;          // 1.5 * 2^84 - 2^32
#if MFHE_FPQ_MODE == 1
    r.k = c;                                             // true conversions of x0, x1: no 2^52 offsets to cancel
#else
    r.k = fma(-4503599627370496.0, r.s0d, fma(-4503599627370496.0, r.s1d, c));
#endif
    return r;
}
#endif
// cneg = -(kFpqOff * (0 - q)) mod 2^64 = kFpqOff * q mod 2^64
__device__ __forceinline__ u64 fpq_cneg(u64 q) { return kFpqOff * q; }
#if MFHE_MODMUL == 1
__device__ __forceinline__ u64 mul_fpq(u64 x, const twf_t& w, u64 nq, u64) { return mul_shoup_lazy4(x, w.t, nq); }
#else
__device__ __forceinline__ u64 mul_fpq(u64 x, const twf_t& w, u64 nq, u64 cneg) {
    const u32 x0 = (u32)x, x1 = (u32)(x >> 32);
#if MFHE_FPQ_MODE == 1
    const double d = fma(__uint2double_rn(x1), w.s0d, fma(__uint2double_rn(x0), w.s1d, w.k));
#elif MFHE_FPQ_MODE == 2
    u32 a0, a1;
    asm("prmt.b32 %0, %1, %1, 0x3210;" : "=r"(a0) : "r"(x0));
    asm("prmt.b32 %0, %1, %1, 0x3210;" : "=r"(a1) : "r"(x1));
    const double m0 = __hiloint2double(0x43300000, (int)a0), m1 = __hiloint2double(0x43300000, (int)a1);
    const double d = fma(m1, w.s0d, fma(m0, w.s1d, w.k));
#else
    const double m0 = __hiloint2double(0x43300000, (int)x0), m1 = __hiloint2double(0x43300000, (int)x1);
    const double d = fma(m1, w.s0d, fma(m0, w.s1d, w.k));
#endif
    const u32 d0 = (u32)__double2loint(d), d1 = (u32)__double2hiint(d);
    const u32 w0 = (u32)w.w, w1 = (u32)(w.w >> 32), n0 = (u32)nq, n1 = (u32)(nq >> 32);
    const u32 c0 = (u32)cneg, c1 = (u32)(cneg >> 32);
    u32 ql, qh, t0, t1;
    asm("mad.lo.cc.u32 %0, %2, %3, %4;\n\t"
        "madc.hi.u32 %1, %2, %3, %5;"
        : "=r"(ql), "=r"(qh) : "r"(x1), "r"(w.s1), "r"(d0), "r"(d1));
    asm("mad.lo.cc.u32 %0, %2, %3, %10;\n\t"
        "madc.hi.u32 %1, %2, %3, %11;\n\t"
        "mad.lo.u32 %1, %2, %4, %1;\n\t"
        "mad.lo.u32 %1, %5, %3, %1;\n\t"
        "mad.lo.cc.u32 %0, %6, %7, %0;\n\t"
        "madc.hi.u32 %1, %6, %7, %1;\n\t"
        "mad.lo.u32 %1, %6, %8, %1;\n\t"
        "mad.lo.u32 %1, %9, %7, %1;"
        : "=&r"(t0), "=&r"(t1)
        : "r"(x0), "r"(w0), "r"(w1), "r"(x1), "r"(ql), "r"(n0), "r"(n1), "r"(qh), "r"(c0), "r"(c1));
    return ((u64)t1 << 32) | t0;
}
#endif

// ---- lazy-bound schedule of the forward (Cooley-Tukey) network -------------
// A CT butterfly  x' = x + t, y' = x + 4q - t  with t = w*y in [0,4q) (approximate quotient) raises the
// bound of both outputs to bound(x) + 4 (in units of q) whatever y was, because the Shoup product accepts
// any 64-bit y.  With q < 2^60 values may grow to 16q, so x needs a conditional subtract only every second
// or third stage instead of every butterfly (ph/include/butterfly.cuh:10-22 subtracts in every butterfly).
// The last stage uses the exact quotient (t < 2q) so that two conditional subtracts canonicalise.
constexpr int kFwdInBound = 2;                       // inputs may be in [0, 2q)
__host__ __device__ constexpr int fwd_xtarget(int B, int s, int logn) {
    if (s == logn - 1) return 2;                     // last stage: leave < 4q for the final csubs
    return (B + 4 > 16) ? 8 : B;
}
__host__ __device__ constexpr int fwd_bound_before(int s, int logn) {
    int B = kFwdInBound;
    for (int i = 0; i < s; ++i) B = fwd_xtarget(B, i, logn) + 4;
    return B;
}

// Forward butterfly at global stage S of a 2^LOGN transform.  `w` is the prepared twiddle (FP64-assisted
// quotient); the last stage takes the raw {w, w'} pair and the exact quotient so that two conditional
// subtracts canonicalise the outputs.
struct qconst_t { u64 q, q2, q4, nq, cneg; };
__device__ __forceinline__ qconst_t make_qconst(u64 q) { return qconst_t{q, 2 * q, 4 * q, 0 - q, fpq_cneg(q)}; }
// per-arithmetic constants of a limb: AR = 0 the integer forms above, AR = 1 the FP64-pipe forms below (q < 2^50)
template <int AR> struct arith_const;

template <int LOGN, int S>
__device__ __forceinline__ void ct_bfly(u64& x, u64& y, const twf_t& w, tw_t raw, const qconst_t& c) {
    constexpr int B = fwd_bound_before(S, LOGN);
    constexpr int T = fwd_xtarget(B, S, LOGN);
    reduce_bound<B, (T < B ? T : B)>(x, c.q);
    if constexpr (S == LOGN - 1) {
        const u64 t = mul_shoup_lazy(y, raw, c.q);
        y = x + c.q2 - t;
        x = x + t;
    } else {
        const u64 t = mul_fpq(y, w, c.nq, c.cneg);
        y = x + c.q4 - t;
        x = x + t;
    }
}

// Inverse (Gentleman-Sande) butterfly; all values stay in [0, 4q): the sum is folded once by 4q, the
// product comes out of the approximate-quotient multiply below 4q.
__device__ __forceinline__ void gs_bfly(u64& x, u64& y, const twf_t& w, const qconst_t& c) {
    const u64 t = x + c.q4 - y;
    x = csub(x + y, c.q4);
    y = mul_fpq(t, w, c.nq, c.cneg);
}

// ---- FP64-pipe arithmetic for moduli below 2^50 -------------------------------------------------------------------
// The fmaheavy pipe binds the 64-bit butterflies above (a modular product is 3 IMAD.WIDE + 2 IMAD.HI + 4 IMAD); the FP64
// pipe of B200 issues DFMA at the IMAD rate and a DFMA multiplies 53-bit mantissas.  For q < 2^50 the whole butterfly
// fits there (profiles/microbench/fp64bfly.cu: 14.0 G butterflies/s/SM against 6.5 for the integer form):
//   * a value is an integer-valued double, |v| <= B q with B tracked at compile time (units of q/16 below);
//   * a twiddle is {w, w/q} with w centred in (-q/2, q/2) -- 16 bytes, same table shape as {w, w'};
//   * t = y w - rint(y w / q) q EXACTLY: h = y*w, l = fma(y, w, -h) (error-free product), k = rint(y * (w/q)) through the
//     1.5 * 2^52 trick, d = fma(-k, q, h) (exact: an integer below 2^53), t = d + l (exact).  6 FP64-pipe operations.
//     |k - y w / q| <= 1/2 + |y w / q| 2^-53, |y w / q| <= B q / 2 < B 2^49, hence |t| <= q (1/2 + B/16); the rounding
//     trick needs |y (w/q)| <= 2^51, i.e. B <= 4.  (Exhaustive-style CPU check of exactly these expressions -- IEEE fma
//     is bit-identical on both machines -- in tests/test_fp64_modmul.py.)
//   * folding a value back to |v| <= q/2 (+ rounding slack) is fma / add / fma.
// Results leave through fp_store as canonical residues, so the two arithmetics are interchangeable word for word.
struct fconst_t { double q, qinv; u64 qi; };
constexpr double kRint = 6755399441055744.0;                 // 1.5 * 2^52
__device__ __forceinline__ fconst_t make_fconst(u64 q) {
    const double qd = (double)(long long)q;
    return fconst_t{qd, __ddiv_rn(1.0, qd), q};
}
__device__ __forceinline__ double d_of(u64 v) { return __longlong_as_double((long long)v); }
__device__ __forceinline__ u64 u_of(double d) { return (u64)__double_as_longlong(d); }
__device__ __forceinline__ double fp_mul(double y, double w, double wq, const fconst_t& c) {
    const double h = __dmul_rn(y, w), l = __fma_rn(y, w, -h);
    const double k = __dadd_rn(__fma_rn(y, wq, kRint), -kRint);
    return __dadd_rn(__fma_rn(-k, c.q, h), l);
}
__device__ __forceinline__ double fp_fold(double x, const fconst_t& c) {
    const double k = __dadd_rn(__fma_rn(x, c.qinv, kRint), -kRint);
    return __fma_rn(-k, c.q, x);
}
// residue below 2^52 -> double (the pattern {x, 0x43300000} is 2^52 + x)
__device__ __forceinline__ double fp_load(u64 x) { return __dadd_rn(d_of(x | 0x4330000000000000ull), -4503599627370496.0); }
// integer-valued |r| < q -> canonical residue: through the integer pattern of r + 1.5 * 2^52 (so -0.0 is 0)
__device__ __forceinline__ u64 fp_store(double r, const fconst_t& c) {
    const long long i = __double_as_longlong(__dadd_rn(r, kRint)) - 0x4338000000000000ll;
    return (u64)(i + ((i >> 63) & (long long)c.qi));
}

// bounds in units of q/16
constexpr int kFpLim = 64;                                   // a multiplied value must stay within 4q
constexpr int kFpFolded = 9;                                 // after fp_fold: q/2 + rounding slack
constexpr int kFpIn = 32;                                    // transform inputs may be in [0, 2q)
__host__ __device__ constexpr int fp_t(int B) { return 8 + (B + 15) / 16; }           // |y w mod q| for |y| <= B
__host__ __device__ constexpr bool fp_fwd_fold(int B) { return B + fp_t(B) > kFpLim; }
__host__ __device__ constexpr int fp_fwd_bound_before(int s) {
    int B = kFpIn;
    for (int i = 0; i < s; ++i) B = (fp_fwd_fold(B) ? kFpFolded : B) + fp_t(B);
    return B;
}
// inverse network: both outputs of a stage enter the next one as either operand
__host__ __device__ constexpr bool fp_inv_fold(int B) { return 2 * B > kFpLim / 2; }
__host__ __device__ constexpr int fp_inv_bound_before(int i) {
    int B = kFpIn;
    for (int j = 0; j < i; ++j) {
        const int xb = fp_inv_fold(B) ? kFpFolded : 2 * B, yb = fp_t(2 * B);
        B = xb > yb ? xb : yb;
    }
    return B;
}

// bound of the values a forward network of 2^logn points leaves when its final fold is skipped
__host__ __device__ constexpr int fp_fwd_out_bound(int logn) { return fp_fwd_bound_before(logn - 1) + fp_t(fp_fwd_bound_before(logn - 1)); }
// product with a NON-constant multiplicand s in [0, q) (no precomputed s/q): the quotient comes from h * (1/q), two
// roundings instead of one, |k - y s / q| <= 1/2 + |y s / q| 2^-52.  Needs |y| <= q (callers fold first); |t| < q.
__device__ __forceinline__ double fp_mul_generic(double y, double s, const fconst_t& c) {
    const double h = __dmul_rn(y, s), l = __fma_rn(y, s, -h);
    const double k = __dadd_rn(__fma_rn(h, c.qinv, kRint), -kRint);
    return __dadd_rn(__fma_rn(-k, c.q, h), l);
}

// Cooley-Tukey butterfly at global stage S on doubles; x, y carry the bit patterns
template <int LOGN, int S>
__device__ __forceinline__ void ct_bfly_fp(u64& x, u64& y, tw_t w, const fconst_t& c) {
    constexpr int B = fp_fwd_bound_before(S);
    static_assert(B <= kFpLim, "forward bound schedule");
    double X = d_of(x);
    if constexpr (fp_fwd_fold(B) && S != LOGN - 1) X = fp_fold(X, c);          // the last stage is folded on the way out
    const double t = fp_mul(d_of(y), d_of(w.x), d_of(w.y), c);
    y = u_of(__dadd_rn(X, -t));
    x = u_of(__dadd_rn(X, t));
}
// Gentleman-Sande butterfly at inverse stage I (global bit I)
template <int LOGN, int I>
__device__ __forceinline__ void gs_bfly_fp(u64& x, u64& y, tw_t w, const fconst_t& c) {
    constexpr int B = fp_inv_bound_before(I);
    static_assert(2 * B <= kFpLim, "inverse bound schedule");
    const double X = d_of(x), Y = d_of(y);
    double sum = __dadd_rn(X, Y);
    if constexpr (fp_inv_fold(B)) sum = fp_fold(sum, c);
    x = u_of(sum);
    y = u_of(fp_mul(__dadd_rn(X, -Y), d_of(w.x), d_of(w.y), c));
}
// last inverse stage: both outputs multiplied (n^-1 and w * n^-1), canonical residues out
template <int LOGN>
__device__ __forceinline__ void gs_last_fp(u64& x, u64& y, tw_t ninv, tw_t w1n, const fconst_t& c) {
    static_assert(2 * fp_inv_bound_before(LOGN - 1) <= kFpLim, "inverse bound schedule");
    const double X = d_of(x), Y = d_of(y);
    x = fp_store(fp_mul(__dadd_rn(X, Y), d_of(ninv.x), d_of(ninv.y), c), c);
    y = fp_store(fp_mul(__dadd_rn(X, -Y), d_of(w1n.x), d_of(w1n.y), c), c);
}

template <> struct arith_const<0> { typedef qconst_t type; static __device__ __forceinline__ type make(u64 q) { return make_qconst(q); } };
template <> struct arith_const<1> { typedef fconst_t type; static __device__ __forceinline__ type make(u64 q) { return make_fconst(q); } };

// generic (non-constant multiplicand) product, Barrett on the 128-bit product.
// mu = floor(2^128 / q) as {lo, hi}; result canonical.
__device__ __forceinline__ u64 mul_barrett(u64 a, u64 b, u64 q, u64 mu_lo, u64 mu_hi) {
    u64 lo = a * b, hi = __umul64hi(a, b);
    // q_hat = floor((hi:lo) * mu / 2^128), computed from the three high partial products
    u64 t1 = __umul64hi(lo, mu_lo);
    u64 m_lo = lo * mu_hi, m_hi = __umul64hi(lo, mu_hi);
    u64 s = t1 + m_lo; u64 c = s < t1;
    u64 n_lo = hi * mu_lo, n_hi = __umul64hi(hi, mu_lo);
    u64 s2 = s + n_lo; c += s2 < s;
    u64 qhat = hi * mu_hi + m_hi + n_hi + c;
    u64 r = lo - qhat * q;
    r = csub(r, 2 * q);
    return csub(r, q);
}

}  // namespace mfhe

// ---- helpers added for the matrix pipeline --------------------------------------------------
namespace mfhe {

// (hi:lo) mod q for any 128-bit value, mu = floor(2^128 / q) as {lo, hi}; canonical result.
__device__ __forceinline__ u64 reduce128(u64 lo, u64 hi, u64 q, u64 mu_lo, u64 mu_hi) {
    u64 t1 = __umul64hi(lo, mu_lo);
    u64 m_lo = lo * mu_hi, m_hi = __umul64hi(lo, mu_hi);
    u64 s = t1 + m_lo; u64 c = s < t1;
    u64 n_lo = hi * mu_lo, n_hi = __umul64hi(hi, mu_lo);
    u64 s2 = s + n_lo; c += s2 < s;
    u64 qhat = hi * mu_hi + m_hi + n_hi + c;
    u64 r = lo - qhat * q;
    r = csub(r, 2 * q);
    return csub(r, q);
}
// One-word Barrett for products of canonical residues (q < 2^62, k = bits of q):
//   mu1 = floor(2^(k+62) / q) = floor(2^128 / q) >> (66 - k),  qhat = mulhi((a b) >> (k-2), mu1) >= floor(a b / q) - 2
// (HAC 14.42: the two truncations cost at most a b / 2^(k+62) <= 1/2 and 2^(k-2) / q < 1/2), so r = a b - qhat q < 3q.
// Two 64x64 high products instead of the four of mul_barrett.
struct barrett1_t { u64 q, q2, mu1; int sh; };
__device__ __forceinline__ barrett1_t make_barrett1(u64 q, u64 mu_lo, u64 mu_hi) {
    const int k = 64 - __clzll((long long)q);
    return barrett1_t{q, 2 * q, (mu_hi << (k - 2)) | (mu_lo >> (66 - k)), k - 2};
}
__device__ __forceinline__ u64 mul_barrett1(u64 a, u64 b, const barrett1_t& c) {
    const u64 lo = a * b, hi = __umul64hi(a, b);
    const u64 x1 = (lo >> c.sh) | (hi << (64 - c.sh));
    const u64 r = lo - __umul64hi(x1, c.mu1) * c.q;
    return csub(csub(r, c.q2), c.q);
}
__device__ __forceinline__ u64 add_mod(u64 a, u64 b, u64 q) { return csub(a + b, q); }
__device__ __forceinline__ u64 sub_mod(u64 a, u64 b, u64 q) { return csub(a + q - b, q); }
__device__ __forceinline__ u64 neg_mod(u64 a, u64 q) { return a ? q - a : 0; }
// x mod q for any 64-bit x
__device__ __forceinline__ u64 reduce64(u64 x, u64 q, u64 mu_hi) {
    // floor(x * floor(2^128/q) / 2^128) >= floor(x/q) - 1 using only the high word of mu
    u64 qhat = __umul64hi(x, mu_hi);
    u64 r = x - qhat * q;
    r = csub(r, 2 * q);
    return csub(r, q);
}
// signed value -> canonical residue (the reference's `v % q; if (v < 0) v += q`)
__device__ __forceinline__ u64 signed_mod(long long v, u64 q, u64 mu_hi) {
    u64 a = v < 0 ? (u64)(-(v + 1)) + 1ull : (u64)v;     // |v| without overflow at INT64_MIN
    u64 r = reduce64(a, q, mu_hi);
    return (v < 0 && r) ? q - r : r;
}

}  // namespace mfhe
