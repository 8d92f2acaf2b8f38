// scheme.cu -- integer kernels of encrypt / decrypt and the small ciphertext-level ops.
//
//   xrow_mul_s   the fused RLWE core: for every row (w, limb, y) of n coefficients
//                    t = INTT_X( NTT_X(a_row) (.) s[w][limb] ),   out = m - t + e   or   out = m + t
//                entirely in registers: one read of a, m (and e), one write of out.  Replaces, per
//                call, src/core/HE.cu:1508-1530 / 1579-1590 = D2D copy + 32 768 launches of Phantom's
//                64-point NTT (ntt_core.cu:443-450) + pointwise_mul_s_kernel (:509-531) + D2D copy +
//                32 768 inverse launches + combine_b_kernel / add_poly_kernel (:535-560) + layout passes.
//   samplers     uniform / gaussian / ternary index-hash streams (HE.cu:564-627, 690-713)
//   layout       matrix-major <-> poly-major (HE.cu:1330-1368)
//   ct ops       add_ct_kernel / mul_tensor_kernel (HE.cu:631-669)
//   trace step   trace.cu:30-161, batched_trace.cu:37-197
//   gl perm      ntt_core.cu:258-269
#include "ctx.h"
#include "ntt_api.h"
#include "ntt_small.cuh"

namespace mfhe {

__device__ __forceinline__ size_t row_offset(int layout, u32 w, u32 l, u32 y, int n, int L) {
    return layout == MFHE_LAYOUT_POLY ? (((size_t)w * n + y) * L + l) * n : (((size_t)w * L + l) * n + y) * n;
}

#ifndef MFHE_XROW_MINB
#define MFHE_XROW_MINB 4
#endif
// AR = 1: the two transforms and the key product run on the FP64 pipe (every modulus below 2^50, T carries the
// {w, w/q} tables; modmath.cuh) -- same words out, the reference's 45- and 35-bit moduli qualify
template <int LOGN, bool DUAL, int AR>
__global__ void __launch_bounds__(256, MFHE_XROW_MINB)
xrow_mul_s_kernel(RowArgs r, SmallTables T, const ulonglong2* __restrict__ mu_tab, size_t rows) {
    using Net = SmallNet<LOGN>;
    constexpr int TPP = Net::TPP, PPB = 256 / TPP, STRIDE = Net::STRIDE;
    __shared__ __align__(16) u64 sm_all[PPB * STRIDE];
    const u32 u = threadIdx.x % TPP, pib = threadIdx.x / TPP;
    const size_t rid = (size_t)blockIdx.x * PPB + pib;          // (w * L + l) * n + y
    if (rid >= rows) return;
    const u32 y = rid % r.n, wl = rid / r.n, l = wl % r.limbs, w = wl / r.limbs;
    const u64 q = T.moduli[l], q2 = 2 * q;
    const ulonglong2 mu = mu_tab[l];
    u64* sm = sm_all + pib * STRIDE;
    const size_t ao = row_offset(r.a_layout, w, l, y, r.n, r.limbs);
    // Two components may come with their own a (decrypt of an (re, im) pair: a1).  encrypt_pair stores the SAME a in
    // both ciphertexts (HE.cu:1538-1545), so the rows are compared while they are loaded and, when every row of the
    // warp agrees, t = INTT(NTT(a) (.) s) is computed once and used for both; any difference -> second pass over a1.
    bool shared_t = true;
    if constexpr (DUAL) {
        const u32 e0 = deposit3(u, Net::FP0);
        bool eq = true;
#pragma unroll
        for (int j = 0; j < 8; ++j) eq &= r.a[ao + e0 + (j << Net::FP0)] == r.a1[ao + e0 + (j << Net::FP0)];
        shared_t = __all_sync(__activemask(), eq);
    }
    const int passes = (DUAL && !shared_t) ? 2 : 1;       // DUAL = false: one pass, compile-time
    for (int pass = 0; pass < passes; ++pass) {
    u64 v[8];
    u32 e;
    if constexpr (AR) {
        const fconst_t fc = make_fconst(q);
        Net::forward_fp(v, ((DUAL && pass) ? r.a1 : r.a) + ao, sm, u, T.tw + ((size_t)l << LOGN), fc);
        const u64* s = r.s + ((size_t)wl << LOGN) + 8 * u;
        const ulonglong2 s0 = __ldg((const ulonglong2*)s), s1 = __ldg((const ulonglong2*)s + 1),
                         s2 = __ldg((const ulonglong2*)s + 2), s3 = __ldg((const ulonglong2*)s + 3);
        const u64 sv[8] = {s0.x, s0.y, s1.x, s1.y, s2.x, s2.y, s3.x, s3.y};
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = u_of(fp_mul_generic(fp_fold(d_of(v[j]), fc), fp_load(sv[j]), fc));
        e = Net::inverse_fp(v, sm, u, T.itw + ((size_t)l << LOGN), fc, T.ninv[l], T.w1n[l]);
    } else {
    Net::forward(v, ((DUAL && pass) ? r.a1 : r.a) + ao, sm, u, T.tw + ((size_t)l << LOGN), q, q2);
    {
        const u64* s = r.s + ((size_t)wl << LOGN) + 8 * u;      // key is stored in the same bit-reversed order
        const ulonglong2 s0 = __ldg((const ulonglong2*)s), s1 = __ldg((const ulonglong2*)s + 1),
                         s2 = __ldg((const ulonglong2*)s + 2), s3 = __ldg((const ulonglong2*)s + 3);
        const barrett1_t bc = make_barrett1(q, mu.x, mu.y);
        v[0] = mul_barrett1(v[0], s0.x, bc); v[1] = mul_barrett1(v[1], s0.y, bc);
        v[2] = mul_barrett1(v[2], s1.x, bc); v[3] = mul_barrett1(v[3], s1.y, bc);
        v[4] = mul_barrett1(v[4], s2.x, bc); v[5] = mul_barrett1(v[5], s2.y, bc);
        v[6] = mul_barrett1(v[6], s3.x, bc); v[7] = mul_barrett1(v[7], s3.y, bc);
    }
    e = Net::inverse(v, sm, u, T.itw + ((size_t)l << LOGN), q, q2, T.ninv[l], T.w1n[l]);
    }
#pragma unroll
    for (int comp = 0; comp < 2; ++comp) {
        if (passes == 2 && comp != pass) continue;
        const u64* m = comp ? r.m1 : r.m0;
        const u64* er = comp ? r.e1 : r.e0;
        u64* o = comp ? r.o1 : r.o0;
        if (!o) continue;
        const size_t mo = row_offset(r.m_layout, w, l, y, r.n, r.limbs), eo = row_offset(r.e_layout, w, l, y, r.n, r.limbs),
                     oo = row_offset(r.o_layout, w, l, y, r.n, r.limbs);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const u32 pos = e + (j << Net::PL);
            u64 x = m[mo + pos];
            if (r.subtract) { x = sub_mod(x, v[j], q); if (er) x = add_mod(x, er[eo + pos], q); }
            else x = add_mod(x, v[j], q);
            o[oo + pos] = x;
        }
    }
    if (pass + 1 < passes) __syncwarp();                       // the shared tile is reused by the second pass
    }
}

cudaError_t xrow_mul_s_launch(const RowArgs& r, const SmallTables& T, const ulonglong2* mu, cudaStream_t st) {
    const size_t rows = (size_t)r.phi * r.limbs * r.n;
#define MFHE_ROW_CASE(LN)                                                                       \
    case LN: {                                                                                  \
        constexpr int PPB = 256 / ((1 << LN) / 8);                                              \
        const unsigned grid = (unsigned)((rows + PPB - 1) / PPB);                               \
        if (T.arith) {                                                                          \
            if (r.a1 && r.o1) xrow_mul_s_kernel<LN, true, 1><<<grid, 256, 0, st>>>(r, T, mu, rows); \
            else xrow_mul_s_kernel<LN, false, 1><<<grid, 256, 0, st>>>(r, T, mu, rows);         \
        } else {                                                                                \
            if (r.a1 && r.o1) xrow_mul_s_kernel<LN, true, 0><<<grid, 256, 0, st>>>(r, T, mu, rows); \
            else xrow_mul_s_kernel<LN, false, 0><<<grid, 256, 0, st>>>(r, T, mu, rows);         \
        }                                                                                       \
        break; }
    switch (T.logn) {
        MFHE_ROW_CASE(4) MFHE_ROW_CASE(5) MFHE_ROW_CASE(6) MFHE_ROW_CASE(7) MFHE_ROW_CASE(8)
        default: return cudaErrorInvalidValue;
    }
#undef MFHE_ROW_CASE
    return cudaGetLastError();
}

// ---- fused decryption of the 4-component matrix product (mfhe_ct_matmul_decrypt, n = 64) ---------------------------
//   out[z][x] = R1[x][z] + (R2[., z] (*)_x s)[x],   R1 = D_bb + D_ba (*)_z s,   R2 = D_ab + D_aa (*)_z s      (re and im alike)
// One CTA per (lane, limb) matrix, a row of 64 per 8 threads as in xrow_mul_s.  The two inner convolutions run along the
// fast index of the product's [x][z] matrices and leave R1, R2 in shared memory; the outer one runs along x, i.e. down
// the COLUMNS of the R2 tile (stride ld in shared memory -- the transposes of the unfused path cost nothing here), and
// its result goes out as rows of [z][x].  Reads the 8 product poly sets once and writes 2, against 26 poly-set passes of
// the unfused sequence (3 row-kernel launches + 4 transposes); same words.
template <int AR>
__global__ void __launch_bounds__(512, 1)
matmul_decrypt_kernel(const u64* __restrict__ D, size_t T, const u64* __restrict__ sk, u64* __restrict__ out_re,
                      u64* __restrict__ out_im, SmallTables Tb, const ulonglong2* __restrict__ mu_tab, int limbs) {
    using Net = SmallNet<6>;
    constexpr int N = 64, LD = N + 1, STRIDE = Net::STRIDE;
    extern __shared__ __align__(16) u64 sm_md[];
    u64* R = sm_md;                                   // [4][64][LD]: R1 re, R1 im, R2 re, R2 im
    u64* snet = sm_md + 4 * N * LD;                   // [64][STRIDE]: the networks' exchange tiles, one per row
    const u32 u = threadIdx.x & 7, row = threadIdx.x >> 3;
    const u32 wl = blockIdx.x, l = wl % limbs;
    const size_t base = (size_t)wl * N * N;
    const u64 q = Tb.moduli[l], q2 = 2 * q;
    const ulonglong2 mu = mu_tab[l];
    u64* sm = snet + row * STRIDE;
    const tw_t* tw = Tb.tw + ((size_t)l << 6);
    const tw_t* itw = Tb.itw + ((size_t)l << 6);
    const tw_t ninv = Tb.ninv[l], w1n = Tb.w1n[l];
    // key words of this thread's network positions 8u .. 8u+7 (stored in the same bit-reversed order)
    u64 sv[8];
    {
        const u64* s = sk + ((size_t)wl << 6) + 8 * u;
        const ulonglong2 s0 = __ldg((const ulonglong2*)s), s1 = __ldg((const ulonglong2*)s + 1),
                         s2 = __ldg((const ulonglong2*)s + 2), s3 = __ldg((const ulonglong2*)s + 3);
        sv[0] = s0.x; sv[1] = s0.y; sv[2] = s1.x; sv[3] = s1.y; sv[4] = s2.x; sv[5] = s2.y; sv[6] = s3.x; sv[7] = s3.y;
    }
    // v <- INTT(NTT(src) (.) s), canonical, v[j] = coefficient  e + (j << PL)  (e returned)
    auto conv = [&](u64 (&v)[8], const u64* src, u32 stride) -> u32 {
        u32 e;
        if constexpr (AR) {
            const fconst_t fc = make_fconst(q);
            Net::forward_fp(v, src, sm, u, tw, fc, stride);
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = u_of(fp_mul_generic(fp_fold(d_of(v[j]), fc), fp_load(sv[j]), fc));
            e = Net::inverse_fp(v, sm, u, itw, fc, ninv, w1n);
        } else {
            Net::forward(v, src, sm, u, tw, q, q2, stride);
            const barrett1_t bc = make_barrett1(q, mu.x, mu.y);
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = mul_barrett1(v[j], sv[j], bc);
            e = Net::inverse(v, sm, u, itw, q, q2, ninv, w1n);
        }
        __syncwarp();                                  // the row's exchange tile is reused by the next transform
        return e;
    };
    // inner convolutions along z: k = 0, 1 -> R1 (re, im) from (D_bb, D_ba); k = 2, 3 -> R2 from (D_ab, D_aa)
#pragma unroll 1
    for (int k = 0; k < 4; ++k) {
        const u64* m = D + (size_t)k * T + base + (size_t)row * N;                         // D_bb re, im | D_ab re, im = sets 0..3
        const u64* a = D + (size_t)(4 + k) * T + base + (size_t)row * N;                   // D_ba re, im | D_aa re, im = sets 4..7
        u64 v[8];
        const u32 e = conv(v, a, 1);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const u32 pos = e + (j << Net::PL);
            R[(k * N + row) * LD + pos] = add_mod(m[pos], v[j], q);
        }
    }
    __syncthreads();
    // outer convolution along x, this thread group's row is now z = row: column z of R2, added to column z of R1
#pragma unroll 1
    for (int ri = 0; ri < 2; ++ri) {
        u64 v[8];
        const u32 e = conv(v, R + (size_t)(2 + ri) * N * LD + row, LD);
        u64* o = (ri ? out_im : out_re) + base + (size_t)row * N;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const u32 pos = e + (j << Net::PL);
            o[pos] = add_mod(R[(ri * N + pos) * LD + row], v[j], q);
        }
    }
}

cudaError_t matmul_decrypt_launch(const u64* D, size_t T, const u64* sk, u64* out_re, u64* out_im, const SmallTables& Tb,
                                  const ulonglong2* mu, int phi, int limbs, cudaStream_t st) {
    constexpr int smem = (4 * 64 * 65 + 64 * SmallNet<6>::STRIDE) * (int)sizeof(u64);
    cudaError_t e;
    if (Tb.arith) {
        e = cudaFuncSetAttribute(matmul_decrypt_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        matmul_decrypt_kernel<1><<<(unsigned)(phi * limbs), 512, smem, st>>>(D, T, sk, out_re, out_im, Tb, mu, limbs);
    } else {
        e = cudaFuncSetAttribute(matmul_decrypt_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        matmul_decrypt_kernel<0><<<(unsigned)(phi * limbs), 512, smem, st>>>(D, T, sk, out_re, out_im, Tb, mu, limbs);
    }
    return cudaGetLastError();
}

// ---- samplers (stand-alone form; encrypt fuses them into the W-CRT loader, wcrt.cu) -------------------
__device__ __forceinline__ u64 splitmix64_s(u64 x) {
    x += 0x9E3779B97F4A7C15ULL;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
    return x ^ (x >> 31);
}
template <int KIND>
__global__ void sample_kernel(u64* __restrict__ out, const u64* __restrict__ moduli, const ulonglong2* __restrict__ mu,
                              int n, int limbs, size_t total) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    if (KIND == 2) {                                   // ternary, [w][limb][x]
        const size_t per = (size_t)limbs * n, off = idx % per, limb = off / n, coeff = off - limb * n, poly = idx / per;
        const u64 t = (u64)poly * 1315423911ULL + (u64)coeff * 2654435761ULL;
        const int r = (int)((t * 11400714819323198485ULL) % 3ULL);
        out[idx] = r == 0 ? 0 : r == 1 ? 1 : moduli[limb] - 1;
        return;
    }
    const size_t n2 = (size_t)n * n, per = (size_t)limbs * n2;
    if (KIND == 0) {                                   // LCG of the flat index, reduced with Barrett (== seed % q)
        const size_t limb = (idx % per) / n2;
        u64 seed = 123456789ULL + idx;
        seed = seed * 6364136223846793005ULL + 1442695040888963407ULL;
        out[idx] = reduce64(seed, moduli[limb], mu[limb].y);
    } else {
        // gaussian: idx runs over the LOGICAL coefficients [w][pos]; the value is drawn once (two fp64 transcendentals)
        // and written to every limb, instead of being recomputed per limb
        const size_t w = idx / n2, pos = idx - w * n2;
        const u64 r1 = splitmix64_s(0xD6E8FEB86659FD93ULL ^ (u64)idx), r2 = splitmix64_s(r1);
        const double inv53 = 1.0 / 9007199254740992.0;
        const double u1 = ((double)(r1 >> 11) + 1.0) * inv53, u2 = ((double)(r2 >> 11) + 1.0) * inv53;
        const double mag = 3.2 * sqrt(-2.0 * log(u1));
        const long long z = llround(mag * cos(6.283185307179586 * u2));
        for (int limb = 0; limb < limbs; ++limb)
            out[w * per + (size_t)limb * n2 + pos] = z >= 0 ? (u64)z : moduli[limb] - (u64)(-z);
    }
}
cudaError_t sample_launch(int kind, u64* out, const u64* moduli, const ulonglong2* mu, int n, int limbs, int phi, cudaStream_t st) {
    const size_t total = kind == 2 ? (size_t)phi * limbs * n : kind == 1 ? (size_t)phi * n * n : (size_t)phi * limbs * n * n;
    const unsigned grid = (unsigned)((total + 255) / 256);
    if (kind == 0) sample_kernel<0><<<grid, 256, 0, st>>>(out, moduli, mu, n, limbs, total);
    else if (kind == 1) sample_kernel<1><<<grid, 256, 0, st>>>(out, moduli, mu, n, limbs, total);
    else sample_kernel<2><<<grid, 256, 0, st>>>(out, moduli, mu, n, limbs, total);
    return cudaGetLastError();
}

// ---- layout -------------------------------------------------------------------------------------------
__global__ void layout_kernel(const u64* __restrict__ in, u64* __restrict__ out, int n, int limbs, size_t total2,
                              int in_layout, int out_layout) {
    // four pairs of words per thread, the four loads in flight before the first store (one pair per thread: 66 % of HBM)
    const size_t i0 = (size_t)blockIdx.x * blockDim.x * 4 + threadIdx.x;
    ulonglong2 v[4]; size_t dst[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const size_t i2 = i0 + (size_t)k * blockDim.x;
        dst[k] = ~(size_t)0;
        if (i2 >= total2) continue;
        const size_t idx = i2 * 2;
        const u32 x = idx % n; const size_t t = idx / n;
        const u32 y = t % n, wl = t / n, l = wl % limbs, w = wl / limbs;
        v[k] = *(const ulonglong2*)(in + row_offset(in_layout, w, l, y, n, limbs) + x);
        dst[k] = row_offset(out_layout, w, l, y, n, limbs) + x;
    }
#pragma unroll
    for (int k = 0; k < 4; ++k)
        if (dst[k] != ~(size_t)0) *(ulonglong2*)(out + dst[k]) = v[k];
}
cudaError_t layout_launch(const u64* in, u64* out, int n, int limbs, int phi, int in_layout, int out_layout, cudaStream_t st) {
    const size_t total2 = (size_t)phi * limbs * n * n / 2;
    layout_kernel<<<(unsigned)((total2 + 1023) / 1024), 256, 0, st>>>(in, out, n, limbs, total2, in_layout, out_layout);
    return cudaGetLastError();
}

// ---- ciphertext add / tensor product -----------------------------------------------------------------
__device__ __forceinline__ int ct_limb(size_t idx, int n, int limbs, int ref_index) {
    // reference kernels: limb = (idx % (limbs*n)) / n, i.e. the POLY formula whatever the data layout
    return ref_index ? (int)((idx / n) % limbs) : (int)((idx / ((size_t)n * n)) % limbs);
}
__global__ void ct_add_kernel(const u64* __restrict__ c1, const u64* __restrict__ c2, u64* __restrict__ res,
                              const u64* __restrict__ moduli, size_t T, int n, int limbs, int ref_index) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= T) return;
    const u64 q = moduli[ct_limb(idx, n, limbs, ref_index)];
    // the reference's add_mod (common.cuh:12-14 -> Phantom add_uint64_uint64_mod) is (a+b) - q*(a+b >= q)
    res[idx] = csub(c1[idx] + c2[idx], q);
    res[T + idx] = csub(c1[T + idx] + c2[T + idx], q);
}
cudaError_t ct_add_launch(const u64* c1, const u64* c2, u64* res, const u64* moduli, size_t T, int n, int limbs,
                          int ref_index, cudaStream_t st) {
    ct_add_kernel<<<(unsigned)((T + 255) / 256), 256, 0, st>>>(c1, c2, res, moduli, T, n, limbs, ref_index);
    return cudaGetLastError();
}
__global__ void ct_mul_kernel(const u64* __restrict__ c1, const u64* __restrict__ c2, u64* __restrict__ d0,
                              u64* __restrict__ d1, u64* __restrict__ d2, const u64* __restrict__ moduli,
                              const ulonglong2* __restrict__ mu_tab, size_t T, int n, int limbs, int ref_index) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= T) return;
    const int l = ct_limb(idx, n, limbs, ref_index);
    const u64 q = moduli[l]; const ulonglong2 mu = mu_tab[l];
    const u64 b1 = c1[idx], a1 = c1[T + idx], b2 = c2[idx], a2 = c2[T + idx];
    if (q < (1ull << 50) && b1 < q && a1 < q && b2 < q && a2 < q) {
        // canonical operands of a modulus below 2^50: the four products on the FP64 pipe (exact, modmath.cuh), which leaves
        // this pass to the memory system (integer form: 56 % of HBM, multiplier-bound)
        const fconst_t fc = make_fconst(q);
        const double B1 = fp_load(b1), A1 = fp_load(a1), B2 = fp_load(b2), A2 = fp_load(a2);
        d0[idx] = fp_store(fp_mul_generic(B1, B2, fc), fc);
        d1[idx] = fp_store(fp_fold(__dadd_rn(fp_mul_generic(B1, A2, fc), fp_mul_generic(A1, B2, fc)), fc), fc);
        d2[idx] = fp_store(fp_mul_generic(A1, A2, fc), fc);
    } else if ((b1 | a1 | b2 | a2) < q || (b1 < q && a1 < q && b2 < q && a2 < q)) {
        const barrett1_t bc = make_barrett1(q, mu.x, mu.y);    // canonical operands: the one-word Barrett is exact
        d0[idx] = mul_barrett1(b1, b2, bc);
        d1[idx] = add_mod(mul_barrett1(b1, a2, bc), mul_barrett1(a1, b2, bc), q);
        d2[idx] = mul_barrett1(a1, a2, bc);
    } else {
        // reference_indexing picks the modulus by the POLY formula whatever the layout, so an operand may belong to a
        // wider limb than q: full 128-bit Barrett (the reference's u128 % accepts any operands)
        d0[idx] = mul_barrett(b1, b2, q, mu.x, mu.y);
        d1[idx] = add_mod(mul_barrett(b1, a2, q, mu.x, mu.y), mul_barrett(a1, b2, q, mu.x, mu.y), q);
        d2[idx] = mul_barrett(a1, a2, q, mu.x, mu.y);
    }
}
cudaError_t ct_mul_launch(const u64* c1, const u64* c2, u64* d0, u64* d1, u64* d2, const u64* moduli, const ulonglong2* mu,
                          size_t T, int n, int limbs, int ref_index, cudaStream_t st) {
    ct_mul_kernel<<<(unsigned)((T + 255) / 256), 256, 0, st>>>(c1, c2, d0, d1, d2, moduli, mu, T, n, limbs, ref_index);
    return cudaGetLastError();
}

// ---- trace step -----------------------------------------------------------------------------------------
// trace.cu:30-62 / batched_trace.cu:37-79:  B'[(-j mod n)][k] = (a, -b) for j = 0, (-b, -a) otherwise
// axis 0: the twisted index is the slow one (the reference's [X power][Y power] convention); axis 1: the fast one (x of the
// ciphertext layout [y][x]) -- the form the encrypted matrix product needs (mfhe_trace_map_x)
__global__ void trace_map_kernel(const u64* __restrict__ Br, const u64* __restrict__ Bi, u64* __restrict__ Bpr,
                                 u64* __restrict__ Bpi, const u64* __restrict__ moduli, int n, int limbs, size_t total, int axis) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const size_t n2 = (size_t)n * n, mat = idx / n2, pos = idx % n2;
    const int j = (int)(pos / n), k = (int)(pos % n), l = (int)(mat % limbs);
    const u64 q = moduli[l];
    const int t = axis ? k : j;
    const size_t dst = mat * n2 + (axis ? (size_t)j * n + ((n - k) & (n - 1)) : (size_t)((n - j) & (n - 1)) * n + k);
    const u64 a = Br[idx], b = Bi[idx];
    if (t == 0) { Bpr[dst] = a; Bpi[dst] = neg_mod(b, q); }
    else { Bpr[dst] = neg_mod(b, q); Bpi[dst] = neg_mod(a, q); }
}
cudaError_t trace_map_launch(const u64* Br, const u64* Bi, u64* Bpr, u64* Bpi, const u64* moduli, int n, int limbs,
                             int batch, cudaStream_t st, int axis) {
    const size_t total = (size_t)batch * limbs * n * n;
    trace_map_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(Br, Bi, Bpr, Bpi, moduli, n, limbs, total, axis);
    return cudaGetLastError();
}

// every n x n matrix of a [count][n][n] array transposed (shared-memory tile, both sides coalesced)
__global__ void __launch_bounds__(256)
transpose_kernel(const u64* __restrict__ in, u64* __restrict__ out, int n) {
    extern __shared__ u64 sm_tp[];
    const int ld = n + 1;
    const size_t base = (size_t)blockIdx.x * n * n;
    for (int i = threadIdx.x; i < n * n; i += blockDim.x) sm_tp[(i / n) * ld + (i % n)] = in[base + i];
    __syncthreads();
    for (int i = threadIdx.x; i < n * n; i += blockDim.x) out[base + i] = sm_tp[(i % n) * ld + (i / n)];
}
cudaError_t transpose_launch(const u64* in, u64* out, int n, size_t count, cudaStream_t st) {
    if (!count) return cudaSuccess;
    transpose_kernel<<<(unsigned)count, 256, (size_t)n * (n + 1) * sizeof(u64), st>>>(in, out, n);
    return cudaGetLastError();
}

// trace.cu:77-120 / batched_trace.cu:99-146:  C = n * A * B'^T (complex, mod q) for one (batch, limb) per CTA.
// Operands are staged transposed in shared memory ([t][row], padded), the four real sums
// sum ar*br, sum ai*bi, sum ar*bi, sum ai*br accumulate lazily in 128 bits (n * q^2 < 2^128) and are
// reduced once -- the same residues as the reference's per-term `u128 %` chain.
struct acc128 { u64 lo, hi; };
__device__ __forceinline__ void mac128(acc128& s, u64 a, u64 b) {
    const u64 lo = a * b, hi = __umul64hi(a, b);
    s.lo += lo; s.hi += hi + (s.lo < lo);
}
__global__ void __launch_bounds__(256)
trace_gemm_kernel(const u64* __restrict__ Ar, const u64* __restrict__ Ai, const u64* __restrict__ Bpr,
                  const u64* __restrict__ Bpi, u64* __restrict__ Cr, u64* __restrict__ Ci,
                  const u64* __restrict__ moduli, const ulonglong2* __restrict__ mu_tab, int n, int limbs, int tr) {
    extern __shared__ u64 sm_tr[];
    const int ld = n + 1;
    u64 *sAr = sm_tr, *sAi = sAr + (size_t)n * ld, *sBr = sAi + (size_t)n * ld, *sBi = sBr + (size_t)n * ld;
    const size_t n2 = (size_t)n * n, base = (size_t)blockIdx.x * n2;
    const int l = blockIdx.x % limbs;
    const u64 q = moduli[l]; const ulonglong2 mu = mu_tab[l];
    for (size_t i = threadIdx.x; i < n2; i += blockDim.x) {
        const int row = tr ? (int)(i % n) : (int)(i / n), t = tr ? (int)(i / n) : (int)(i % n);   // tr: operands are K-major
        sAr[t * ld + row] = Ar[base + i]; sAi[t * ld + row] = Ai[base + i];
        sBr[t * ld + row] = Bpr[base + i]; sBi[t * ld + row] = Bpi[base + i];
    }
    __syncthreads();
    const u64 nm = reduce64((u64)n, q, mu.y);
    for (size_t o = threadIdx.x; o < n2; o += blockDim.x) {
        const int row = (int)(o / n), col = (int)(o % n);
        acc128 rr = {0, 0}, ii = {0, 0}, ri = {0, 0}, ir = {0, 0};
        for (int t = 0; t < n; ++t) {
            const u64 ar = sAr[t * ld + row], ai = sAi[t * ld + row], br = sBr[t * ld + col], bi = sBi[t * ld + col];
            mac128(rr, ar, br); mac128(ii, ai, bi); mac128(ri, ar, bi); mac128(ir, ai, br);
        }
        const u64 vr = sub_mod(reduce128(rr.lo, rr.hi, q, mu.x, mu.y), reduce128(ii.lo, ii.hi, q, mu.x, mu.y), q);
        const u64 vi = add_mod(reduce128(ri.lo, ri.hi, q, mu.x, mu.y), reduce128(ir.lo, ir.hi, q, mu.x, mu.y), q);
        Cr[base + o] = mul_barrett(vr, nm, q, mu.x, mu.y);
        Ci[base + o] = mul_barrett(vi, nm, q, mu.x, mu.y);
    }
}
cudaError_t trace_gemm_launch(const u64* Ar, const u64* Ai, const u64* Bpr, const u64* Bpi, u64* Cr, u64* Ci,
                              const u64* moduli, const ulonglong2* mu, int n, int limbs, int batch, cudaStream_t st,
                              int transposed) {
    const size_t smem = 4 * (size_t)n * (n + 1) * sizeof(u64);
    if (smem > 200 * 1024) return cudaErrorInvalidValue;
    {   // set on every call: cheap, and right for any device / thread (no process-wide one-shot flag)
        cudaError_t e = cudaFuncSetAttribute(trace_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e != cudaSuccess) return e;
    }
    trace_gemm_kernel<<<(unsigned)(batch * limbs), 256, smem, st>>>(Ar, Ai, Bpr, Bpi, Cr, Ci, moduli, mu, n, limbs, transposed);
    return cudaGetLastError();
}

// trace.cu:132-161 / batched_trace.cu:163-197: limbs >= 3 are multiplied by 0, as in the reference
__global__ void trace_rescale_kernel(u64* __restrict__ Cr, u64* __restrict__ Ci, const u64* __restrict__ moduli,
                                     const ulonglong2* __restrict__ mu_tab, int n, int limbs, size_t total, u64 inv0,
                                     u64 inv1, u64 inv2) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const int l = (int)((idx / ((size_t)n * n)) % limbs);
    const u64 q = moduli[l]; const ulonglong2 mu = mu_tab[l];
    const u64 inv = l == 0 ? inv0 : l == 1 ? inv1 : l == 2 ? inv2 : 0;
    const u64 cr = Cr[idx], ci = Ci[idx];
    if (q < (1ull << 50) && inv < q && cr < q && ci < q) {      // canonical operands, modulus below 2^50: FP64-pipe products (exact)
        const fconst_t fc = make_fconst(q);
        const double K = fp_load(inv);
        Cr[idx] = fp_store(fp_mul_generic(fp_load(cr), K, fc), fc);
        Ci[idx] = fp_store(fp_mul_generic(fp_load(ci), K, fc), fc);
    } else if (inv < q && cr < q && ci < q) {                   // canonical operands: one-word Barrett
        const barrett1_t bc = make_barrett1(q, mu.x, mu.y);
        Cr[idx] = mul_barrett1(cr, inv, bc);
        Ci[idx] = mul_barrett1(ci, inv, bc);
    } else {
        Cr[idx] = mul_barrett(Cr[idx], inv, q, mu.x, mu.y);
        Ci[idx] = mul_barrett(Ci[idx], inv, q, mu.x, mu.y);
    }
}
// every limb multiplied by its own constant (the repaired form of the rescale step: the reference's kernel only knows
// three constants, a leftover of a 3-limb design -- SURVEY H7)
__global__ void trace_rescale_limbs_kernel(u64* __restrict__ Cr, u64* __restrict__ Ci, const u64* __restrict__ moduli,
                                           const ulonglong2* __restrict__ mu_tab, int n, int limbs, size_t total,
                                           const __grid_constant__ RescaleConsts k) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const int l = (int)((idx / ((size_t)n * n)) % limbs);
    const u64 q = moduli[l]; const ulonglong2 mu = mu_tab[l];
    const u64 cr = Cr[idx], ci = Ci[idx], kv = k.inv[l];
    if (q < (1ull << 50) && cr < q && ci < q && kv < q) {      // FP64-pipe products (exact) for moduli below 2^50
        const fconst_t fc = make_fconst(q);
        const double K = fp_load(kv);
        Cr[idx] = fp_store(fp_mul_generic(fp_load(cr), K, fc), fc);
        Ci[idx] = fp_store(fp_mul_generic(fp_load(ci), K, fc), fc);
    } else {
        Cr[idx] = mul_barrett(cr, kv, q, mu.x, mu.y);
        Ci[idx] = mul_barrett(ci, kv, q, mu.x, mu.y);
    }
}
cudaError_t trace_rescale_limbs_launch(u64* Cr, u64* Ci, const u64* moduli, const ulonglong2* mu, int n, int limbs, int batch,
                                       const RescaleConsts& k, cudaStream_t st) {
    const size_t total = (size_t)batch * limbs * n * n;
    trace_rescale_limbs_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(Cr, Ci, moduli, mu, n, limbs, total, k);
    return cudaGetLastError();
}

// RNS rescale by the last limb, final step: out[w][l][c] = (in[w][l][c] - out[w][l][c]) * q_last^-1 mod q_l, the input
// having `limbs` limbs per lane and the output limbs - 1 (out holds the W-CRT of the centred last-limb coefficients)
__global__ void rescale_combine_kernel(const u64* __restrict__ in, u64* __restrict__ out, const u64* __restrict__ moduli,
                                       const ulonglong2* __restrict__ mu_tab, int cols, int limbs, size_t total,
                                       const __grid_constant__ RescaleConsts k) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;       // over [w][limbs - 1][cols]
    if (idx >= total) return;
    const size_t c = idx % cols, wl = idx / cols, l = wl % (limbs - 1), w = wl / (limbs - 1);
    const u64 q = moduli[l]; const ulonglong2 mu = mu_tab[l];
    const u64 x = in[(w * limbs + l) * cols + c];
    out[idx] = mul_barrett(sub_mod(x, out[idx], q), k.inv[l], q, mu.x, mu.y);
}
cudaError_t rescale_combine_launch(const u64* in, u64* out, const u64* moduli, const ulonglong2* mu, int cols, int limbs, int phi,
                                   const RescaleConsts& k, cudaStream_t st) {
    const size_t total = (size_t)phi * (limbs - 1) * cols;
    rescale_combine_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(in, out, moduli, mu, cols, limbs, total, k);
    return cudaGetLastError();
}

cudaError_t trace_rescale_launch(u64* Cr, u64* Ci, const u64* moduli, const ulonglong2* mu, int n, int limbs, int batch,
                                 u64 inv0, u64 inv1, u64 inv2, cudaStream_t st) {
    const size_t total = (size_t)batch * limbs * n * n;
    trace_rescale_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(Cr, Ci, moduli, mu, n, limbs, total, inv0, inv1, inv2);
    return cudaGetLastError();
}

// ---- GL permutation: ntt_core.cu:258-269  out[perm[x]] = in[x] along x ---------------------------------
__global__ void gl_perm_kernel(const u64* __restrict__ in, u64* __restrict__ out, const uint32_t* __restrict__ perm,
                               int n, size_t total) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const size_t base = idx - idx % n;
    out[base + perm[idx % n]] = in[idx];
}
cudaError_t gl_perm_launch(const u64* in, u64* out, const uint32_t* perm, int limbs, int batch, int n, cudaStream_t st) {
    const size_t total = (size_t)batch * limbs * n;
    gl_perm_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(in, out, perm, n, total);
    return cudaGetLastError();
}

}  // namespace mfhe
