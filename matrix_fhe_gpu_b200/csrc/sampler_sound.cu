// sampler_sound.cu -- the cryptographically sound sampler option.
//
// The reference's samplers (src/core/HE.cu:564-627, 690-713) are hashes of the flat index with fixed constants: `a` is one
// LCG step of the index, the noise seed is a literal, the ternary key is a function of (w, x) only -- reproducible, which
// is what the parity tests need, and unusable as encryption randomness.  Behind mfhe_ctx_set_sampler(ctx, 1, seed) the
// same three streams come from Philox4x32-10 (Salmon et al., SC'11; the counter-based generator cuRAND and Phantom's
// ph/src/prng.cu's role -- that one is blake2-based), keyed by a 128-bit context seed:
//     counter = (index lo, index hi, draw number lo, stream tag | draw number hi),   key = (seed[0] ^ seed[1] folded)
// so every word is an independent function of (seed, stream, draw, index): no state, any launch shape, and a draw
// number that the context increments per sampled array, so two encryptions never share randomness.
//   uniform   128 random bits reduced mod q_l (bias < q / 2^128)
//   gaussian  discrete gaussian, sigma = 3.2, by inversion of a 64-bit cumulative table over z in [-41, 41] (tail
//             beyond 12.8 sigma has mass < 2^-118; table rounding bounds the statistical distance by ~2^-57); one draw
//             per logical coefficient [w][y][x], the same integer in every limb (as HE.cu:590-627 does)
//   ternary   uniform on {-1, 0, 1}: (r0 + r1 + r2 + r3) mod 3 == the 128-bit value mod 3 (2^32 = 1 mod 3), bias 2^-128
#include "ctx.h"

namespace mfhe {

__host__ __device__ inline void philox4x32_10(u32 c[4], u32 k0, u32 k1) {
    for (int r = 0; r < 10; ++r) {
        const u64 p0 = (u64)0xD2511F53u * c[0], p1 = (u64)0xCD9E8D57u * c[2];
        const u32 n0 = (u32)(p1 >> 32) ^ c[1] ^ k0, n1 = (u32)p1, n2 = (u32)(p0 >> 32) ^ c[3] ^ k1, n3 = (u32)p0;
        c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
}

struct SoundArgs { u64 seed0, seed1, draw; };
__device__ __forceinline__ void draw128(const SoundArgs& a, u32 tag, u64 index, u32 out[4]) {
    out[0] = (u32)index; out[1] = (u32)(index >> 32); out[2] = (u32)a.draw; out[3] = (tag << 24) | ((u32)(a.draw >> 32) & 0xFFFFFFu);
    // a 128-bit seed enters as key (low halves) and as a pre-whitening of the counter (high halves)
    out[0] ^= (u32)(a.seed1); out[1] ^= (u32)(a.seed1 >> 32);
    philox4x32_10(out, (u32)a.seed0, (u32)(a.seed0 >> 32));
}

constexpr int kCdtHalf = 41, kCdtSize = 2 * kCdtHalf + 1;
struct Cdt { u64 c[kCdtSize]; };           // c[i] = floor(2^64 * P(Z <= i - 41)) (last entry saturated)

template <int KIND>
__global__ void sound_kernel(u64* __restrict__ out, const u64* __restrict__ moduli, const ulonglong2* __restrict__ mu, int n,
                             int limbs, size_t total, SoundArgs a, const __grid_constant__ Cdt cdt) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    u32 r[4];
    if (KIND == 2) {                                   // ternary [w][limb][x]: one draw per (w, x), every limb
        const size_t w = idx / n, x = idx % n;
        draw128(a, 3u, idx, r);
        const u32 t = (u32)(((u64)r[0] + r[1] + r[2] + r[3]) % 3u);
        for (int l = 0; l < limbs; ++l) out[(w * limbs + l) * n + x] = t == 0 ? 0 : t == 1 ? 1 : moduli[l] - 1;
        return;
    }
    const size_t n2 = (size_t)n * n, per = (size_t)limbs * n2;
    if (KIND == 0) {                                   // uniform [w][limb][y][x]
        const size_t limb = (idx % per) / n2;
        draw128(a, 1u, idx, r);
        const u64 lo = ((u64)r[1] << 32) | r[0], hi = ((u64)r[3] << 32) | r[2];
        out[idx] = reduce128(lo, hi, moduli[limb], mu[limb].x, mu[limb].y);
    } else {                                           // gaussian: idx over logical coefficients [w][pos]
        const size_t w = idx / n2, pos = idx - w * n2;
        draw128(a, 2u, idx, r);
        const u64 u = ((u64)r[1] << 32) | r[0];
        int lo_i = 0, hi_i = kCdtSize - 1;             // smallest i with u < c[i]
        while (lo_i < hi_i) { const int mid = (lo_i + hi_i) >> 1; if (u < cdt.c[mid]) hi_i = mid; else lo_i = mid + 1; }
        const int z = lo_i - kCdtHalf;
        for (int l = 0; l < limbs; ++l) out[w * per + (size_t)l * n2 + pos] = z >= 0 ? (u64)z : moduli[l] - (u64)(-z);
    }
}

static Cdt make_cdt() {
    Cdt t;
    long double w[kCdtSize], sum = 0;
    for (int i = 0; i < kCdtSize; ++i) { const long double z = i - kCdtHalf; w[i] = expl(-z * z / (2.0L * 3.2L * 3.2L)); sum += w[i]; }
    long double acc = 0;
    for (int i = 0; i < kCdtSize; ++i) {
        acc += w[i] / sum;
        const long double v = ldexpl(acc, 64);
        t.c[i] = (i == kCdtSize - 1 || v >= 18446744073709551615.0L) ? ~0ull : (u64)v;
    }
    return t;
}

cudaError_t sound_sample_launch(int kind, u64* out, const u64* moduli, const ulonglong2* mu, int n, int limbs, int phi,
                                u64 seed0, u64 seed1, u64 draw, cudaStream_t st) {
    static const Cdt cdt = make_cdt();
    const size_t total = kind == 2 ? (size_t)phi * n : kind == 1 ? (size_t)phi * n * n : (size_t)phi * limbs * n * n;
    const unsigned grid = (unsigned)((total + 255) / 256);
    const SoundArgs a{seed0, seed1, draw};
    if (kind == 0) sound_kernel<0><<<grid, 256, 0, st>>>(out, moduli, mu, n, limbs, total, a, cdt);
    else if (kind == 1) sound_kernel<1><<<grid, 256, 0, st>>>(out, moduli, mu, n, limbs, total, a, cdt);
    else sound_kernel<2><<<grid, 256, 0, st>>>(out, moduli, mu, n, limbs, total, a, cdt);
    return cudaGetLastError();
}

void sound_cdt_copy(u64* h_out) { const Cdt t = make_cdt(); for (int i = 0; i < kCdtSize; ++i) h_out[i] = t.c[i]; }

}  // namespace mfhe
