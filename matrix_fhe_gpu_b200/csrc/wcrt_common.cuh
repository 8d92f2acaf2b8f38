// wcrt_common.cuh -- addressing, the index-hash samplers and the single-element fused loader shared by the
// CUDA-core (wcrt.cu) and tensor-core (wcrt_tc.cu) W-CRT kernels.
#pragma once
#include "ctx.h"

namespace mfhe {

__device__ __forceinline__ size_t poly_set_addr(int layout, int w, int l, int c, int n, int cols, int L) {
    if (layout == MFHE_LAYOUT_POLY) {
        const int y = c / n, x = c - y * n;
        return (((size_t)w * n + y) * L + l) * n + x;
    }
    return ((size_t)w * L + l) * cols + c;
}

__device__ __forceinline__ u64 splitmix64_dev(u64 x) {
    x += 0x9E3779B97F4A7C15ULL;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
    return x ^ (x >> 31);
}
// src/core/HE.cu:590-627: one centred integer per logical coefficient [w][pos], same in every limb
__device__ __forceinline__ long long gaussian_noise_value(u64 coeff_id) {
    const u64 r1 = splitmix64_dev(0xD6E8FEB86659FD93ULL ^ coeff_id), r2 = splitmix64_dev(r1);
    const double inv53 = 1.0 / 9007199254740992.0;
    const double u1 = ((double)(r1 >> 11) + 1.0) * inv53, u2 = ((double)(r2 >> 11) + 1.0) * inv53;
    const double mag = 3.2 * sqrt(-2.0 * log(u1));
    return llround(mag * cos(6.283185307179586 * u2));
}
// src/core/HE.cu:564-578
__device__ __forceinline__ u64 uniform_value(u64 idx, u64 q, u64 mu_hi) {
    u64 seed = 123456789ULL + idx;
    seed = seed * 6364136223846793005ULL + 1442695040888963407ULL;
    return reduce64(seed, q, mu_hi);
}

// one canonical residue of input row i, column c (the sources of WcrtSource)
template <int SRC>
__device__ __forceinline__ u64 wcrt_load_one(const WcrtArgs& a, int i, int l, int c, u64 q, u64 mu_hi) {
    if constexpr (SRC == kSrcMem) {
        return __ldg((const u64*)a.in + poly_set_addr(a.in_layout, i, l, c, a.n, a.cols, a.limbs));
    } else if constexpr (SRC == kSrcI64) {
        return signed_mod(((const long long*)a.in)[(size_t)i * a.cols + c], q, mu_hi);
    } else if constexpr (SRC == kSrcQuantRe || SRC == kSrcQuantIm) {
        const double2 z = ((const double2*)a.in)[(size_t)i * a.cols + c];
        return signed_mod(llround((SRC == kSrcQuantRe ? z.x : z.y) * a.delta), q, mu_hi);
    } else if constexpr (SRC == kSrcUniform) {
        return uniform_value(((u64)i * a.limbs + l) * a.cols + c, q, mu_hi);
    } else {
        // kSrcGauss: a.in is the int8 table of centred noise values [phi][cols] (gauss_table_launch); the
        // reference draws ONE value per logical coefficient and uses it in every limb (HE.cu:590-627)
        const int z = ((const signed char*)a.in)[(size_t)i * a.cols + c];
        return z >= 0 ? (u64)z : q - (u64)(-z);
    }
}

}  // namespace mfhe
