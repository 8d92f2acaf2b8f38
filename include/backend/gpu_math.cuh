// gpu_math.cuh -- Gaussian integer a + b i held as RNS residues, array-of-structures
// (reference include/backend/gpu_math.cuh:11-91: struct GaussianIntRNS with add / mul / conj / mul_minus_i).
#pragma once
#include <cuda_runtime.h>
#include "../core/config.h"

namespace matrix_fhe {

struct __align__(16) GaussianIntRNS {
    uint64_t x[RNS_NUM_LIMBS];   // real part, one residue per limb
    uint64_t y[RNS_NUM_LIMBS];   // imaginary part

    __device__ static inline uint64_t add_mod(uint64_t a, uint64_t b, uint64_t q) { const uint64_t s = a + b; return s >= q ? s - q : s; }
    __device__ static inline uint64_t sub_mod(uint64_t a, uint64_t b, uint64_t q) { return a >= b ? a - b : a + q - b; }
    __device__ static inline uint64_t mul_mod(uint64_t a, uint64_t b, uint64_t q) { return (uint64_t)(((unsigned __int128)a * b) % q); }
    __device__ static inline uint64_t neg_mod(uint64_t a, uint64_t q) { return a ? q - a : 0; }

    __device__ static GaussianIntRNS add(const GaussianIntRNS& A, const GaussianIntRNS& B, const uint64_t* moduli) {
        GaussianIntRNS r;
#pragma unroll
        for (int l = 0; l < RNS_NUM_LIMBS; ++l) { r.x[l] = add_mod(A.x[l], B.x[l], moduli[l]); r.y[l] = add_mod(A.y[l], B.y[l], moduli[l]); }
        return r;
    }
    // (a + bi)(c + di) = (ac - bd) + (ad + bc) i
    __device__ static GaussianIntRNS mul(const GaussianIntRNS& A, const GaussianIntRNS& B, const uint64_t* moduli) {
        GaussianIntRNS r;
#pragma unroll
        for (int l = 0; l < RNS_NUM_LIMBS; ++l) {
            const uint64_t q = moduli[l];
            r.x[l] = sub_mod(mul_mod(A.x[l], B.x[l], q), mul_mod(A.y[l], B.y[l], q), q);
            r.y[l] = add_mod(mul_mod(A.x[l], B.y[l], q), mul_mod(A.y[l], B.x[l], q), q);
        }
        return r;
    }
    __device__ static GaussianIntRNS conj(const GaussianIntRNS& A, const uint64_t* moduli) {
        GaussianIntRNS r;
#pragma unroll
        for (int l = 0; l < RNS_NUM_LIMBS; ++l) { r.x[l] = A.x[l]; r.y[l] = neg_mod(A.y[l], moduli[l]); }
        return r;
    }
    // -i (a + bi) = b - a i
    __device__ static GaussianIntRNS mul_minus_i(const GaussianIntRNS& A, const uint64_t* moduli) {
        GaussianIntRNS r;
#pragma unroll
        for (int l = 0; l < RNS_NUM_LIMBS; ++l) { r.x[l] = A.y[l]; r.y[l] = neg_mod(A.x[l], moduli[l]); }
        return r;
    }
};

}  // namespace matrix_fhe
