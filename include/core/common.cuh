// common.cuh -- device-side modular helpers under the reference's names
// (include/core/common.cuh:12-39).  Self-contained: no Phantom include.  Results are canonical.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace matrix_fhe {

static __device__ __forceinline__ uint64_t add_mod(uint64_t a, uint64_t b, uint64_t q) {
    const uint64_t s = a + b, t = s - q;
    return t + (t >> 63) * q;                 // one conditional subtraction, as Phantom's csub_q
}
static __device__ __forceinline__ uint64_t sub_mod(uint64_t a, uint64_t b, uint64_t q) {
    const uint64_t t = a - b;                 // top bit set <=> a < b (operands below 2^63)
    return t + (t >> 63) * q;
}
static __device__ __forceinline__ uint64_t mul_mod(uint64_t a, uint64_t b, uint64_t q) {
    return (uint64_t)(((unsigned __int128)a * b) % q);
}
static __device__ __forceinline__ uint64_t pow_mod(uint64_t base, uint64_t exp, uint64_t q) {
    uint64_t r = 1;
    for (base %= q; exp; exp >>= 1) {
        if (exp & 1) r = mul_mod(r, base, q);
        base = mul_mod(base, base, q);
    }
    return r;
}
static __device__ __forceinline__ uint64_t inv_mod(uint64_t n, uint64_t q) { return pow_mod(n, q - 2, q); }

}  // namespace matrix_fhe
