// phantom_math.cuh -- Shoup companion word and Shoup product (reference include/backend/phantom_math.cuh:9-28).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace phantom_math {

// floor(operand * 2^64 / modulus)
__host__ __device__ inline uint64_t compute_shoup(uint64_t operand, uint64_t modulus) {
    return (uint64_t)((((unsigned __int128)operand) << 64) / modulus);
}
// a * b mod q given b_shoup = compute_shoup(b, q); canonical result
__device__ inline uint64_t multiply_and_reduce_shoup(uint64_t a, uint64_t b, uint64_t b_shoup, uint64_t q) {
    const uint64_t r = a * b - __umul64hi(a, b_shoup) * q;
    return r >= q ? r - q : r;
}

}  // namespace phantom_math
