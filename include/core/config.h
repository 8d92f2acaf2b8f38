// config.h -- parameter contract of the matrix scheme (drop-in for the reference's
// include/core/config.h:7-52; same names, same values, nothing else).
// Ring: Z_Q[X,Y,W] / (X^n - i, Y^n - i, Phi_p(W)),  n = MATRIX_N,  p = BATCH_PRIME_P = 3 * 257.
#pragma once
#include <cstdint>

namespace matrix_fhe {

static constexpr int LOG_N = 16;
static constexpr int HE_N = 1 << LOG_N;

static constexpr int MATRIX_N = 64;
static constexpr int BATCH_PRIME_P = 771;
static constexpr int BATCH_SIZE = 512;                 // Euler phi(771)
static constexpr int PACK_N = MATRIX_N * BATCH_SIZE;
static constexpr int POLY_N = PACK_N;

static constexpr int RNS_NUM_LIMBS = 11;
static constexpr int P_NUM_LIMBS = 3;
static constexpr double SCALING_FACTOR = 34359738368.0;   // 2^35

// every modulus is 1 mod lcm(4 * 64, 771) = 197376: one 45-bit prime, then ten 35-bit primes
static constexpr uint64_t RNS_MODULI[RNS_NUM_LIMBS] = {
    17592186435073ULL, 17182765057ULL, 17184541441ULL, 17186120449ULL, 17186515201ULL, 17186909953ULL,
    17188883713ULL,    17190462721ULL, 17190857473ULL, 17191844353ULL, 17192831233ULL};

// extension base reserved by the reference for key switching (55, 40, 40 bits)
static constexpr uint64_t P_MODULI[P_NUM_LIMBS] = {18014398515156481ULL, 549757491457ULL, 549759662593ULL};

}  // namespace matrix_fhe
