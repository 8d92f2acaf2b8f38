// trace.cuh -- single-lane trace step under the reference's names (include/core/trace.cuh:11-25).
// Data: [limb][n][n] real part and imaginary part, residues mod q_limb.
#pragma once
#include <cstdint>

namespace matrix_fhe {

// B'[(-j mod n)][k] = conj-and-twist of B[j][k]
void map_B_to_Bprime_Xinv_twist(const uint64_t* B_real, const uint64_t* B_imag, uint64_t* Bp_real, uint64_t* Bp_imag, int n,
                                int rns_limbs);
// C = n * A * B'^T (complex, per limb)
void trace_gemm_ABpT_rns(const uint64_t* A_real, const uint64_t* A_imag, const uint64_t* Bp_real, const uint64_t* Bp_imag,
                         uint64_t* C_real, uint64_t* C_imag, int n, int rns_limbs);
// C <- C * inv_l for limbs 0..2 (limbs >= 3 are multiplied by 0, as in the reference)
void rescale_by_delta_rns(uint64_t* C_real, uint64_t* C_imag, int n, int rns_limbs, uint64_t inv0, uint64_t inv1, uint64_t inv2);

}  // namespace matrix_fhe
