// batched_trace.cuh -- the trace step over a batch of lanes (include/core/batched_trace.cuh:18-49).
// Data: [batch][limb][n][n].
#pragma once
#include <cstdint>

namespace matrix_fhe {

void map_B_to_Bprime_batched(const uint64_t* B_real, const uint64_t* B_imag, uint64_t* Bp_real, uint64_t* Bp_imag, int n,
                             int rns_limbs, int batch_size);
void trace_gemm_batched(const uint64_t* A_real, const uint64_t* A_imag, const uint64_t* Bp_real, const uint64_t* Bp_imag,
                        uint64_t* C_real, uint64_t* C_imag, int n, int rns_limbs, int batch_size);
void rescale_by_delta_batched(uint64_t* C_real, uint64_t* C_imag, int n, int rns_limbs, int batch_size, uint64_t inv0,
                              uint64_t inv1, uint64_t inv2);

}  // namespace matrix_fhe
