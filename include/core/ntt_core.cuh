// ntt_core.cuh -- X-axis transforms under the reference's names (include/core/ntt_core.cuh:8-53).
// Three conventions, each ONE launch for the whole batch:
//   *_phantom : X^n + 1, minimal primitive 2n-th root, natural in -> bit-reversed out
//   *_gl      : X^n - i, A[k] = sum_j a_j psi^j omega^(jk), natural order both ways (tmp is unused)
//   custom_*  : X^n - 1 (cyclic), natural order both ways
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace matrix_fhe {

// kept for source compatibility; the tables themselves live inside libmfhe_b200.so
struct NTTTable {
    uint64_t* d_psi_powers;
    uint64_t* d_psi_inv_powers;
    uint64_t* d_twist_powers;
    uint64_t* d_twist_inv_powers;
    uint64_t* d_n_inv;
    int modulus_count;
    int n;
};

void init_ntt_moduli_manual(const uint64_t* h_moduli);
void init_ntt_tables_manual(int n, int limbs);
void init_gl_twist_tables(int n, int limbs);
void init_gl_perm_tables(int n);
const NTTTable& get_manual_ntt_table();
const uint32_t* get_gl_perm();
const uint32_t* get_gl_inv_perm();

void apply_gl_perm(const uint64_t* in, uint64_t* out, int limbs, int batch_count, int n, bool inverse, cudaStream_t stream = 0);

void xy_ntt_forward_phantom(uint64_t* data, int limbs, int batch_count, int n, cudaStream_t stream = 0);
void xy_ntt_backward_phantom(uint64_t* data, int limbs, int batch_count, int n, cudaStream_t stream = 0);
void xy_ntt_forward_gl(uint64_t* data, uint64_t* tmp, int limbs, int batch_count, int n, cudaStream_t stream = 0);
void xy_ntt_backward_gl(uint64_t* data, uint64_t* tmp, int limbs, int batch_count, int n, cudaStream_t stream = 0);
void custom_ntt_forward(uint64_t* data, int limbs, int batch_count, int n, cudaStream_t stream = 0);
void custom_ntt_backward(uint64_t* data, int limbs, int batch_count, int n, cudaStream_t stream = 0);

}  // namespace matrix_fhe
