// encoder.cuh -- single-lane XY canonical embedding and exact CRT helpers under the reference's
// names (include/core/encoder.cuh:10-65).  The three __global__ entry points are part of the
// reference's public header and are kept launchable; the class forwards to the batched kernels
// of libmfhe_b200.so with lanes = 1.
#pragma once
#include <cstdint>
#include <vector>
#include <cuComplex.h>
#include <cuda_runtime.h>

namespace matrix_fhe {

__global__ void mat_mul_kernel_complex(const cuDoubleComplex* A, const cuDoubleComplex* B, cuDoubleComplex* C, int n);
__global__ void crt_compose_centerlift_kernel(const uint64_t* d_in_rns, int64_t* d_out_centered, int n2, int limbs);
__global__ void dequantize_exact_kernel(const uint64_t* d_real_rns, const uint64_t* d_imag_rns, cuDoubleComplex* d_out,
                                        int n2, double scaling_factor, int limbs);

void crt_compose_centerlift_big(const uint64_t* d_in_rns, uint64_t* d_out_mag, uint8_t* d_out_neg, int n2, int limbs,
                                cudaStream_t stream = 0);

class Encoder {
public:
    int n;
    cuDoubleComplex* d_V_cx;         // V[j][k] = zeta^(5^j k)
    cuDoubleComplex* d_V_cx_T;
    cuDoubleComplex* d_V_inv_cx;     // conj(V)^T / n
    cuDoubleComplex* d_V_inv_cx_T;

    Encoder(int n);
    ~Encoder();
    void idft2(const cuDoubleComplex* d_eval_xy, cuDoubleComplex* d_coeff_xy);
    void encode(const cuDoubleComplex* d_msg, uint64_t* d_real_rns, uint64_t* d_imag_rns);
    void decode_from_eval_complex(const cuDoubleComplex* d_eval, cuDoubleComplex* d_msg);
    void decode_lane_from_rns_eval(const uint64_t* d_real_rns, const uint64_t* d_imag_rns, cuDoubleComplex* d_msg);

private:
    void init_complex_matrices();
};

}  // namespace matrix_fhe
