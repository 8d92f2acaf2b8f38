// HE.cuh -- scheme-level API of the matrix FHE hot path, signature-compatible with the reference's
// include/core/HE.cuh:11-112 so that its drivers (src/main.cu, test/*.cu) compile unchanged.
// Every function forwards to libmfhe_b200.so through the C ABI (include/mfhe.h) using one
// process-global context created on first use (the reference keeps file-static singletons,
// src/core/HE.cu:21-44).  Pointers are device pointers owned by the caller; work is asynchronous on
// the given stream (legacy default stream when none is given); a failing call prints to stderr and
// exit(1)s, like the reference's fatal paths.
#pragma once
#include <cstdint>
#include <vector>
#include <cuComplex.h>
#include <cuda_runtime.h>
#include "config.h"

class DNTTTable;   // opaque here; Phantom's class in the reference

namespace matrix_fhe {

// (b, a) of T = phi * L * n * n words each, matrix-major [w][limb][y][x]; a starts at data + T.
struct RLWECiphertext {
    uint64_t* data = nullptr;
    int num_limbs = 0;
    bool is_ntt = false;
};
// [w][limb][x], W-eval + X-NTT domain (Phantom bit-reversed order)
struct SecretKey {
    uint64_t* data;
    int num_limbs;
};

void init_he_backend();
void copy_device_moduli(uint64_t* h_out, int count);
const DNTTTable& get_ntt_table();       // aborts, as the reference does (HE.cu:424-427)
const DNTTTable& get_xy_ntt_table();

// W-CRT over Phi_p(W)
void wntt_forward_matrix(const uint64_t* in, uint64_t* out, int n, int limbs, int phi, cudaStream_t stream = 0);
void wntt_inverse_matrix(const uint64_t* in_eval, uint64_t* out_coeff, int n, int limbs, int phi, cudaStream_t stream = 0);
void wntt_forward_centered(const int64_t* in_coeff_centered, int64_t* out_eval_centered, int n, int phi, cudaStream_t stream = 0);
void wntt_inverse_centered(const int64_t* in_eval_centered, int64_t* out_coeff_centered, int n, int phi, cudaStream_t stream = 0);
// complex W-DFT on split planes
void wdft_forward_centered_pair(const int64_t* in_re_centered, const int64_t* in_im_centered, double* out_re_eval,
                                double* out_im_eval, int n, int phi, cudaStream_t stream = 0);
void wdft_inverse_pair(const double* in_re_eval, const double* in_im_eval, double* out_re_coeff, double* out_im_coeff,
                       int n, int phi, cudaStream_t stream = 0);

void allocate_ciphertext(RLWECiphertext& ct, int limbs);
void free_ciphertext(RLWECiphertext& ct);
void generate_secret_key(SecretKey& sk, int limbs);

void encrypt(const uint64_t* message_coeffs, const SecretKey& sk, RLWECiphertext& ct);
void encrypt_pair(const uint64_t* msg_re, const uint64_t* msg_im, const SecretKey& sk, RLWECiphertext& ct_re,
                  RLWECiphertext& ct_im);
void decrypt_and_decode(const RLWECiphertext& ct_re, const RLWECiphertext& ct_im, const SecretKey& sk,
                        cuDoubleComplex* output_msg);
void decrypt_to_eval_matrix(const RLWECiphertext& ct, const SecretKey& sk, uint64_t* out_eval_matrix);

void add_ciphertexts(const RLWECiphertext& ct1, const RLWECiphertext& ct2, RLWECiphertext& res);
void multiply_ciphertexts_raw(const RLWECiphertext& ct1, const RLWECiphertext& ct2, uint64_t* d0, uint64_t* d1, uint64_t* d2);

}  // namespace matrix_fhe
