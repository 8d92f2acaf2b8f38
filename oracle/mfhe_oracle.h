/*
 * mfhe_oracle.h -- CPU restatement of the reference's hot-path arithmetic.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may build, load or call it.  The product path
 * (matrix_fhe_gpu_b200/, include/) never links or falls back to this code.
 *
 * Each function cites the reference file:line it restates.  "ph/" abbreviates
 * extern/phantom-fhe/ inside the reference tree.
 *
 * Parity pinning: tests/test_oracle_*.py check this oracle against
 *   (i)  the known answers of the reference's own drivers
 *        (test/test_custom_ntt_roundtrip.cu, test/test_wcrt_roundtrip.cu),
 *   (ii) constants derived from the reference's algorithms in SURVEY.md 8(c),
 *   (iii) Phantom's own host table builder compiled from /root/reference into
 *        oracle/_ref (fixtures committed under tests/golden/).
 */
#ifndef MFHE_ORACLE_H
#define MFHE_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- modular arithmetic: include/core/common.cuh:12-39 ------------------ */
uint64_t mfo_addmod(uint64_t a, uint64_t b, uint64_t q);
uint64_t mfo_submod(uint64_t a, uint64_t b, uint64_t q);
uint64_t mfo_mulmod(uint64_t a, uint64_t b, uint64_t q);
uint64_t mfo_powmod(uint64_t b, uint64_t e, uint64_t q);
uint64_t mfo_invmod(uint64_t a, uint64_t q);
/* include/backend/phantom_math.cuh:9-28 */
uint64_t mfo_shoup(uint64_t w, uint64_t q);
uint64_t mfo_mulmod_shoup(uint64_t x, uint64_t w, uint64_t w_shoup, uint64_t q);
uint32_t mfo_bitrev(uint32_t x, int bits);
int      mfo_is_prime(uint64_t n);

/* ---- Phantom host number theory ----------------------------------------- */
/* ph/src/host/numth.cu:207-230 get_primes: 2^bits - k*2N + 1, descending.   */
int mfo_get_primes(uint64_t ntt_size, int bit_size, size_t count, uint64_t *out);
/* ph/src/host/modulus.cu:80-112 CoeffModulus::Create for one bit size:
 * the k-th found prime is returned LAST.                                     */
int mfo_coeff_modulus_create(uint64_t n, int bit_size, size_t count, uint64_t *out);
/* ph/src/host/numth.cu:265-332 minimal primitive degree-th root (degree 2^k) */
uint64_t mfo_min_primitive_root(uint64_t degree, uint64_t q);

/* ph/src/host/ntt.cu:11-56: bit-reversed psi-power tables with Shoup words;
 * n^-1 folded into itw[1].  All arrays have N = 1<<logn entries.            */
int mfo_phantom_tables(int logn, uint64_t q,
                       uint64_t *tw, uint64_t *tw_shoup,
                       uint64_t *itw, uint64_t *itw_shoup,
                       uint64_t *n_inv, uint64_t *n_inv_shoup);

/* ph/src/ntt/ntt_1d.cu:17-76 + ph/include/butterfly.cuh:10-22 (forward, lazy
 * CT, natural in -> bit-reversed out, canonical) and ntt_1d.cu:194-256 +
 * butterfly.cuh:28-37 (inverse GS, bit-reversed in -> natural out).          */
void mfo_negacyclic_forward(uint64_t *a, int logn, uint64_t q,
                            const uint64_t *tw, const uint64_t *tw_shoup);
void mfo_negacyclic_inverse(uint64_t *a, int logn, uint64_t q,
                            const uint64_t *itw, const uint64_t *itw_shoup,
                            uint64_t n_inv, uint64_t n_inv_shoup);
/* definitional O(N^2): out[bitrev(k)] = sum_j a_j psi^((2k+1) j)             */
void mfo_negacyclic_forward_def(const uint64_t *a, uint64_t *out, int logn,
                                uint64_t q, uint64_t psi);
/* schoolbook c = a*b in Z_q[X]/(X^N - wrap); wrap = q-1 for X^N+1, psi^N for
 * X^N - i (test/test_custom_ntt_roundtrip.cu:263-280)                        */
void mfo_schoolbook_mul(const uint64_t *a, const uint64_t *b, uint64_t *c,
                        size_t n, uint64_t q, uint64_t wrap);

/* FP64-pipe arithmetic of the CUDA kernels for q < 2^50 (csrc/modmath.cuh), same expressions on the CPU: exactness
 * of the modular product and the compile-time bound schedule of the two networks (ntt_oracle.c) */
long mfo_fp64_modmul_check(uint64_t q, long count, uint64_t seed, int bound16, double *max_ratio);
int mfo_fp64_negacyclic_forward(uint64_t *a, int logn, uint64_t q, const uint64_t *tw, double *stage_max);
int mfo_fp64_negacyclic_inverse(uint64_t *a, int logn, uint64_t q, const uint64_t *itw, uint64_t n_inv, double *stage_max);

/* flat RNS batch [batch][L][N], OpenMP over limb-polys: the timed CPU
 * baseline.  tables are [L][N] as produced by mfo_phantom_tables per limb.   */
void mfo_negacyclic_forward_batch(uint64_t *data, size_t batch, size_t L, int logn,
                                  const uint64_t *moduli,
                                  const uint64_t *tw, const uint64_t *tw_shoup);
void mfo_negacyclic_inverse_batch(uint64_t *data, size_t batch, size_t L, int logn,
                                  const uint64_t *moduli,
                                  const uint64_t *itw, const uint64_t *itw_shoup,
                                  const uint64_t *n_inv, const uint64_t *n_inv_shoup);
int mfo_num_threads(void);
void mfo_set_num_threads(int n);
/* sampler_oracle.c: the sound sampler option */
void mfo_philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1);
void mfo_sound_cdt(uint64_t *out83);
void mfo_sound_sample(int kind, uint64_t *out, const uint64_t *moduli, int n, int limbs, int phi, uint64_t seed0, uint64_t seed1,
                      uint64_t draw);

/* ---- reference "gl" convention, ring Z_q[X]/(X^n - i) ------------------- */
/* src/core/ntt_core.cu:49-70 get_psi: smallest base >= 2 giving order 4n.   */
uint64_t mfo_gl_get_psi(uint64_t q, int n);
/* src/core/ntt_core.cu:394-431,462-481: twist, bit-reverse, CT stages with
 * omega^(j n/m); natural order in and out.  A[k] = sum_j a_j psi^j w^(jk).   */
void mfo_gl_forward(uint64_t *a, int n, uint64_t q);
void mfo_gl_backward(uint64_t *a, int n, uint64_t q);
/* custom_ntt_forward/backward alone (no twist): cyclic DFT, natural order.   */
void mfo_cyclic_forward(uint64_t *a, int n, uint64_t q);
void mfo_cyclic_backward(uint64_t *a, int n, uint64_t q);
/* src/core/ntt_core.cu:150-173 init_gl_perm_tables                          */
void mfo_gl_perm(int n, uint32_t *perm, uint32_t *inv_perm);

#ifdef __cplusplus
}
#endif
#endif
