/*
 * mfhe.h -- C ABI of the B200-native Matrix-FHE hot path (libmfhe_b200.so).
 *
 * Plain C: pointers, sizes and ints only.  Every function returns 0 on success and
 * a negative MFHE_E* code otherwise; it never throws, exits or prints.
 * mfhe_last_error() returns the message of the last failure on the calling thread.
 * Unless a name ends in _host, data pointers are DEVICE pointers owned by the caller
 * and the work is enqueued asynchronously on `stream` (a cudaStream_t passed as
 * void*; NULL = the legacy default stream, as the reference uses).  The data path
 * allocates nothing on the device after plan/context creation; two exceptions, both
 * outside the timed paths: the *_host entry points allocate (and free) their own
 * staging buffers, and mfhe_apply_gl_perm uploads its tiny permutation table on first
 * use of a given n.
 *
 * Streams and threads.  A mfhe_ctx owns ONE scratch arena.  Calls on the same context
 * may come from different streams and different host threads: the enqueue is
 * serialised by a mutex in the context, and a call on a different stream than the
 * previous arena user first waits on the device (cudaStreamWaitEvent) for that user
 * to finish, so live scratch is never shared -- such calls are correct but do not
 * overlap; use one context per stream for concurrency.  Plans (mfhe_ntt_plan,
 * mfhe_bconv_plan) hold read-only tables and may be used from any stream / thread.
 * One process drives one device (the torchrun model): kernel attributes and a few
 * tiny tables are cached per process for the device that first used them.
 *
 * Each entry point cites the reference interface it replaces ("ph/" =
 * extern/phantom-fhe/).  INTEGRATION.md shows the binding a maintainer of the
 * reference adds on their side.
 */
#ifndef MFHE_H
#define MFHE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MFHE_OK            0
#define MFHE_EINVAL       -1   /* bad argument (size, modulus without the needed root, ...) */
#define MFHE_ECUDA        -2   /* CUDA runtime error; see mfhe_last_error()                 */
#define MFHE_ENOMEM       -3
#define MFHE_EUNSUPPORTED -4

const char *mfhe_last_error(void);
/* library version string, e.g. "mfhe-b200 0.1 (sm_100a)" */
const char *mfhe_version(void);

/* ------------------------------------------------------------------------ *
 * Flat RNS negacyclic NTT over [batch][L][N] u64 (family A of SURVEY.md).
 * ------------------------------------------------------------------------ */
typedef struct mfhe_ntt_plan mfhe_ntt_plan;
typedef struct mfhe_bconv_plan mfhe_bconv_plan;      /* base conversion plan, declared below */

/* X-axis conventions of the reference */
#define MFHE_NTT_PHANTOM_NEGACYCLIC 0 /* X^N + 1, natural in -> bit-reversed out
                                         ph/src/host/ntt.cu:11-56 tables, minimal primitive 2N-th root */
#define MFHE_NTT_GL_XN_MINUS_I      1 /* X^n - i, natural in/out: src/core/ntt_core.cu:462-481          */
#define MFHE_NTT_CYCLIC_NATURAL     2 /* X^n - 1, natural in/out: src/core/ntt_core.cu:394-431          */

/* Builds per-limb twiddle tables on the host and uploads them to the current device.
 * Replaces PhantomContext + DNTTTable construction (ph/src/context.cu:121-183,
 * src/core/HE.cu:327-336) and init_ntt_tables_manual / init_gl_twist_tables
 * (src/core/ntt_core.cu:75-198).  logn in 4..8 and 12..17; moduli < 2^60.
 * GL / cyclic conventions are available for logn <= 8 only. */
int mfhe_ntt_plan_create(int logn, int num_limbs, const uint64_t *moduli, int convention,
                         mfhe_ntt_plan **out);
void mfhe_ntt_plan_destroy(mfhe_ntt_plan *plan);
int mfhe_ntt_plan_logn(const mfhe_ntt_plan *plan);
int mfhe_ntt_plan_limbs(const mfhe_ntt_plan *plan);
/* psi used for limb l (the root the convention prescribes), for cross-checking */
uint64_t mfhe_ntt_plan_root(const mfhe_ntt_plan *plan, int limb);
/* copy the forward / inverse table words {w, w_shoup} of one limb to host arrays of N u64
 * each (index order tw[m + i]); inverse != 0 selects the inverse table.  Debug / parity aid. */
int mfhe_ntt_plan_get_table(const mfhe_ntt_plan *plan, int limb, int inverse, uint64_t *w,
                            uint64_t *w_shoup);

/* In-place transforms of `batch` polys of plan->L limbs each, layout [batch][L][N].
 * Replaces nwt_2d_radix8_forward_inplace / backward_inplace (ph/src/ntt/fntt_2d.cu:620-653,
 * ph/src/ntt/intt_2d.cu:724-757; one call per poly there, one call for the batch here)
 * and xy_ntt_forward/backward_phantom, xy_ntt_forward/backward_gl, custom_ntt_forward/
 * backward (src/core/ntt_core.cu:394-481). */
int mfhe_ntt_forward(const mfhe_ntt_plan *plan, uint64_t *d_data, size_t batch, void *stream);
int mfhe_ntt_inverse(const mfhe_ntt_plan *plan, uint64_t *d_data, size_t batch, void *stream);
/* Same on the limb window [limb_start, limb_start+limb_count): data is
 * [batch][limb_count][N] (Phantom's `start_modulus_idx` / `coeff_modulus_size`). */
int mfhe_ntt_forward_limbs(const mfhe_ntt_plan *plan, uint64_t *d_data, size_t batch,
                           int limb_start, int limb_count, void *stream);
int mfhe_ntt_inverse_limbs(const mfhe_ntt_plan *plan, uint64_t *d_data, size_t batch,
                           int limb_start, int limb_count, void *stream);
/* Host-buffer convenience (what a cgo/ctypes caller with host memory uses): H2D copy,
 * transform, D2H copy, synchronous.  Uses an internal device staging buffer that grows
 * to the largest request. */
int mfhe_ntt_forward_host(mfhe_ntt_plan *plan, uint64_t *h_data, size_t batch);
int mfhe_ntt_inverse_host(mfhe_ntt_plan *plan, uint64_t *h_data, size_t batch);

/* Tuning knob: number of limb-polys processed per column+row kernel pair so that the
 * intermediate stays in L2 (0 = whole batch in one pair of launches). */
int mfhe_ntt_plan_set_chunk(mfhe_ntt_plan *plan, size_t chunk_limb_polys);

/* Arithmetic of the transforms (flat N = 2^12..2^17 and the n <= 256 networks, all conventions).  0: 64-bit integer
 * butterflies (any modulus below 2^60).
 * 1: butterflies on the FP64 pipe -- exact, same results word for word -- available when every modulus of the plan is
 * below 2^50 (SURVEY config 2's 50-bit run, the 35/45-bit moduli of the matrix scheme) and then the default; inputs
 * must be below 2q.  MFHE_NTT_ARITH=int in the environment keeps new plans on 0.  set_arith returns MFHE_EUNSUPPORTED
 * when 1 is asked of a plan that cannot run it. */
int mfhe_ntt_plan_arith(const mfhe_ntt_plan *plan);
int mfhe_ntt_plan_set_arith(mfhe_ntt_plan *plan, int arith);

/* Pointwise ops on [batch][L][N] in the plan's moduli (canonical in, canonical out).
 * Replaces pointwise_mul kernels of the reference (src/core/HE.cu:509-531 semantics with
 * a full-size second operand) and Phantom's multiply_rns_poly. */
int mfhe_poly_mul(const mfhe_ntt_plan *plan, const uint64_t *d_a, const uint64_t *d_b,
                  uint64_t *d_out, size_t batch, void *stream);
int mfhe_poly_add(const mfhe_ntt_plan *plan, const uint64_t *d_a, const uint64_t *d_b,
                  uint64_t *d_out, size_t batch, void *stream);
int mfhe_poly_sub(const mfhe_ntt_plan *plan, const uint64_t *d_a, const uint64_t *d_b,
                  uint64_t *d_out, size_t batch, void *stream);

/* Ring product of every limb-poly of a with that of b:  a <- INTT(NTT(a) (.) NTT(b))  (X^N + 1 for the Phantom
 * convention; the plan's ring otherwise), canonical residues in and out, coefficient order in and out.
 * b is transformed in place and LEFT in NTT form (pass b_is_ntt != 0 if it already is: one operand reused for
 * many products).  For N >= 2^12 the pointwise product is applied while the first inverse pass loads its data,
 * so there is no pass of its own over a and b.  Replaces the sequence nwt_2d_radix8_forward_inplace x2 /
 * multiply_rns_poly / nwt_2d_radix8_backward_inplace (ph/src/ntt/fntt_2d.cu:620-653, intt_2d.cu:724-757) and, for
 * n <= 256, the reference's NTT_X / pointwise_mul / INTT_X chains (src/core/HE.cu:1508-1530). */
int mfhe_ntt_mul_intt(const mfhe_ntt_plan *plan, uint64_t *d_a, uint64_t *d_b, int b_is_ntt,
                      size_t batch, void *stream);


/* ======================================================================== *
 * Matrix pipeline (family B of SURVEY.md): the reference's native ring
 * Z_Q[X,Y,W]/(X^n - i, Y^n - i, Phi_p(W)), p = 3 f.  Reference shape: n = 64,
 * f = 257 (p = 771, phi = 512 lanes), L = 11, delta = 2^35 (include/core/config.h:7-44).
 *
 * Layouts of an RNS "poly set" of T = phi*L*n*n words (src/core/HE.cu:1330-1368):
 *   MATRIX  [w][limb][y][x]            (ciphertexts, encoder output, samplers)
 *   POLY    [(w*n + y)][limb][x]       (what the reference's W-CRT kernels emit)
 * Complex buffers are interleaved (re, im) doubles == cuDoubleComplex.
 * ======================================================================== */
typedef struct mfhe_ctx mfhe_ctx;
#define MFHE_LAYOUT_MATRIX 0
#define MFHE_LAYOUT_POLY   1

typedef struct mfhe_params {
    int n;                   /* MATRIX_N: 16, 32 or 64                                             */
    int f;                   /* odd prime; p = 3 f, phi = 2 (f - 1).  Reference: 257               */
    int num_limbs;           /* RNS_NUM_LIMBS                                                      */
    const uint64_t *moduli;  /* each = 1 mod lcm(2n, p), < 2^59                                    */
    double delta;            /* SCALING_FACTOR                                                     */
    size_t arena_bytes;      /* device scratch reserved at creation; 0 = what the pipeline needs   */
} mfhe_params;

/* Replaces init_he_backend (src/core/HE.cu:318-408): W-CRT / W-DFT tables, XY Vandermonde,
 * CRT constants, n-point Phantom NTT tables, scratch arena -- everything the hot path needs, so
 * that no later call allocates. */
int mfhe_ctx_create(const mfhe_params *params, mfhe_ctx **out);
void mfhe_ctx_destroy(mfhe_ctx *ctx);
/* {n, f, p, phi, L, big_words} */
int mfhe_ctx_info(const mfhe_ctx *ctx, int out[6]);
/* copy_device_moduli (src/core/HE.cu:410-422): host out[count] */
int mfhe_ctx_get_moduli(const mfhe_ctx *ctx, uint64_t *h_out, int count);
uint64_t mfhe_ctx_eta(const mfhe_ctx *ctx, int limb);          /* h_find_eta, HE.cu:119-133 */
/* host copy of the reference's device tables d_wntt_powers / d_wntt_inv_powers for one limb:
 * out[w*phi + r] = V[w][r]  (inverse: V^-1[r][w]), HE.cu:258-266.  Parity aid. */
int mfhe_ctx_get_wcrt_table(const mfhe_ctx *ctx, int limb, int inverse, uint64_t *h_out);
/* CRT constants as the reference uploads them (encoder.cu:341-421): which = 0 Q[W], 1 Q/2[W],
 * 2 M[L][W], 3 inv[L] */
int mfhe_ctx_get_crt(const mfhe_ctx *ctx, int which, uint64_t *h_out);
/* the n-point X-axis plan (Phantom negacyclic) owned by the context: get_xy_ntt_table, HE.cu:429-435 */
const mfhe_ntt_plan *mfhe_ctx_xy_plan(const mfhe_ctx *ctx);
/* device copy of the XY Vandermonde matrices the reference keeps as public Encoder members
 * (encoder.cuh:46-49; encoder.cu:425-444): which = 0 V, 1 V^T, 2 V^-1, 3 (V^-1)^T; d_out n*n complex */
int mfhe_ctx_copy_xy_matrix(const mfhe_ctx *ctx, int which, double *d_out, void *stream);

/* ---- W-CRT: wntt_forward_matrix / wntt_inverse_matrix (HE.cu:437-452, 716-781) with explicit
 * layouts (reference: forward MATRIX -> POLY, inverse POLY -> MATRIX); out must not alias in.
 * Inputs are canonical residues (< q of their limb), as every producer on the path emits them. */
int mfhe_wcrt_forward(mfhe_ctx *ctx, const uint64_t *d_in, uint64_t *d_out, int in_layout, int out_layout, void *stream);
int mfhe_wcrt_inverse(mfhe_ctx *ctx, const uint64_t *d_in, uint64_t *d_out, int in_layout, int out_layout, void *stream);
/* wntt_forward_vector_kernel (HE.cu:1245-1270): [w][limb][x] */
int mfhe_wcrt_forward_vector(mfhe_ctx *ctx, const uint64_t *d_in, uint64_t *d_out, void *stream);
/* wntt_forward_centered / wntt_inverse_centered (HE.cu:454-470, 1029-1114): int64 [w][y][x] */
int mfhe_wcrt_forward_centered(mfhe_ctx *ctx, const int64_t *d_in, int64_t *d_out, void *stream);
int mfhe_wcrt_inverse_centered(mfhe_ctx *ctx, const int64_t *d_in, int64_t *d_out, void *stream);
/* matrix_to_poly_kernel / poly_to_matrix_kernel (HE.cu:1330-1368, batched_encoder.cu:83-102) */
int mfhe_layout_convert(mfhe_ctx *ctx, const uint64_t *d_in, uint64_t *d_out, int in_layout, int out_layout, void *stream);

/* ---- complex W-DFT (HE.cu:472-502, 1116-1202; batched_encoder.cu:104-123): [phi][n*n] ------ */
int mfhe_wdft_forward_complex(mfhe_ctx *ctx, const double *d_in, double *d_out, void *stream);
int mfhe_wdft_inverse_complex(mfhe_ctx *ctx, const double *d_in, double *d_out, void *stream);
int mfhe_wdft_forward_centered_pair(mfhe_ctx *ctx, const int64_t *d_re, const int64_t *d_im,
                                    double *d_out_re, double *d_out_im, void *stream);
int mfhe_wdft_inverse_pair(mfhe_ctx *ctx, const double *d_re, const double *d_im,
                           double *d_out_re, double *d_out_im, void *stream);

/* ---- XY canonical embedding, quantise, CRT (src/core/encoder.cu) ------------------------------ */
/* Encoder::idft2 = Vinv M Vinv^T (encoder.cu:460-467) / decode_from_eval_complex = V E V^T
 * (encoder.cu:492-501) for `lanes` n x n complex matrices in one launch */
int mfhe_xy_idft2(mfhe_ctx *ctx, const double *d_in, double *d_out, int lanes, void *stream);
int mfhe_xy_dft2(mfhe_ctx *ctx, const double *d_in, double *d_out, int lanes, void *stream);
/* mat_mul_kernel_complex (encoder.cu:318-326): C = A B, n x n complex */
int mfhe_matmul_complex(const double *d_A, const double *d_B, double *d_C, int n, void *stream);
/* quantize_coeff_to_rns_kernel (batched_encoder.cu:125-152; lanes = phi) and quantize_soa_kernel
 * (encoder.cu:36-50; lanes = 1): llround(z*delta) mod q_l -> [lane][limb][n*n] */
int mfhe_quantize(mfhe_ctx *ctx, const double *d_in, uint64_t *d_re, uint64_t *d_im, int lanes, void *stream);
/* crt_compose_centerlift_big (encoder.cu:191-245): in [lane][limbs][n2] -> mag [lane][n2][big_words],
 * neg [lane][n2]; lanes > 1 replaces the reference's per-lane host loop (HE.cu:1653-1668) */
int mfhe_crt_compose_big(mfhe_ctx *ctx, const uint64_t *d_in, uint64_t *d_mag, uint8_t *d_neg,
                         int n2, int limbs, int lanes, void *stream);
/* crt_compose_centerlift_kernel (encoder.cu:152-189; keeps the low word like the reference) */
int mfhe_crt_compose_i64(mfhe_ctx *ctx, const uint64_t *d_in, int64_t *d_out, int n2, int limbs, void *stream);
/* dequantize_exact_kernel (encoder.cu:112-150) */
int mfhe_dequantize_exact(mfhe_ctx *ctx, const uint64_t *d_re, const uint64_t *d_im, double *d_out,
                          int n2, int limbs, void *stream);
/* compose_big_pair_to_complex_by_delta_kernel (HE.cu:1007-1027) */
int mfhe_big_pair_to_complex(mfhe_ctx *ctx, const uint64_t *d_re_mag, const uint8_t *d_re_neg,
                             const uint64_t *d_im_mag, const uint8_t *d_im_neg, double *d_out,
                             size_t total, void *stream);
/* BatchedEncoder::encode_to_wntt_eval (batched_encoder.cu:161-228): msg [phi][n*n] complex ->
 * out_re, out_im MATRIX layout */
int mfhe_encode(mfhe_ctx *ctx, const double *d_msg, uint64_t *d_out_re, uint64_t *d_out_im, void *stream);
/* Encoder::encode (encoder.cu:446-456), one lane -> [limb][n*n] */
int mfhe_encode_lane(mfhe_ctx *ctx, const double *d_msg, uint64_t *d_re, uint64_t *d_im, void *stream);
/* decode_eval_pair_to_complex (HE.cu:1619-1689): evals in `layout` -> out [phi][n*n] complex */
int mfhe_decode_eval_pair(mfhe_ctx *ctx, const uint64_t *d_eval_re, const uint64_t *d_eval_im, int layout,
                          double *d_out, void *stream);
/* Encoder::decode_lane_from_rns_eval (encoder.cu:469-490) */
int mfhe_decode_lane_from_rns_eval(mfhe_ctx *ctx, const uint64_t *d_re, const uint64_t *d_im, double *d_out, void *stream);

/* ---- samplers, keys, encrypt / decrypt (src/core/HE.cu) ------------------------------------------ */
/* uniform_random_kernel :564-578, gaussian_noise_kernel :590-627 (MATRIX, W-coeff),
 * ternary_secret_kernel :690-713 ([w][limb][x]) */
/* Sampler option.  mode 0 (default): the reference's index-hash "samplers" (HE.cu:564-627, 690-713),
 * reproducible and NOT encryption randomness -- the parity mode.  mode 1: Philox4x32-10 keyed by the 128-bit `seed`
 * (uniform by 128-bit reduction, discrete gaussian sigma = 3.2 by a 64-bit cumulative table, uniform ternary key), every
 * sampled array under a fresh draw number; affects mfhe_sample_*, mfhe_keygen, mfhe_encrypt*, mfhe_encode_encrypt_pair.
 * Setting the mode resets the draw number (same seed => same streams: testable against the oracle). */
int mfhe_ctx_set_sampler(mfhe_ctx *ctx, int mode, const uint64_t seed[2]);
/* the 83-entry cumulative table of the discrete gaussian (z = -41 .. 41), for cross-checks */
int mfhe_sampler_cdt(uint64_t *h_out, int count);
int mfhe_sample_uniform(mfhe_ctx *ctx, uint64_t *d_out, void *stream);
int mfhe_sample_gaussian(mfhe_ctx *ctx, uint64_t *d_out, void *stream);
int mfhe_sample_ternary(mfhe_ctx *ctx, uint64_t *d_out, void *stream);
/* generate_secret_key (HE.cu:1272-1307): d_sk [w][limb][x], W-eval + X-NTT (Phantom order) */
int mfhe_keygen(mfhe_ctx *ctx, uint64_t *d_sk, void *stream);
/* encrypt_pair (HE.cu:1455-1552).  d_ct_* hold [b | a], T words each, MATRIX layout. */
int mfhe_encrypt_pair(mfhe_ctx *ctx, const uint64_t *d_m_re, const uint64_t *d_m_im, const uint64_t *d_sk,
                      uint64_t *d_ct_re, uint64_t *d_ct_im, void *stream);
/* the same with caller-supplied a (W-coeff, MATRIX) and errors (W-coeff, MATRIX): the
 * parity-checked integer core */
int mfhe_encrypt_pair_with(mfhe_ctx *ctx, const uint64_t *d_m_re, const uint64_t *d_m_im, const uint64_t *d_sk,
                           const uint64_t *d_a_coeff, const uint64_t *d_e_re, const uint64_t *d_e_im,
                           uint64_t *d_ct_re, uint64_t *d_ct_im, void *stream);
/* encode_to_wntt_eval followed by encrypt_pair (batched_encoder.cu:161-228 + HE.cu:1455-1552) as ONE call with
 * identical ciphertext words: message and error are summed in the W-coeff domain and share a W-CRT (3 instead
 * of 4 per pair).  d_msg: phi * n^2 complex fp64. */
int mfhe_encode_encrypt_pair(mfhe_ctx *ctx, const double *d_msg, const uint64_t *d_sk, uint64_t *d_ct_re,
                             uint64_t *d_ct_im, void *stream);
/* encrypt (HE.cu:1370-1453) */
int mfhe_encrypt(mfhe_ctx *ctx, const uint64_t *d_m, const uint64_t *d_sk, uint64_t *d_ct, void *stream);
/* decrypt_to_eval (HE.cu:1553-1601; out_layout POLY) / decrypt_to_eval_matrix (:1603-1617; MATRIX) */
int mfhe_decrypt_to_eval(mfhe_ctx *ctx, const uint64_t *d_ct, const uint64_t *d_sk, uint64_t *d_out,
                         int out_layout, void *stream);
/* decrypt_and_decode (HE.cu:1691-1708) */
int mfhe_decrypt_decode(mfhe_ctx *ctx, const uint64_t *d_ct_re, const uint64_t *d_ct_im, const uint64_t *d_sk,
                        double *d_out, void *stream);
/* add_ciphertexts / multiply_ciphertexts_raw (HE.cu:631-669, 1710-1740).  reference_indexing != 0
 * takes the modulus of flat word idx as limb (idx / n) % L exactly as the reference kernels do
 * (they index MATRIX data with the POLY formula); 0 uses the MATRIX layout's true limb. */
int mfhe_ct_add(mfhe_ctx *ctx, const uint64_t *d_ct1, const uint64_t *d_ct2, uint64_t *d_res,
                int reference_indexing, void *stream);
int mfhe_ct_mul_raw(mfhe_ctx *ctx, const uint64_t *d_ct1, const uint64_t *d_ct2, uint64_t *d_d0, uint64_t *d_d1,
                    uint64_t *d_d2, int reference_indexing, void *stream);

/* ---- trace step (src/core/trace.cu:30-161, batched_trace.cu:37-197): [batch][limb][n][n] -------- */
int mfhe_trace_map(mfhe_ctx *ctx, const uint64_t *d_Br, const uint64_t *d_Bi, uint64_t *d_Bpr, uint64_t *d_Bpi,
                   int batch, void *stream);
int mfhe_trace_gemm(mfhe_ctx *ctx, const uint64_t *d_Ar, const uint64_t *d_Ai, const uint64_t *d_Bpr,
                    const uint64_t *d_Bpi, uint64_t *d_Cr, uint64_t *d_Ci, int batch, void *stream);
int mfhe_trace_rescale(mfhe_ctx *ctx, uint64_t *d_Cr, uint64_t *d_Ci, int batch, uint64_t inv0, uint64_t inv1,
                       uint64_t inv2, void *stream);

/* ---- GL permutation (src/core/ntt_core.cu:150-173, 433-441) --------------------------------------- */
/* ---- encrypted matrix product: SURVEY 8(f) rank 1.  The reference ships the pieces (multiply_ciphertexts_raw
 * HE.cu:1727-1740, the trace step trace.cu:30-161 / batched_trace.cu:37-197) but no caller; these entry points wire the
 * trace GEMM onto ciphertexts.  Layout of every operand: poly sets [w][l][y][x] (MFHE_LAYOUT_MATRIX). ---- */
/* trace.cu:30-62 with the twist along x (the fast index of this layout) instead of the slow one: applied to an ENCODED
 * plaintext before encryption; batch = number of (lane) matrices per limb set, i.e. phi for a full poly set */
int mfhe_trace_map_x(mfhe_ctx *ctx, const uint64_t *d_Br, const uint64_t *d_Bi, uint64_t *d_Bpr, uint64_t *d_Bpi, int batch, void *stream);
/* trace.cu:77-120 contracting over the SLOW index: D[x][z] = n * sum_y U[y][x] V[y][z] (complex, mod q), tensor cores */
int mfhe_trace_gemm_ty(mfhe_ctx *ctx, const uint64_t *d_Ur, const uint64_t *d_Ui, const uint64_t *d_Vr, const uint64_t *d_Vi,
                       uint64_t *d_Dr, uint64_t *d_Di, int batch, void *stream);
/* every n x n matrix of a [count][n][n] array transposed */
int mfhe_transpose(mfhe_ctx *ctx, const uint64_t *d_in, uint64_t *d_out, size_t count, void *stream);
/* trace.cu:132-161 repaired: limb l is multiplied by h_inv[l] (HOST array of L constants) */
int mfhe_trace_rescale_limbs(mfhe_ctx *ctx, uint64_t *d_Cr, uint64_t *d_Ci, int batch, const uint64_t *h_inv, void *stream);
/* the four component GEMMs of two ciphertext pairs; d_out = 8 poly sets [bb re, im | ab re, im | ba re, im | aa re, im],
 * each [w][l][x][z].  n = 64 and moduli below 2^48: one tcgen05 kernel (csrc/ct_matmul_umma.cu); otherwise four GEMMs of
 * the mfhe_trace_gemm_ty kind (MFHE_MATMUL_PATH=m forces those). */
int mfhe_ct_matmul(mfhe_ctx *ctx, const uint64_t *d_ctU_re, const uint64_t *d_ctU_im, const uint64_t *d_ctV_re,
                   const uint64_t *d_ctV_im, uint64_t *d_out, void *stream);
/* decrypts the 4-component product; out = evals [w][l][z][x] of n * sum_y U V'.  d_prod may be used as scratch (it is for
 * n != 64; the fused n = 64 kernel only reads it): treat it as consumed */
int mfhe_ct_matmul_decrypt(mfhe_ctx *ctx, uint64_t *d_prod, const uint64_t *d_sk, uint64_t *d_out_re, uint64_t *d_out_im, void *stream);

/* ---- multi-GPU: limb-sharded base conversion (BASELINE config 5).  One process per GPU.  The exchange
 * is fused into the base-conversion kernel (its stores go to peer memory over NVLink); the communicator only holds the
 * double-buffered receive buffers and barrier flags, shared between the processes as CUDA IPC handles that the HOST program
 * moves (MFHE_COMM_HANDLE_BYTES per rank, any transport).  No NCCL, no collective. ---- */
typedef struct mfhe_comm mfhe_comm;
#define MFHE_COMM_HANDLE_BYTES 192
size_t mfhe_bconv_sharded_recv_words(int world, int l_out, size_t N, size_t batch);     /* receive-buffer size for create */
int mfhe_comm_create(int rank, int world, size_t recv_words, mfhe_comm **out);
void mfhe_comm_destroy(mfhe_comm *comm);
int mfhe_comm_export(const mfhe_comm *comm, void *h_handle /* MFHE_COMM_HANDLE_BYTES */);
int mfhe_comm_import(mfhe_comm *comm, const void *h_all_handles /* world x MFHE_COMM_HANDLE_BYTES, rank order */);
int mfhe_comm_barrier(mfhe_comm *comm, void *stream);            /* stream-ordered, release / acquire at system scope */
/* batch slices per step (default 1): with g > 1 the step is software-pipelined -- slice s's reduce + NTT run on a side stream
 * while slice s+1's INTT and NVLink-bound scatter run on the caller's stream, one barrier per slice */
int mfhe_comm_set_groups(mfhe_comm *comm, int groups);
/* INTT(own Q-limbs) -> partial sums stored into the owners' receive buffers -> barrier -> reduce -> NTT(own P-limbs).
 * plan_q / plan_p / bconv cover the FULL bases; d_x_local [batch][q_count][N] (destroyed), d_out [batch][p_count][N];
 * rank r owns the r-th balanced contiguous window of each base (Phantom: nwt_2d_radix8_backward_inplace + bConv_BEHZ +
 * nwt_2d_radix8_forward_inplace on one GPU, ph/src/rns_bconv.cu:212-229). */
int mfhe_bconv_sharded_step(mfhe_comm *comm, const mfhe_ntt_plan *plan_q, const mfhe_ntt_plan *plan_p,
                            const mfhe_bconv_plan *bconv, uint64_t *d_x_local, uint64_t *d_out, size_t batch, void *stream);

/* ---- mod-up / mod-down with an auxiliary base P (config.h:22,46-52 reserves P_MODULI, Phantom:
 * ph/src/ntt/ntt_modup.cu, ntt_moddown.cu, ph/src/rns_bconv.cu:530-830).  `qp` is a flat plan whose limbs are Q then P.
 * mod-up:   d_x [batch][LQ][N] COEFFICIENT form over Q -> d_out [batch][LQ+LP][N] NTT form over Q u P (q_to_p: Q -> P).
 * mod-down: d_x [batch][LQ+LP][N] NTT form -> d_out [batch][LQ][N] NTT form, (x - [x]_P) / P up to the base conversion's
 *           multiple-of-P slack (|error| < LP); p_to_q: P -> Q; d_scratch: batch * LP * N words. ---- */
int mfhe_modup(const mfhe_ntt_plan *qp, const mfhe_bconv_plan *q_to_p, const uint64_t *d_x, uint64_t *d_out, size_t batch,
               void *stream);
int mfhe_moddown(const mfhe_ntt_plan *qp, const mfhe_bconv_plan *p_to_q, const uint64_t *d_x, uint64_t *d_out,
                 uint64_t *d_scratch, size_t batch, void *stream);
/* mod-down that allocates nothing and asks no scratch of its caller.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same arguments and the same words out as mfhe_moddown; the scratch is a static device array
 * inside the library (held by every GPU that loads it until exit).  Limit: batch * LP * N <= MFHE_MODDOWN_STATIC_WORDS
 * (16 MiB), beyond it MFHE_EINVAL with the sizes in mfhe_last_error().  All calls share that one array: calls that overlap
 * on the GPU (two streams, two unordered threads) race -- keep them on one stream or order them with events.  The first
 * call on a device resolves the array's device address (do it once before a graph capture); no call uses cudaMalloc,
 * cudaFree or an allocating library routine. */
#define MFHE_MODDOWN_STATIC_WORDS ((size_t)1 << 21)
int mfhe_moddown_static(const mfhe_ntt_plan *qp, const mfhe_bconv_plan *p_to_q, const uint64_t *d_x, uint64_t *d_out,
                        size_t batch, void *stream);

/* ---- RNS rescale of the matrix scheme (the reference reserves the chain -- delta = 2^35 next to
 * ten 35-bit primes, config.h:25-44 -- and ships only rescale_by_delta_rns, trace.cu:132-161, a multiplication by constants).
 * X' = (X - [X]_q_last) / q_last on the integer W-coefficients; in [w][L][n^2] -> out [w][L-1][n^2] (a poly set of a context
 * created on moduli[0 .. L-2]).  Works on plaintext evals and on each half / component of a ciphertext alike. */
int mfhe_rescale_last(mfhe_ctx *ctx, const uint64_t *d_in, uint64_t *d_out, void *stream);

int mfhe_gl_perm_tables(int n, uint32_t *h_perm, uint32_t *h_inv_perm);
int mfhe_apply_gl_perm(const uint64_t *d_in, uint64_t *d_out, int limbs, int batch_count, int n, int inverse, void *stream);

/* ---- RNS base conversion (Phantom DBaseConverter::bConv_BEHZ, ph/src/rns_bconv.cu:212-229) -------- */
int mfhe_bconv_plan_create(int l_in, const uint64_t *q_in, int l_out, const uint64_t *p_out, mfhe_bconv_plan **out);
void mfhe_bconv_plan_destroy(mfhe_bconv_plan *plan);
/* d_in [batch][l_in][N] -> d_out [batch][l_out][N] */
int mfhe_bconv_plan_sizes(const mfhe_bconv_plan *plan, int *l_in, int *l_out);
int mfhe_bconv(const mfhe_bconv_plan *plan, const uint64_t *d_in, uint64_t *d_out, size_t N, size_t batch, void *stream);
/* limb-sharded form: this rank holds input limbs [in_start, in_start+in_count); writes the partial
 * sums mod p_j of ALL output limbs, each < p_j, to d_partial [batch][l_out][N]; summing the ranks'
 * partials mod p_j (mfhe_bconv_reduce after an all-gather / reduce-scatter) equals mfhe_bconv. */
int mfhe_bconv_partial(const mfhe_bconv_plan *plan, const uint64_t *d_in_local, int in_start, int in_count,
                       uint64_t *d_partial, size_t N, size_t batch, void *stream);
/* fused partial + exchange: the same sums as mfhe_bconv_partial, but output limb j of batch item b is
 * stored at dst[j] + b * dst_batch_stride[j] (words; dst / dst_batch_stride are HOST arrays of l_out
 * entries, passed to the kernel by value).  dst[j] may point into a PEER GPU's memory (NVLink P2P /
 * CUDA symmetric memory): the kernel then writes every partial sum straight into the receive buffer of
 * the rank that owns output limb j -- layout [src_rank][batch][out_count][N] there, which is what
 * mfhe_bconv_reduce consumes -- so the exchange of ph/src/rns_bconv.cu:212-229's cross-limb sum rides
 * on the compute kernel's stores and no collective follows (the caller orders steps with a barrier). */
int mfhe_bconv_partial_scatter(const mfhe_bconv_plan *plan, const uint64_t *d_in_local, int in_start, int in_count,
                               uint64_t *const *dst, const size_t *dst_batch_stride, size_t N, size_t batch,
                               void *stream);
/* d_parts [nparts][batch][out_count][N] -> d_out [batch][out_count][N], output limbs
 * [out_start, out_start+out_count) */
int mfhe_bconv_reduce(const mfhe_bconv_plan *plan, const uint64_t *d_parts, int nparts, int out_start, int out_count,
                      uint64_t *d_out, size_t N, size_t batch, void *stream);
/* the same on a batch slice of a larger [nparts][batch][count][N] buffer: part_stride words between consecutive parts */
int mfhe_bconv_reduce_strided(const mfhe_bconv_plan *plan, const uint64_t *d_parts, int nparts, size_t part_stride,
                              int out_start, int out_count, uint64_t *d_out, size_t N, size_t batch, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* MFHE_H */
