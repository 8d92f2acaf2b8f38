/*
 * scheme_oracle.h -- CPU restatement of the reference's matrix pipeline (family B of SURVEY.md):
 * W-CRT, complex W-DFT, XY canonical embedding, quantise / CRT compose, samplers, keygen,
 * encrypt / decrypt, ciphertext ops, trace step, base conversion.
 *
 * TEST INFRASTRUCTURE ONLY (see mfhe_oracle.h).  Every function cites the reference
 * file:line it restates.  The ring is Z_Q[X,Y,W]/(X^n - i, Y^n - i, Phi_p(W)) with
 * p = 3*f (f an odd prime; the reference has f = 257, p = 771, phi = 512, n = 64, L = 11);
 * the generalisation to other (n, f, L) only re-reads the reference's constants as
 * parameters so that the parity tests can also run at shapes that finish in milliseconds.
 */
#ifndef MFHE_SCHEME_ORACLE_H
#define MFHE_SCHEME_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MFO_MAX_LIMBS 64
#define MFO_MAX_BIG   32      /* 64-bit words of the CRT big integer */

typedef struct { double re, im; } mfo_cx;

typedef struct mfo_ctx {
    int n, logn, f, p, phi, L, big_words;
    double delta;
    uint64_t moduli[MFO_MAX_LIMBS];
    uint64_t eta[MFO_MAX_LIMBS];
    uint16_t *wexp;          /* [phi]  k_wntt_exp                       HE.cu:72-105          */
    uint64_t *wntt;          /* [L][w][r]   V_l[w][r]                   HE.cu:258-262         */
    uint64_t *wntt_inv;      /* [L][w][r] = V_l^-1[r][w]                HE.cu:263-266         */
    mfo_cx *wdft;            /* [w][r]      e^(2 pi i exp[w] r / p)     HE.cu:275-310         */
    mfo_cx *wdft_inv;        /* [w][r] = V^-1[r][w]                                          */
    mfo_cx *xyV, *xyVT, *xyVinv, *xyVinvT;   /* [n][n]                  encoder.cu:425-444    */
    uint64_t crt_Q[MFO_MAX_BIG], crt_Qhalf[MFO_MAX_BIG];                /* encoder.cu:341-421 */
    uint64_t crt_M[MFO_MAX_LIMBS * MFO_MAX_BIG];   /* stride big_words */
    uint64_t crt_inv[MFO_MAX_LIMBS];
    /* Phantom n-point tables per limb (HE.cu:327-336 -> ph/src/host/ntt.cu:11-56) */
    uint64_t *tw, *tws, *itw, *itws;     /* [L][n] */
    uint64_t ninv[MFO_MAX_LIMBS], ninvs[MFO_MAX_LIMBS];
} mfo_ctx;

/* gauss_jordan != 0: invert V_l and the complex V by the reference's literal Gauss-Jordan
 * (HE.cu:135-235); 0: closed form (inverse DFT restricted to the units, reduced mod Phi_p),
 * which is the same matrix mod q and the exact inverse in fp64. */
mfo_ctx *mfo_ctx_create(int n, int f, int L, const uint64_t *moduli, double delta, int gauss_jordan);
void mfo_ctx_destroy(mfo_ctx *c);
const uint64_t *mfo_ctx_wntt(const mfo_ctx *c, int inverse);
const uint16_t *mfo_ctx_wexp(const mfo_ctx *c);
uint64_t mfo_ctx_eta(const mfo_ctx *c, int limb);
int mfo_ctx_big_words(const mfo_ctx *c);
const uint64_t *mfo_ctx_crt(const mfo_ctx *c, int which);   /* 0 Q, 1 Qhalf, 2 M, 3 inv */
const mfo_cx *mfo_ctx_cx(const mfo_ctx *c, int which);      /* 0 wdft 1 wdft_inv 2 V 3 VT 4 Vinv 5 VinvT */

/* HE.cu:119-133 */
uint64_t mfo_find_eta(uint64_t q, int f);
/* HE.cu:135-185 */
int mfo_matrix_inverse_mod(const uint64_t *m, uint64_t *inv, int dim, uint64_t q);

/* ---- W-CRT --------------------------------------------------------------------------- */
/* HE.cu:716-747: in matrix-major [r][l][y][x] -> out poly-major [(w n + y)][l][x] */
void mfo_wntt_forward_matrix(const mfo_ctx *c, const uint64_t *in, uint64_t *out);
/* HE.cu:751-781: in poly-major -> out matrix-major */
void mfo_wntt_inverse_matrix(const mfo_ctx *c, const uint64_t *in, uint64_t *out);
/* HE.cu:1245-1270: [w][l][x] */
void mfo_wntt_forward_vector(const mfo_ctx *c, const uint64_t *in, uint64_t *out);
/* HE.cu:1029-1114 */
void mfo_wntt_forward_centered(const mfo_ctx *c, const int64_t *in, int64_t *out);
void mfo_wntt_inverse_centered(const mfo_ctx *c, const int64_t *in, int64_t *out);
/* HE.cu:1330-1368, batched_encoder.cu:83-102 */
void mfo_matrix_to_poly(const mfo_ctx *c, const uint64_t *in, uint64_t *out);
void mfo_poly_to_matrix(const mfo_ctx *c, const uint64_t *in, uint64_t *out);

/* ---- complex W-DFT -------------------------------------------------------------------- */
/* HE.cu:1147-1172 */
void mfo_wdft_forward_complex(const mfo_ctx *c, const mfo_cx *in, mfo_cx *out);
/* batched_encoder.cu:104-123 */
void mfo_w_idft(const mfo_ctx *c, const mfo_cx *in, mfo_cx *out);
/* HE.cu:1116-1145, 1174-1202 */
void mfo_wdft_forward_pair(const mfo_ctx *c, const int64_t *re, const int64_t *im, double *ore, double *oim);
void mfo_wdft_inverse_pair(const mfo_ctx *c, const double *re, const double *im, double *ore, double *oim);

/* ---- XY embedding, quantise, CRT ------------------------------------------------------- */
/* encoder.cu:460-467 (idft2: Vinv * M * VinvT) and :492-501 (V * E * VT), `lanes` matrices */
void mfo_xy_idft2(const mfo_ctx *c, const mfo_cx *in, mfo_cx *out, int lanes);
void mfo_xy_dft2(const mfo_ctx *c, const mfo_cx *in, mfo_cx *out, int lanes);
/* batched_encoder.cu:125-152 (lanes = phi) and encoder.cu:36-50 (lanes = 1): out [lane][l][n2] */
void mfo_quantize(const mfo_ctx *c, const mfo_cx *in, uint64_t *ore, uint64_t *oim, int lanes);
/* encoder.cu:191-230: in [l][n2] (one lane) -> mag [n2][big_words], neg [n2] */
void mfo_crt_compose_big(const mfo_ctx *c, const uint64_t *in, uint64_t *mag, uint8_t *neg, int n2, int limbs);
/* encoder.cu:152-189 (truncates to the low word) */
void mfo_crt_compose_i64(const mfo_ctx *c, const uint64_t *in, int64_t *out, int n2, int limbs);
/* encoder.cu:112-150 */
void mfo_dequantize_exact(const mfo_ctx *c, const uint64_t *re, const uint64_t *im, mfo_cx *out, int n2, int limbs);
/* HE.cu:1007-1027 */
void mfo_big_pair_to_complex(const mfo_ctx *c, const uint64_t *re_mag, const uint8_t *re_neg,
                             const uint64_t *im_mag, const uint8_t *im_neg, mfo_cx *out, size_t total);
/* batched_encoder.cu:161-228 */
void mfo_encode(const mfo_ctx *c, const mfo_cx *msg, uint64_t *out_re, uint64_t *out_im);
/* encoder.cu:446-456 (single lane) */
void mfo_encode_lane(const mfo_ctx *c, const mfo_cx *msg, uint64_t *out_re, uint64_t *out_im);
/* HE.cu:1619-1689: inputs poly-major evals */
void mfo_decode_eval_pair(const mfo_ctx *c, const uint64_t *eval_re, const uint64_t *eval_im, mfo_cx *out);
/* encoder.cu:469-490 */
void mfo_decode_lane_from_rns_eval(const mfo_ctx *c, const uint64_t *re, const uint64_t *im, mfo_cx *out);

/* ---- samplers, keys, encrypt / decrypt --------------------------------------------------- */
/* HE.cu:564-578, 590-627, 690-713 */
void mfo_sample_uniform(const mfo_ctx *c, uint64_t *out);
void mfo_sample_gaussian(const mfo_ctx *c, uint64_t *out);
void mfo_sample_ternary(const mfo_ctx *c, uint64_t *out);     /* [w][l][x] */
/* HE.cu:1272-1307: sk [w][l][x] in W-eval + X-NTT (Phantom order) */
void mfo_keygen(const mfo_ctx *c, uint64_t *sk);
/* ntt_core.cu:443-460: in-place n-point Phantom NTT of cnt polys [cnt][L][n] */
void mfo_xy_ntt_forward(const mfo_ctx *c, uint64_t *data, size_t cnt);
void mfo_xy_ntt_backward(const mfo_ctx *c, uint64_t *data, size_t cnt);
/* HE.cu:1455-1552 with caller-supplied a (W-coeff, matrix-major) and errors; ct = [b | a] */
void mfo_encrypt_pair_with(const mfo_ctx *c, const uint64_t *m_re, const uint64_t *m_im, const uint64_t *sk,
                           const uint64_t *a_coeff, const uint64_t *e_re, const uint64_t *e_im,
                           uint64_t *ct_re, uint64_t *ct_im);
void mfo_encrypt_pair(const mfo_ctx *c, const uint64_t *m_re, const uint64_t *m_im, const uint64_t *sk,
                      uint64_t *ct_re, uint64_t *ct_im);
/* HE.cu:1370-1453 */
void mfo_encrypt(const mfo_ctx *c, const uint64_t *m, const uint64_t *sk, uint64_t *ct);
/* HE.cu:1553-1601 (poly-major out) */
void mfo_decrypt_to_eval(const mfo_ctx *c, const uint64_t *ct, const uint64_t *sk, uint64_t *out_poly);
/* HE.cu:1691-1708 */
void mfo_decrypt_and_decode(const mfo_ctx *c, const uint64_t *ct_re, const uint64_t *ct_im, const uint64_t *sk, mfo_cx *out);
/* HE.cu:631-669, 1710-1740 (modulus index exactly as the reference computes it: (idx / n) % L) */
void mfo_ct_add(const mfo_ctx *c, const uint64_t *ct1, const uint64_t *ct2, uint64_t *res);
void mfo_ct_mul_raw(const mfo_ctx *c, const uint64_t *ct1, const uint64_t *ct2, uint64_t *d0, uint64_t *d1, uint64_t *d2);

/* ---- trace step: trace.cu:30-161, batched_trace.cu:37-197 -------------------------------- */
void mfo_trace_map(const mfo_ctx *c, const uint64_t *Br, const uint64_t *Bi, uint64_t *Bpr, uint64_t *Bpi, int batch);
void mfo_trace_gemm(const mfo_ctx *c, const uint64_t *Ar, const uint64_t *Ai, const uint64_t *Bpr, const uint64_t *Bpi,
                    uint64_t *Cr, uint64_t *Ci, int batch);
void mfo_trace_rescale(const mfo_ctx *c, uint64_t *Cr, uint64_t *Ci, int batch, uint64_t inv0, uint64_t inv1, uint64_t inv2);
void mfo_rescale_last(const mfo_ctx *c, const uint64_t *in, uint64_t *out);
/* encrypted matrix product */
void mfo_trace_map_x(const mfo_ctx *c, const uint64_t *Br, const uint64_t *Bi, uint64_t *Bpr, uint64_t *Bpi, int batch);
void mfo_trace_gemm_ty(const mfo_ctx *c, const uint64_t *Ur, const uint64_t *Ui, const uint64_t *Vr, const uint64_t *Vi,
                       uint64_t *Dr, uint64_t *Di, int batch);
void mfo_ct_matmul(const mfo_ctx *c, const uint64_t *ctU_re, const uint64_t *ctU_im, const uint64_t *ctV_re,
                   const uint64_t *ctV_im, uint64_t *out);
void mfo_ct_matmul_decrypt(const mfo_ctx *c, const uint64_t *prod, const uint64_t *sk, uint64_t *out_re, uint64_t *out_im);

/* ---- Phantom base conversion: ph/src/rns_bconv.cu:22-29,109-131,212-229 ------------------ */
/* in [L_in][N], out [L_out][N];  qhat_inv[i] = (Q/q_i)^-1 mod q_i, qhat_mod_p[j][i] = Q/q_i mod p_j */
void mfo_bconv(const uint64_t *in, uint64_t *out, size_t N, int L_in, const uint64_t *q_in,
               int L_out, const uint64_t *p_out);

#ifdef __cplusplus
}
#endif
#endif
