/*
 * scheme_oracle.c -- XY embedding, quantise / CRT, encode / decode, samplers, keygen,
 * encrypt / decrypt, ciphertext ops, trace step, base conversion.
 * TEST INFRASTRUCTURE ONLY (see mfhe_oracle.h / scheme_oracle.h).
 */
#include "mfhe_oracle.h"
#include "scheme_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

typedef unsigned __int128 u128;
int mfo_big_cmp(const uint64_t *a, const uint64_t *b, int W);
void mfo_big_compose(const mfo_ctx *c, const uint64_t *res, int limbs, uint64_t *mag, int *neg);

/* encoder.cu:318-326 mat_mul_kernel_complex: C = A * B, cuCadd(sum, cuCmul(a, b)) in k order */
static void matmul_cx(const mfo_cx *A, const mfo_cx *B, mfo_cx *C, int n) {
    for (int r = 0; r < n; ++r)
        for (int col = 0; col < n; ++col) {
            double sr = 0, si = 0;
            for (int k = 0; k < n; ++k) {
                mfo_cx a = A[r * n + k], b = B[k * n + col];
                double pr = a.re * b.re - a.im * b.im, pi = a.re * b.im + a.im * b.re;
                sr += pr; si += pi;
            }
            C[r * n + col].re = sr; C[r * n + col].im = si;
        }
}

/* encoder.cu:460-467 */
void mfo_xy_idft2(const mfo_ctx *c, const mfo_cx *in, mfo_cx *out, int lanes) {
    const int n = c->n; const size_t nn = (size_t)n * n;
#pragma omp parallel for schedule(static)
    for (long ell = 0; ell < lanes; ++ell) {
        mfo_cx *T = (mfo_cx *)malloc(nn * sizeof(mfo_cx));
        matmul_cx(c->xyVinv, in + (size_t)ell * nn, T, n);
        matmul_cx(T, c->xyVinvT, out + (size_t)ell * nn, n);
        free(T);
    }
}
/* encoder.cu:492-501 */
void mfo_xy_dft2(const mfo_ctx *c, const mfo_cx *in, mfo_cx *out, int lanes) {
    const int n = c->n; const size_t nn = (size_t)n * n;
#pragma omp parallel for schedule(static)
    for (long ell = 0; ell < lanes; ++ell) {
        mfo_cx *T = (mfo_cx *)malloc(nn * sizeof(mfo_cx));
        matmul_cx(c->xyV, in + (size_t)ell * nn, T, n);
        matmul_cx(T, c->xyVT, out + (size_t)ell * nn, n);
        free(T);
    }
}

/* batched_encoder.cu:125-152 / encoder.cu:36-50 */
void mfo_quantize(const mfo_ctx *c, const mfo_cx *in, uint64_t *ore, uint64_t *oim, int lanes) {
    const size_t n2 = (size_t)c->n * c->n, L = (size_t)c->L;
#pragma omp parallel for schedule(static)
    for (long r = 0; r < lanes; ++r)
        for (size_t pos = 0; pos < n2; ++pos) {
            mfo_cx z = in[(size_t)r * n2 + pos];
            int64_t xr = llround(z.re * c->delta), xi = llround(z.im * c->delta);
            for (size_t l = 0; l < L; ++l) {
                int64_t q = (int64_t)c->moduli[l], mr = xr % q, mi = xi % q;
                if (mr < 0) mr += q;
                if (mi < 0) mi += q;
                ore[((size_t)r * L + l) * n2 + pos] = (uint64_t)mr;
                oim[((size_t)r * L + l) * n2 + pos] = (uint64_t)mi;
            }
        }
}

/* encoder.cu:191-230 */
void mfo_crt_compose_big(const mfo_ctx *c, const uint64_t *in, uint64_t *mag, uint8_t *neg, int n2, int limbs) {
    const int W = c->big_words;
#pragma omp parallel for schedule(static)
    for (long idx = 0; idx < n2; ++idx) {
        uint64_t res[MFO_MAX_LIMBS], m[MFO_MAX_BIG]; int ng;
        for (int k = 0; k < limbs; ++k) res[k] = in[(size_t)k * n2 + idx];
        mfo_big_compose(c, res, limbs, m, &ng);
        memcpy(mag + (size_t)idx * W, m, (size_t)W * 8);
        neg[idx] = (uint8_t)ng;
    }
}
/* encoder.cu:152-189: keeps only the low word */
void mfo_crt_compose_i64(const mfo_ctx *c, const uint64_t *in, int64_t *out, int n2, int limbs) {
    for (long idx = 0; idx < n2; ++idx) {
        uint64_t res[MFO_MAX_LIMBS], m[MFO_MAX_BIG]; int ng;
        for (int k = 0; k < limbs; ++k) res[k] = in[(size_t)k * n2 + idx];
        mfo_big_compose(c, res, limbs, m, &ng);
        int64_t v = (int64_t)m[0];
        out[idx] = ng ? -v : v;
    }
}
/* encoder.cu:103-110 / HE.cu:917-924: Horner in 2^64 */
static double big_to_f64(const uint64_t *a, int W) {
    const double two64 = 18446744073709551616.0;
    double v = 0.0;
    for (int i = W - 1; i >= 0; --i) v = v * two64 + (double)a[i];
    return v;
}
/* encoder.cu:112-150 */
void mfo_dequantize_exact(const mfo_ctx *c, const uint64_t *re, const uint64_t *im, mfo_cx *out, int n2, int limbs) {
    const int W = c->big_words;
#pragma omp parallel for schedule(static)
    for (long idx = 0; idx < n2; ++idx) {
        uint64_t res[MFO_MAX_LIMBS], m[MFO_MAX_BIG]; int ng;
        for (int k = 0; k < limbs; ++k) res[k] = re[(size_t)k * n2 + idx];
        mfo_big_compose(c, res, limbs, m, &ng);
        double v = big_to_f64(m, W) / c->delta; out[idx].re = ng ? -v : v;
        for (int k = 0; k < limbs; ++k) res[k] = im[(size_t)k * n2 + idx];
        mfo_big_compose(c, res, limbs, m, &ng);
        v = big_to_f64(m, W) / c->delta; out[idx].im = ng ? -v : v;
    }
}
/* HE.cu:1007-1027 */
void mfo_big_pair_to_complex(const mfo_ctx *c, const uint64_t *re_mag, const uint8_t *re_neg,
                             const uint64_t *im_mag, const uint8_t *im_neg, mfo_cx *out, size_t total) {
    const int W = c->big_words;
    for (size_t i = 0; i < total; ++i) {
        double vr = big_to_f64(re_mag + i * W, W), vi = big_to_f64(im_mag + i * W, W);
        if (re_neg[i]) vr = -vr;
        if (im_neg[i]) vi = -vi;
        out[i].re = vr / c->delta; out[i].im = vi / c->delta;
    }
}

/* batched_encoder.cu:161-228 */
void mfo_encode(const mfo_ctx *c, const mfo_cx *msg, uint64_t *out_re, uint64_t *out_im) {
    const size_t n2 = (size_t)c->n * c->n, phi = (size_t)c->phi, T = phi * (size_t)c->L * n2;
    mfo_cx *xy = (mfo_cx *)malloc(phi * n2 * sizeof(mfo_cx)), *wc = (mfo_cx *)malloc(phi * n2 * sizeof(mfo_cx));
    uint64_t *cr = (uint64_t *)malloc(T * 8), *ci = (uint64_t *)malloc(T * 8), *ev = (uint64_t *)malloc(T * 8);
    mfo_xy_idft2(c, msg, xy, (int)phi);
    mfo_w_idft(c, xy, wc);
    mfo_quantize(c, wc, cr, ci, (int)phi);
    mfo_wntt_forward_matrix(c, cr, ev); mfo_poly_to_matrix(c, ev, out_re);
    mfo_wntt_forward_matrix(c, ci, ev); mfo_poly_to_matrix(c, ev, out_im);
    free(xy); free(wc); free(cr); free(ci); free(ev);
}
/* encoder.cu:446-456 */
void mfo_encode_lane(const mfo_ctx *c, const mfo_cx *msg, uint64_t *out_re, uint64_t *out_im) {
    const size_t n2 = (size_t)c->n * c->n;
    mfo_cx *P = (mfo_cx *)malloc(n2 * sizeof(mfo_cx));
    mfo_xy_idft2(c, msg, P, 1);
    mfo_quantize(c, P, out_re, out_im, 1);
    free(P);
}
/* HE.cu:1619-1689 */
void mfo_decode_eval_pair(const mfo_ctx *c, const uint64_t *eval_re, const uint64_t *eval_im, mfo_cx *out) {
    const size_t n2 = (size_t)c->n * c->n, phi = (size_t)c->phi, L = (size_t)c->L, T = phi * L * n2;
    const int W = c->big_words;
    uint64_t *cr = (uint64_t *)malloc(T * 8), *ci = (uint64_t *)malloc(T * 8);
    uint64_t *mr = (uint64_t *)malloc(phi * n2 * W * 8), *mi = (uint64_t *)malloc(phi * n2 * W * 8);
    uint8_t *nr = (uint8_t *)malloc(phi * n2), *ni = (uint8_t *)malloc(phi * n2);
    mfo_cx *cc = (mfo_cx *)malloc(phi * n2 * sizeof(mfo_cx)), *ec = (mfo_cx *)malloc(phi * n2 * sizeof(mfo_cx));
    mfo_wntt_inverse_matrix(c, eval_re, cr);
    mfo_wntt_inverse_matrix(c, eval_im, ci);
    for (size_t ell = 0; ell < phi; ++ell) {
        mfo_crt_compose_big(c, cr + ell * L * n2, mr + ell * n2 * W, nr + ell * n2, (int)n2, (int)L);
        mfo_crt_compose_big(c, ci + ell * L * n2, mi + ell * n2 * W, ni + ell * n2, (int)n2, (int)L);
    }
    mfo_big_pair_to_complex(c, mr, nr, mi, ni, cc, phi * n2);
    mfo_wdft_forward_complex(c, cc, ec);
    mfo_xy_dft2(c, ec, out, (int)phi);
    free(cr); free(ci); free(mr); free(mi); free(nr); free(ni); free(cc); free(ec);
}
/* encoder.cu:469-490 */
void mfo_decode_lane_from_rns_eval(const mfo_ctx *c, const uint64_t *re, const uint64_t *im, mfo_cx *out) {
    const size_t n2 = (size_t)c->n * c->n;
    mfo_cx *P = (mfo_cx *)malloc(n2 * sizeof(mfo_cx));
    mfo_dequantize_exact(c, re, im, P, (int)n2, c->L);
    mfo_xy_dft2(c, P, out, 1);
    free(P);
}

/* ---- samplers ------------------------------------------------------------------------------ */
/* HE.cu:564-578 */
void mfo_sample_uniform(const mfo_ctx *c, uint64_t *out) {
    const size_t n2 = (size_t)c->n * c->n, L = (size_t)c->L, total = (size_t)c->phi * L * n2;
    for (size_t idx = 0; idx < total; ++idx) {
        uint64_t q = c->moduli[(idx % (L * n2)) / n2];
        uint64_t seed = 123456789ULL + idx;
        seed = seed * 6364136223846793005ULL + 1442695040888963407ULL;
        out[idx] = seed % q;
    }
}
static uint64_t splitmix64(uint64_t x) {       /* HE.cu:581-586 */
    x += 0x9E3779B97F4A7C15ULL;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
    return x ^ (x >> 31);
}
/* HE.cu:590-627 */
void mfo_sample_gaussian(const mfo_ctx *c, uint64_t *out) {
    const size_t n2 = (size_t)c->n * c->n, L = (size_t)c->L, total = (size_t)c->phi * L * n2;
    for (size_t idx = 0; idx < total; ++idx) {
        size_t w = idx / (L * n2), off = idx % (L * n2), l = off / n2, pos = off % n2;
        uint64_t q = c->moduli[l];
        uint64_t coeff_id = (uint64_t)(w * n2 + pos);
        uint64_t r1 = splitmix64(0xD6E8FEB86659FD93ULL ^ coeff_id), r2 = splitmix64(r1);
        const double inv53 = 1.0 / 9007199254740992.0;
        double u1 = ((double)(r1 >> 11) + 1.0) * inv53, u2 = ((double)(r2 >> 11) + 1.0) * inv53;
        double mag = 3.2 * sqrt(-2.0 * log(u1));
        double z = mag * cos(6.283185307179586 * u2);
        int64_t noise = llround(z);
        out[idx] = noise >= 0 ? (uint64_t)noise : q - (uint64_t)(-noise);
    }
}
/* HE.cu:690-713: layout [poly = w][limb][coeff] */
void mfo_sample_ternary(const mfo_ctx *c, uint64_t *out) {
    const size_t n = (size_t)c->n, L = (size_t)c->L, total = (size_t)c->phi * L * n;
    for (size_t idx = 0; idx < total; ++idx) {
        size_t off = idx % (L * n), limb = off / n, coeff = off - limb * n, poly = idx / (L * n);
        uint64_t t = (uint64_t)poly * 1315423911ULL + (uint64_t)coeff * 2654435761ULL;
        int r = (int)((t * 11400714819323198485ULL) % 3ULL);
        out[idx] = r == 0 ? 0 : r == 1 ? 1 : c->moduli[limb] - 1;
    }
}

/* ntt_core.cu:443-460 -> ph/src/ntt/ntt_1d.cu */
void mfo_xy_ntt_forward(const mfo_ctx *c, uint64_t *data, size_t cnt) {
    const size_t n = (size_t)c->n, L = (size_t)c->L;
#pragma omp parallel for schedule(static)
    for (long i = 0; i < (long)(cnt * L); ++i) {
        size_t l = (size_t)i % L;
        mfo_negacyclic_forward(data + (size_t)i * n, c->logn, c->moduli[l], c->tw + l * n, c->tws + l * n);
    }
}
void mfo_xy_ntt_backward(const mfo_ctx *c, uint64_t *data, size_t cnt) {
    const size_t n = (size_t)c->n, L = (size_t)c->L;
#pragma omp parallel for schedule(static)
    for (long i = 0; i < (long)(cnt * L); ++i) {
        size_t l = (size_t)i % L;
        mfo_negacyclic_inverse(data + (size_t)i * n, c->logn, c->moduli[l], c->itw + l * n, c->itws + l * n,
                               c->ninv[l], c->ninvs[l]);
    }
}
/* HE.cu:1272-1307 */
void mfo_keygen(const mfo_ctx *c, uint64_t *sk) {
    const size_t total = (size_t)c->phi * c->L * c->n;
    uint64_t *s = (uint64_t *)malloc(total * 8);
    mfo_sample_ternary(c, s);
    mfo_wntt_forward_vector(c, s, sk);
    mfo_xy_ntt_forward(c, sk, (size_t)c->phi);
    free(s);
}

/* ---- encrypt / decrypt ------------------------------------------------------------------------ */
/* HE.cu:509-531: t = a_ntt (.) s_ntt, s broadcast over y; poly-major data */
static void pointwise_mul_s(const mfo_ctx *c, const uint64_t *a, const uint64_t *s, uint64_t *t) {
    const size_t n = (size_t)c->n, L = (size_t)c->L, total = (size_t)c->phi * n * L * n;
#pragma omp parallel for schedule(static)
    for (long i = 0; i < (long)total; ++i) {
        size_t idx = (size_t)i, off = idx % (L * n), l = off / n, x = off - l * n, poly = idx / (L * n), w = poly / n;
        t[idx] = mfo_mulmod(a[idx], s[(w * L + l) * n + x], c->moduli[l]);
    }
}

/* HE.cu:1455-1552 with caller-supplied a (W-coeff, matrix-major) and errors (W-coeff) */
void mfo_encrypt_pair_with(const mfo_ctx *c, const uint64_t *m_re, const uint64_t *m_im, const uint64_t *sk,
                           const uint64_t *a_coeff, const uint64_t *e_re, const uint64_t *e_im,
                           uint64_t *ct_re, uint64_t *ct_im) {
    const size_t n = (size_t)c->n, L = (size_t)c->L, T = (size_t)c->phi * L * n * n;
    uint64_t *mp = (uint64_t *)malloc(T * 8), *a_eval = (uint64_t *)malloc(T * 8), *t = (uint64_t *)malloc(T * 8);
    uint64_t *e_eval = (uint64_t *)malloc(T * 8), *b = (uint64_t *)malloc(T * 8);
    mfo_wntt_forward_matrix(c, a_coeff, a_eval);
    memcpy(t, a_eval, T * 8);
    mfo_xy_ntt_forward(c, t, (size_t)c->phi * n);
    pointwise_mul_s(c, t, sk, t);
    mfo_xy_ntt_backward(c, t, (size_t)c->phi * n);
    for (int part = 0; part < 2; ++part) {
        const uint64_t *m = part ? m_im : m_re, *e = part ? e_im : e_re;
        uint64_t *ct = part ? ct_im : ct_re;
        if (!ct) continue;
        mfo_matrix_to_poly(c, m, mp);
        mfo_wntt_forward_matrix(c, e, e_eval);
        for (size_t idx = 0; idx < T; ++idx) {           /* HE.cu:535-547 combine_b */
            uint64_t q = c->moduli[(idx % (L * n)) / n];
            b[idx] = mfo_addmod(mfo_submod(mp[idx], t[idx], q), e_eval[idx], q);
        }
        mfo_poly_to_matrix(c, b, ct);
        mfo_poly_to_matrix(c, a_eval, ct + T);
    }
    free(mp); free(a_eval); free(t); free(e_eval); free(b);
}
void mfo_encrypt_pair(const mfo_ctx *c, const uint64_t *m_re, const uint64_t *m_im, const uint64_t *sk,
                      uint64_t *ct_re, uint64_t *ct_im) {
    const size_t T = (size_t)c->phi * c->L * c->n * c->n;
    uint64_t *a = (uint64_t *)malloc(T * 8), *e = (uint64_t *)malloc(T * 8);
    mfo_sample_uniform(c, a);
    mfo_sample_gaussian(c, e);           /* the same stream for e_re and e_im: HE.cu:1516-1517 */
    mfo_encrypt_pair_with(c, m_re, m_im, sk, a, e, e, ct_re, ct_im);
    free(a); free(e);
}
/* HE.cu:1370-1453 */
void mfo_encrypt(const mfo_ctx *c, const uint64_t *m, const uint64_t *sk, uint64_t *ct) {
    const size_t T = (size_t)c->phi * c->L * c->n * c->n;
    uint64_t *a = (uint64_t *)malloc(T * 8), *e = (uint64_t *)malloc(T * 8);
    mfo_sample_uniform(c, a);
    mfo_sample_gaussian(c, e);
    mfo_encrypt_pair_with(c, m, NULL, sk, a, e, NULL, ct, NULL);
    free(a); free(e);
}
/* HE.cu:1553-1601 */
void mfo_decrypt_to_eval(const mfo_ctx *c, const uint64_t *ct, const uint64_t *sk, uint64_t *out_poly) {
    const size_t n = (size_t)c->n, L = (size_t)c->L, T = (size_t)c->phi * L * n * n;
    uint64_t *bp = (uint64_t *)malloc(T * 8), *t = (uint64_t *)malloc(T * 8);
    mfo_matrix_to_poly(c, ct, bp);
    mfo_matrix_to_poly(c, ct + T, t);
    mfo_xy_ntt_forward(c, t, (size_t)c->phi * n);
    pointwise_mul_s(c, t, sk, t);
    mfo_xy_ntt_backward(c, t, (size_t)c->phi * n);
    for (size_t idx = 0; idx < T; ++idx) out_poly[idx] = mfo_addmod(bp[idx], t[idx], c->moduli[(idx % (L * n)) / n]);
    free(bp); free(t);
}
/* HE.cu:1691-1708 */
void mfo_decrypt_and_decode(const mfo_ctx *c, const uint64_t *ct_re, const uint64_t *ct_im, const uint64_t *sk, mfo_cx *out) {
    const size_t T = (size_t)c->phi * c->L * c->n * c->n;
    uint64_t *er = (uint64_t *)malloc(T * 8), *ei = (uint64_t *)malloc(T * 8);
    mfo_decrypt_to_eval(c, ct_re, sk, er);
    mfo_decrypt_to_eval(c, ct_im, sk, ei);
    mfo_decode_eval_pair(c, er, ei, out);
    free(er); free(ei);
}

/* HE.cu:631-644 + 1710-1725.  The kernels take the modulus of flat index idx as
 * limb = (idx % (L n)) / n, whatever the layout of the data is. */
void mfo_ct_add(const mfo_ctx *c, const uint64_t *ct1, const uint64_t *ct2, uint64_t *res) {
    const size_t n = (size_t)c->n, L = (size_t)c->L, T = (size_t)c->phi * L * n * n;
    for (size_t idx = 0; idx < T; ++idx) {
        uint64_t q = c->moduli[(idx % (L * n)) / n];
        res[idx] = mfo_addmod(ct1[idx], ct2[idx], q);
        res[T + idx] = mfo_addmod(ct1[T + idx], ct2[T + idx], q);
    }
}
/* HE.cu:647-669 + 1727-1740 */
void mfo_ct_mul_raw(const mfo_ctx *c, const uint64_t *ct1, const uint64_t *ct2, uint64_t *d0, uint64_t *d1, uint64_t *d2) {
    const size_t n = (size_t)c->n, L = (size_t)c->L, T = (size_t)c->phi * L * n * n;
#pragma omp parallel for schedule(static)
    for (long i = 0; i < (long)T; ++i) {
        size_t idx = (size_t)i;
        uint64_t q = c->moduli[(idx % (L * n)) / n];
        uint64_t b1 = ct1[idx], a1 = ct1[T + idx], b2 = ct2[idx], a2 = ct2[T + idx];
        d0[idx] = mfo_mulmod(b1, b2, q);
        d1[idx] = mfo_addmod(mfo_mulmod(b1, a2, q), mfo_mulmod(a1, b2, q), q);
        d2[idx] = mfo_mulmod(a1, a2, q);
    }
}

/* ---- trace step --------------------------------------------------------------------------------- */
static uint64_t neg_mod(uint64_t a, uint64_t q) { return a == 0 ? 0 : q - a; }
/* trace.cu:30-62, batched_trace.cu:37-79: data [batch][L][n][n] */
void mfo_trace_map(const mfo_ctx *c, const uint64_t *Br, const uint64_t *Bi, uint64_t *Bpr, uint64_t *Bpi, int batch) {
    const int n = c->n; const size_t n2 = (size_t)n * n, L = (size_t)c->L;
    for (size_t b = 0; b < (size_t)batch; ++b)
        for (size_t l = 0; l < L; ++l) {
            uint64_t q = c->moduli[l]; size_t base = (b * L + l) * n2;
            for (int j = 0; j < n; ++j)
                for (int k = 0; k < n; ++k) {
                    size_t src = base + (size_t)j * n + k, dst = base + (size_t)((n - j) & (n - 1)) * n + k;
                    uint64_t a = Br[src], bb = Bi[src];
                    if (j == 0) { Bpr[dst] = a; Bpi[dst] = neg_mod(bb, q); }
                    else { Bpr[dst] = neg_mod(bb, q); Bpi[dst] = neg_mod(a, q); }
                }
        }
}
/* trace.cu:77-120, batched_trace.cu:99-146: C = n * A * Bp^T (complex), per limb */
void mfo_trace_gemm(const mfo_ctx *c, const uint64_t *Ar, const uint64_t *Ai, const uint64_t *Bpr, const uint64_t *Bpi,
                    uint64_t *Cr, uint64_t *Ci, int batch) {
    const int n = c->n; const size_t n2 = (size_t)n * n, L = (size_t)c->L;
#pragma omp parallel for schedule(static)
    for (long bl = 0; bl < (long)((size_t)batch * L); ++bl) {
        size_t l = (size_t)bl % L, base = (size_t)bl * n2;
        uint64_t q = c->moduli[l], nm = (uint64_t)n % q;
        for (int row = 0; row < n; ++row)
            for (int col = 0; col < n; ++col) {
                uint64_t ar_ = 0, ai_ = 0;
                for (int t = 0; t < n; ++t) {
                    uint64_t ar = Ar[base + (size_t)row * n + t], ai = Ai[base + (size_t)row * n + t];
                    uint64_t br = Bpr[base + (size_t)col * n + t], bi = Bpi[base + (size_t)col * n + t];
                    uint64_t pr = mfo_submod(mfo_mulmod(ar, br, q), mfo_mulmod(ai, bi, q), q);
                    uint64_t pi = mfo_addmod(mfo_mulmod(ar, bi, q), mfo_mulmod(ai, br, q), q);
                    ar_ = mfo_addmod(ar_, pr, q); ai_ = mfo_addmod(ai_, pi, q);
                }
                Cr[base + (size_t)row * n + col] = mfo_mulmod(ar_, nm, q);
                Ci[base + (size_t)row * n + col] = mfo_mulmod(ai_, nm, q);
            }
    }
}
/* trace.cu:132-161, batched_trace.cu:163-197: limbs >= 3 are multiplied by 0 */
void mfo_trace_rescale(const mfo_ctx *c, uint64_t *Cr, uint64_t *Ci, int batch, uint64_t inv0, uint64_t inv1, uint64_t inv2) {
    const size_t n2 = (size_t)c->n * c->n, L = (size_t)c->L;
    for (size_t bl = 0; bl < (size_t)batch * L; ++bl) {
        size_t l = bl % L; uint64_t q = c->moduli[l], inv = l == 0 ? inv0 : l == 1 ? inv1 : l == 2 ? inv2 : 0;
        for (size_t i = 0; i < n2; ++i) {
            Cr[bl * n2 + i] = mfo_mulmod(Cr[bl * n2 + i], inv, q);
            Ci[bl * n2 + i] = mfo_mulmod(Ci[bl * n2 + i], inv, q);
        }
    }
}

/* ---- encrypted matrix product: SURVEY 8(f) rank 1 (no reference caller exists; restated from the definitions the
 * reference's pieces give: trace.cu:30-62 for the twist, trace.cu:77-120 for the Gaussian-integer GEMM and its factor n,
 * HE.cu:1553-1601 for "m = b + a (*) s" as the meaning of a ciphertext).  Operands: poly sets [w][l][y][x]. ---------- */
/* the map of trace.cu:30-62 with the twisted index being x (fast): B'[y][(-x) mod n] = conj(B[y][x]) * (x == 0 ? 1 : -i) */
void mfo_trace_map_x(const mfo_ctx *c, const uint64_t *Br, const uint64_t *Bi, uint64_t *Bpr, uint64_t *Bpi, int batch) {
    const int n = c->n; const size_t n2 = (size_t)n * n, L = (size_t)c->L;
    for (size_t b = 0; b < (size_t)batch; ++b)
        for (size_t l = 0; l < L; ++l) {
            uint64_t q = c->moduli[l]; size_t base = (b * L + l) * n2;
            for (int y = 0; y < n; ++y)
                for (int x = 0; x < n; ++x) {
                    size_t src = base + (size_t)y * n + x, dst = base + (size_t)y * n + ((n - x) & (n - 1));
                    uint64_t a = Br[src], bb = Bi[src];
                    if (x == 0) { Bpr[dst] = a; Bpi[dst] = neg_mod(bb, q); }
                    else { Bpr[dst] = neg_mod(bb, q); Bpi[dst] = neg_mod(a, q); }
                }
        }
}
/* D[x][z] = n * sum_y U[y][x] * V[y][z]  over Z_q[i], per (batch, limb) */
void mfo_trace_gemm_ty(const mfo_ctx *c, const uint64_t *Ur, const uint64_t *Ui, const uint64_t *Vr, const uint64_t *Vi,
                       uint64_t *Dr, uint64_t *Di, int batch) {
    const int n = c->n; const size_t n2 = (size_t)n * n, L = (size_t)c->L;
#pragma omp parallel for schedule(static)
    for (long bl = 0; bl < (long)((size_t)batch * L); ++bl) {
        size_t l = (size_t)bl % L, base = (size_t)bl * n2;
        uint64_t q = c->moduli[l], nm = (uint64_t)n % q;
        for (int x = 0; x < n; ++x)
            for (int z = 0; z < n; ++z) {
                uint64_t sr = 0, si = 0;
                for (int y = 0; y < n; ++y) {
                    uint64_t ur = Ur[base + (size_t)y * n + x], ui = Ui[base + (size_t)y * n + x];
                    uint64_t vr = Vr[base + (size_t)y * n + z], vi = Vi[base + (size_t)y * n + z];
                    sr = mfo_addmod(sr, mfo_submod(mfo_mulmod(ur, vr, q), mfo_mulmod(ui, vi, q), q), q);
                    si = mfo_addmod(si, mfo_addmod(mfo_mulmod(ur, vi, q), mfo_mulmod(ui, vr, q), q), q);
                }
                Dr[base + (size_t)x * n + z] = mfo_mulmod(sr, nm, q);
                Di[base + (size_t)x * n + z] = mfo_mulmod(si, nm, q);
            }
    }
}
/* out = 8 poly sets [bb re, im | ab re, im | ba re, im | aa re, im] */
void mfo_ct_matmul(const mfo_ctx *c, const uint64_t *ctU_re, const uint64_t *ctU_im, const uint64_t *ctV_re,
                   const uint64_t *ctV_im, uint64_t *out) {
    const size_t T = (size_t)c->phi * c->L * c->n * c->n;
    for (int k = 0; k < 4; ++k) {
        size_t uo = (k & 1) ? T : 0, vo = (k & 2) ? T : 0;
        mfo_trace_gemm_ty(c, ctU_re + uo, ctU_im + uo, ctV_re + vo, ctV_im + vo, out + (size_t)(2 * k) * T,
                          out + (size_t)(2 * k + 1) * T, c->phi);
    }
}
/* o = m + a (*) s along the fast index of matrix-major data, s = key [w][l][x] in NTT form (HE.cu:1579-1590) */
static void rows_add_as(const mfo_ctx *c, const uint64_t *m, const uint64_t *a, const uint64_t *sk, uint64_t *o) {
    const size_t n = (size_t)c->n, L = (size_t)c->L, rows = (size_t)c->phi * L * n;
#pragma omp parallel for schedule(static)
    for (long r = 0; r < (long)rows; ++r) {
        size_t wl = (size_t)r / n, l = wl % L;
        uint64_t q = c->moduli[l], t[256];
        memcpy(t, a + (size_t)r * n, n * 8);
        mfo_negacyclic_forward(t, c->logn, q, c->tw + l * n, c->tws + l * n);
        for (size_t x = 0; x < n; ++x) t[x] = mfo_mulmod(t[x], sk[wl * n + x], q);
        mfo_negacyclic_inverse(t, c->logn, q, c->itw + l * n, c->itws + l * n, c->ninv[l], c->ninvs[l]);
        for (size_t x = 0; x < n; ++x) o[(size_t)r * n + x] = mfo_addmod(m[(size_t)r * n + x], t[x], q);
    }
}
static void transpose_set(const mfo_ctx *c, const uint64_t *in, uint64_t *out) {
    const size_t n = (size_t)c->n, mats = (size_t)c->phi * c->L;
    for (size_t mtx = 0; mtx < mats; ++mtx)
        for (size_t i = 0; i < n; ++i)
            for (size_t j = 0; j < n; ++j) out[(mtx * n + j) * n + i] = in[(mtx * n + i) * n + j];
}
/* D_bb + D_ab (*)_x s + D_ba (*)_z s + D_aa (*)_x s (*)_z s, returned as [w][l][z][x] */
void mfo_ct_matmul_decrypt(const mfo_ctx *c, const uint64_t *prod, const uint64_t *sk, uint64_t *out_re, uint64_t *out_im) {
    const size_t T = (size_t)c->phi * c->L * c->n * c->n;
    uint64_t *acc = (uint64_t *)malloc(T * 8), *e = (uint64_t *)malloc(T * 8), *at = (uint64_t *)malloc(T * 8), *et = (uint64_t *)malloc(T * 8);
    for (int part = 0; part < 2; ++part) {
        rows_add_as(c, prod + (size_t)(0 + part) * T, prod + (size_t)(4 + part) * T, sk, acc);   /* bb + ba (*)_z s */
        rows_add_as(c, prod + (size_t)(2 + part) * T, prod + (size_t)(6 + part) * T, sk, e);     /* ab + aa (*)_z s */
        transpose_set(c, acc, at); transpose_set(c, e, et);
        rows_add_as(c, at, et, sk, part ? out_im : out_re);
    }
    free(acc); free(e); free(at); free(et);
}

/* ---- RNS rescale of the matrix scheme: X' = (X - [X]_{q_last}) / q_last on the W-coefficients,
 * stated per column with the definitions of the W-CRT (HE.cu:716-781): in [w][L][n^2] -> out [w][L-1][n^2] ---------- */
void mfo_rescale_last(const mfo_ctx *c, const uint64_t *in, uint64_t *out) {
    const size_t phi = (size_t)c->phi, L = (size_t)c->L, cols = (size_t)c->n * c->n;
    const uint64_t ql = c->moduli[L - 1];
    const uint64_t *Vi = c->wntt_inv + (L - 1) * phi * phi;          /* [w][r] = V^-1[r][w] */
#pragma omp parallel for schedule(static)
    for (long col = 0; col < (long)cols; ++col) {
        int64_t cen[1024];
        for (size_t r = 0; r < phi; ++r) {
            uint64_t acc = 0;
            for (size_t w = 0; w < phi; ++w)
                acc = mfo_addmod(acc, mfo_mulmod(in[(w * L + (L - 1)) * cols + (size_t)col], Vi[w * phi + r], ql), ql);
            cen[r] = acc > (ql >> 1) ? (int64_t)acc - (int64_t)ql : (int64_t)acc;
        }
        for (size_t l = 0; l + 1 < L; ++l) {
            const uint64_t q = c->moduli[l], qinv = mfo_invmod(ql % q, q);
            const uint64_t *V = c->wntt + l * phi * phi;             /* [w][r] */
            for (size_t w = 0; w < phi; ++w) {
                uint64_t acc = 0;
                for (size_t r = 0; r < phi; ++r) {
                    int64_t m = cen[r] % (int64_t)q; if (m < 0) m += (int64_t)q;
                    acc = mfo_addmod(acc, mfo_mulmod((uint64_t)m, V[w * phi + r], q), q);
                }
                out[(w * (L - 1) + l) * cols + (size_t)col] = mfo_mulmod(mfo_submod(in[(w * L + l) * cols + (size_t)col], acc, q), qinv, q);
            }
        }
    }
}

/* ---- Phantom bConv_BEHZ: ph/src/rns_bconv.cu:22-29, 109-131, 212-229 ------------------------------ */
void mfo_bconv(const uint64_t *in, uint64_t *out, size_t N, int L_in, const uint64_t *q_in, int L_out, const uint64_t *p_out) {
    uint64_t *qhat_inv = (uint64_t *)malloc((size_t)L_in * 8), *qhat_mod_p = (uint64_t *)malloc((size_t)L_in * L_out * 8);
    for (int i = 0; i < L_in; ++i) {
        uint64_t pr = 1;
        for (int k = 0; k < L_in; ++k) if (k != i) pr = mfo_mulmod(pr, q_in[k] % q_in[i], q_in[i]);
        qhat_inv[i] = mfo_invmod(pr, q_in[i]);
        for (int j = 0; j < L_out; ++j) {
            uint64_t pj = 1;
            for (int k = 0; k < L_in; ++k) if (k != i) pj = mfo_mulmod(pj, q_in[k] % p_out[j], p_out[j]);
            qhat_mod_p[(size_t)j * L_in + i] = pj;
        }
    }
#pragma omp parallel for schedule(static)
    for (long d = 0; d < (long)N; ++d) {
        uint64_t y[MFO_MAX_LIMBS * 2];
        for (int i = 0; i < L_in; ++i) y[i] = mfo_mulmod(in[(size_t)i * N + d], qhat_inv[i], q_in[i]);
        for (int j = 0; j < L_out; ++j) {
            u128 acc = 0;
            for (int i = 0; i < L_in; ++i) {
                acc += (u128)y[i] * qhat_mod_p[(size_t)j * L_in + i];
                if ((i & 7) == 7) acc %= p_out[j];     /* same residue; keeps the oracle valid for 60-bit bases */
            }
            out[(size_t)j * N + d] = (uint64_t)(acc % p_out[j]);
        }
    }
    free(qhat_inv); free(qhat_mod_p);
}
