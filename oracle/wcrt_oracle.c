/*
 * wcrt_oracle.c -- oracle context, W-CRT tables and transforms, complex W-DFT.
 * TEST INFRASTRUCTURE ONLY (see mfhe_oracle.h / scheme_oracle.h).  Plain C, gcc,
 * compiled with -ffp-contract=off so fp64 sequences round like the reference's host code.
 */
#include "mfhe_oracle.h"
#include "scheme_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

typedef unsigned __int128 u128;

/* HE.cu:119-133 h_find_eta: smallest g >= 2 with g^((q-1)/p) of exact order p = 3 f */
uint64_t mfo_find_eta(uint64_t q, int f) {
    const uint64_t p = 3ull * (uint64_t)f;
    if ((q - 1) % p) return 0;
    const uint64_t e = (q - 1) / p;
    for (uint64_t g = 2; g < q; ++g) {
        uint64_t eta = mfo_powmod(g, e, q);
        if (eta == 1) continue;
        if (mfo_powmod(eta, p, q) != 1) continue;
        if (mfo_powmod(eta, p / 3, q) == 1) continue;
        if (mfo_powmod(eta, p / (uint64_t)f, q) == 1) continue;
        return eta;
    }
    return 0;
}

/* HE.cu:135-185 matrix_inverse_mod: Gauss-Jordan, first non-zero pivot */
int mfo_matrix_inverse_mod(const uint64_t *m, uint64_t *inv, int dim, uint64_t q) {
    size_t d = (size_t)dim;
    uint64_t *a = (uint64_t *)malloc(d * d * sizeof(uint64_t));
    memcpy(a, m, d * d * sizeof(uint64_t));
    memset(inv, 0, d * d * sizeof(uint64_t));
    for (size_t i = 0; i < d; ++i) inv[i * d + i] = 1;
    for (size_t i = 0; i < d; ++i) {
        size_t piv = i;
        while (piv < d && a[piv * d + i] == 0) ++piv;
        if (piv == d) { free(a); return -1; }
        if (piv != i)
            for (size_t j = 0; j < d; ++j) {
                uint64_t t = a[i * d + j]; a[i * d + j] = a[piv * d + j]; a[piv * d + j] = t;
                t = inv[i * d + j]; inv[i * d + j] = inv[piv * d + j]; inv[piv * d + j] = t;
            }
        uint64_t pinv = mfo_invmod(a[i * d + i], q);
        for (size_t j = 0; j < d; ++j) {
            a[i * d + j] = mfo_mulmod(a[i * d + j], pinv, q);
            inv[i * d + j] = mfo_mulmod(inv[i * d + j], pinv, q);
        }
#pragma omp parallel for schedule(static)
        for (long r = 0; r < (long)d; ++r) {
            if ((size_t)r == i) continue;
            uint64_t fac = a[(size_t)r * d + i];
            if (!fac) continue;
            for (size_t c = 0; c < d; ++c) {
                a[(size_t)r * d + c] = mfo_submod(a[(size_t)r * d + c], mfo_mulmod(fac, a[i * d + c], q), q);
                inv[(size_t)r * d + c] = mfo_submod(inv[(size_t)r * d + c], mfo_mulmod(fac, inv[i * d + c], q), q);
            }
        }
    }
    free(a);
    return 0;
}

/* HE.cu:187-235 matrix_inverse_complex: Gauss-Jordan with partial pivoting.
 * std::complex<double> division is restated as the textbook formula (libstdc++ uses the
 * same through __divdc3 up to scaling for finite, well-scaled operands). */
static mfo_cx cx_mul(mfo_cx a, mfo_cx b) { mfo_cx r = {a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re}; return r; }
static mfo_cx cx_div(mfo_cx a, mfo_cx b) {
    double d = b.re * b.re + b.im * b.im;
    mfo_cx r = {(a.re * b.re + a.im * b.im) / d, (a.im * b.re - a.re * b.im) / d};
    return r;
}
static double cx_abs(mfo_cx a) { return hypot(a.re, a.im); }
static int matrix_inverse_complex(const mfo_cx *m, mfo_cx *inv, int dim) {
    size_t d = (size_t)dim;
    mfo_cx *a = (mfo_cx *)malloc(d * d * sizeof(mfo_cx));
    memcpy(a, m, d * d * sizeof(mfo_cx));
    memset(inv, 0, d * d * sizeof(mfo_cx));
    for (size_t i = 0; i < d; ++i) inv[i * d + i].re = 1.0;
    for (size_t i = 0; i < d; ++i) {
        size_t piv = i; double best = cx_abs(a[i * d + i]);
        for (size_t r = i + 1; r < d; ++r) { double c = cx_abs(a[r * d + i]); if (c > best) { best = c; piv = r; } }
        if (best < 1e-18) { free(a); return -1; }
        if (piv != i)
            for (size_t j = 0; j < d; ++j) {
                mfo_cx t = a[i * d + j]; a[i * d + j] = a[piv * d + j]; a[piv * d + j] = t;
                t = inv[i * d + j]; inv[i * d + j] = inv[piv * d + j]; inv[piv * d + j] = t;
            }
        mfo_cx pv = a[i * d + i];
        for (size_t j = 0; j < d; ++j) { a[i * d + j] = cx_div(a[i * d + j], pv); inv[i * d + j] = cx_div(inv[i * d + j], pv); }
#pragma omp parallel for schedule(static)
        for (long r = 0; r < (long)d; ++r) {
            if ((size_t)r == i) continue;
            mfo_cx fac = a[(size_t)r * d + i];
            if (cx_abs(fac) < 1e-18) continue;
            for (size_t c = 0; c < d; ++c) {
                mfo_cx t = cx_mul(fac, a[i * d + c]);
                a[(size_t)r * d + c].re -= t.re; a[(size_t)r * d + c].im -= t.im;
                t = cx_mul(fac, inv[i * d + c]);
                inv[(size_t)r * d + c].re -= t.re; inv[(size_t)r * d + c].im -= t.im;
            }
        }
    }
    free(a);
    return 0;
}

/* Phi_p(W) for p = 3 f as integer coefficients (degree phi = 2 (f-1)):
 * Phi_p = (W^p - 1)(W - 1) / ((W^3 - 1)(W^f - 1)). */
static void cyclotomic_3f(int f, int *phi_out /* [2(f-1)+1] */) {
    int p = 3 * f, len = p + 2;
    long *a = (long *)calloc((size_t)len + 1, sizeof(long));
    long *t = (long *)calloc((size_t)len + 1, sizeof(long));
    a[p] = 1; a[0] = -1;
    int deg = p;
    /* divide by W^f - 1 */
    memset(t, 0, ((size_t)len + 1) * sizeof(long));
    for (int d = deg; d >= f; --d) { long c = a[d]; if (!c) continue; t[d - f] += c; a[d] -= c; a[d - f] += c; }
    memcpy(a, t, ((size_t)len + 1) * sizeof(long)); deg -= f;
    /* multiply by W - 1 */
    for (int d = deg + 1; d >= 1; --d) a[d] = a[d - 1] - a[d];
    a[0] = -a[0]; deg += 1;
    /* divide by W^3 - 1 */
    memset(t, 0, ((size_t)len + 1) * sizeof(long));
    for (int d = deg; d >= 3; --d) { long c = a[d]; if (!c) continue; t[d - 3] += c; a[d] -= c; a[d - 3] += c; }
    deg -= 3;
    for (int i = 0; i <= deg; ++i) phi_out[i] = (int)t[i];
    free(a); free(t);
}

static void fill_wexp(mfo_ctx *c) {
    /* HE.cu:72-105 k_wntt_exp == (a*257 + b*3) mod 771, a in {1,2} outer, b in 1..256 inner
     * (batched_encoder.cu:276-282 builds it by this very rule) */
    int idx = 0;
    for (int a = 1; a <= 2; ++a)
        for (int b = 1; b <= c->f - 1; ++b) c->wexp[idx++] = (uint16_t)((a * c->f + b * 3) % c->p);
}

static void build_wntt(mfo_ctx *c, int gj) {
    const size_t phi = (size_t)c->phi;
    int *Phi = (int *)calloc(phi + 1, sizeof(int));
    cyclotomic_3f(c->f, Phi);
    for (int l = 0; l < c->L; ++l) {
        const uint64_t q = c->moduli[l];
        const uint64_t eta = mfo_find_eta(q, c->f);
        c->eta[l] = eta;
        uint64_t *V = c->wntt + (size_t)l * phi * phi;       /* [w][r], HE.cu:245-254 */
        for (size_t w = 0; w < phi; ++w) {
            uint64_t root = mfo_powmod(eta, c->wexp[w], q), cur = 1;
            for (size_t r = 0; r < phi; ++r) { V[w * phi + r] = cur; cur = mfo_mulmod(cur, root, q); }
        }
        uint64_t *VI = c->wntt_inv + (size_t)l * phi * phi;  /* [w][r] = V^-1[r][w] */
        if (gj) {
            uint64_t *inv = (uint64_t *)malloc(phi * phi * sizeof(uint64_t));
            mfo_matrix_inverse_mod(V, inv, (int)phi, q);
            for (size_t w = 0; w < phi; ++w)
                for (size_t r = 0; r < phi; ++r) VI[w * phi + r] = inv[r * phi + w];
            free(inv);
        } else {
            const uint64_t pinv = mfo_invmod((uint64_t)c->p % q, q), ieta = mfo_invmod(eta, q);
#pragma omp parallel for schedule(dynamic, 4)
            for (long w = 0; w < (long)phi; ++w) {
                uint64_t g[4096];
                uint64_t root = mfo_powmod(ieta, c->wexp[w], q), cur = pinv;
                for (int j = 0; j < c->p; ++j) { g[j] = cur; cur = mfo_mulmod(cur, root, q); }
                for (int d = c->p - 1; d >= (int)phi; --d) {
                    uint64_t co = g[d];
                    if (!co) continue;
                    for (size_t i = 0; i < phi; ++i) {      /* Phi monic: g -= co * W^(d-phi) * Phi */
                        int pc = Phi[i];
                        if (!pc) continue;
                        uint64_t v = pc == 1 || pc == -1 ? co : mfo_mulmod(co, (uint64_t)(pc < 0 ? -pc : pc), q);
                        size_t k = (size_t)d - phi + i;
                        g[k] = pc > 0 ? mfo_submod(g[k], v, q) : mfo_addmod(g[k], v, q);
                    }
                }
                for (size_t r = 0; r < phi; ++r) VI[(size_t)w * phi + r] = g[r];
            }
        }
    }
    free(Phi);
}

static void build_wdft(mfo_ctx *c, int gj) {
    const size_t phi = (size_t)c->phi;
    const double two_pi = 6.283185307179586476925286766559;
    /* HE.cu:282-292 */
    for (size_t w = 0; w < phi; ++w) {
        double ang = two_pi * (double)c->wexp[w] / (double)c->p;
        mfo_cx root = {cos(ang), sin(ang)}, cur = {1.0, 0.0};
        for (size_t r = 0; r < phi; ++r) { c->wdft[w * phi + r] = cur; cur = cx_mul(cur, root); }
    }
    if (gj) {
        mfo_cx *inv = (mfo_cx *)malloc(phi * phi * sizeof(mfo_cx));
        matrix_inverse_complex(c->wdft, inv, (int)phi);
        for (size_t w = 0; w < phi; ++w)
            for (size_t r = 0; r < phi; ++r) c->wdft_inv[w * phi + r] = inv[r * phi + w];
        free(inv);
    } else {
        int *Phi = (int *)calloc(phi + 1, sizeof(int));
        cyclotomic_3f(c->f, Phi);
#pragma omp parallel for schedule(dynamic, 4)
        for (long w = 0; w < (long)phi; ++w) {
            long double gr[4096], gi[4096];
            for (int j = 0; j < c->p; ++j) {
                long k = ((long)c->wexp[w] * j) % c->p;
                long double ang = -6.283185307179586476925286766559L * (long double)k / (long double)c->p;
                gr[j] = cosl(ang) / c->p; gi[j] = sinl(ang) / c->p;
            }
            for (int d = c->p - 1; d >= (int)phi; --d)
                for (size_t i = 0; i < phi; ++i) {
                    int pc = Phi[i];
                    if (!pc) continue;
                    size_t k = (size_t)d - phi + i;
                    gr[k] -= pc * gr[d]; gi[k] -= pc * gi[d];
                }
            for (size_t r = 0; r < phi; ++r) { c->wdft_inv[(size_t)w * phi + r].re = (double)gr[r]; c->wdft_inv[(size_t)w * phi + r].im = (double)gi[r]; }
        }
        free(Phi);
    }
}

/* encoder.cu:425-444 init_complex_matrices, literally (same operation order) */
static void build_xy(mfo_ctx *c) {
    const int n = c->n;
    const double PI = 3.141592653589793;
    for (int j = 0; j < n; ++j) {
        uint64_t e = 1, b5 = 5; int p = j;
        while (p > 0) { if (p & 1) e = (e * b5) % (uint64_t)(4 * n); b5 = (b5 * b5) % (uint64_t)(4 * n); p >>= 1; }
        double ang = 2.0 * PI * e / (4.0 * n);
        mfo_cx z = {cos(ang), sin(ang)}, zi = {z.re, -z.im}, cc = {1, 0}, ci = {1, 0};
        for (int k = 0; k < n; ++k) {
            c->xyV[j * n + k] = cc;
            mfo_cx s = {1.0 / n, 0};
            c->xyVinv[k * n + j] = cx_mul(ci, s);
            cc = cx_mul(cc, z); ci = cx_mul(ci, zi);
        }
    }
    for (int r = 0; r < n; ++r)
        for (int k = 0; k < n; ++k) { c->xyVT[k * n + r] = c->xyV[r * n + k]; c->xyVinvT[k * n + r] = c->xyVinv[r * n + k]; }
}

/* encoder.cu:341-421: Q = prod q_i, Q/2, M_i = Q/q_i, inv_i = (M_i mod q_i)^-1 */
static void build_crt(mfo_ctx *c) {
    int bits = 0;
    for (int l = 0; l < c->L; ++l) { int b = 0; while ((c->moduli[l] >> b) != 0) ++b; bits += b; }
    int W = (bits + 63) / 64;
    if (W < 1) W = 1;
    c->big_words = W;
    uint64_t Q[MFO_MAX_BIG + 1] = {0};
    Q[0] = 1;
    for (int l = 0; l < c->L; ++l) {
        u128 carry = 0;
        for (int i = 0; i < W; ++i) { u128 pr = (u128)Q[i] * c->moduli[l] + carry; Q[i] = (uint64_t)pr; carry = pr >> 64; }
    }
    memcpy(c->crt_Q, Q, (size_t)W * 8);
    uint64_t cy = 0;
    for (int i = W - 1; i >= 0; --i) { c->crt_Qhalf[i] = (Q[i] >> 1) | (cy << 63); cy = Q[i] & 1; }
    for (int l = 0; l < c->L; ++l) {
        uint64_t qi = c->moduli[l];
        u128 rem = 0;
        uint64_t *M = c->crt_M + (size_t)l * W;
        for (int i = W - 1; i >= 0; --i) { u128 cur = (rem << 64) | Q[i]; M[i] = (uint64_t)(cur / qi); rem = cur % qi; }
        rem = 0;
        for (int i = W - 1; i >= 0; --i) { u128 cur = (rem << 64) | M[i]; rem = cur % qi; }
        c->crt_inv[l] = mfo_invmod((uint64_t)rem, qi);
    }
}

mfo_ctx *mfo_ctx_create(int n, int f, int L, const uint64_t *moduli, double delta, int gj) {
    if (L < 1 || L > MFO_MAX_LIMBS || n < 2 || (n & (n - 1)) || 3 * f > 4095) return NULL;
    mfo_ctx *c = (mfo_ctx *)calloc(1, sizeof(mfo_ctx));
    c->n = n; c->f = f; c->p = 3 * f; c->phi = 2 * (f - 1); c->L = L; c->delta = delta;
    c->logn = 0; while ((1 << c->logn) < n) ++c->logn;
    memcpy(c->moduli, moduli, (size_t)L * 8);
    for (int l = 0; l < L; ++l)
        if ((moduli[l] - 1) % (uint64_t)c->p || (moduli[l] - 1) % (uint64_t)(2 * n)) { free(c); return NULL; }
    const size_t phi = (size_t)c->phi, nn = (size_t)n * n;
    c->wexp = (uint16_t *)calloc(phi, sizeof(uint16_t));
    c->wntt = (uint64_t *)malloc((size_t)L * phi * phi * 8);
    c->wntt_inv = (uint64_t *)malloc((size_t)L * phi * phi * 8);
    c->wdft = (mfo_cx *)malloc(phi * phi * sizeof(mfo_cx));
    c->wdft_inv = (mfo_cx *)malloc(phi * phi * sizeof(mfo_cx));
    c->xyV = (mfo_cx *)malloc(nn * sizeof(mfo_cx)); c->xyVT = (mfo_cx *)malloc(nn * sizeof(mfo_cx));
    c->xyVinv = (mfo_cx *)malloc(nn * sizeof(mfo_cx)); c->xyVinvT = (mfo_cx *)malloc(nn * sizeof(mfo_cx));
    c->tw = (uint64_t *)malloc((size_t)L * n * 8); c->tws = (uint64_t *)malloc((size_t)L * n * 8);
    c->itw = (uint64_t *)malloc((size_t)L * n * 8); c->itws = (uint64_t *)malloc((size_t)L * n * 8);
    fill_wexp(c);
    build_wntt(c, gj);
    build_wdft(c, gj);
    build_xy(c);
    build_crt(c);
    for (int l = 0; l < L; ++l)
        mfo_phantom_tables(c->logn, moduli[l], c->tw + (size_t)l * n, c->tws + (size_t)l * n,
                           c->itw + (size_t)l * n, c->itws + (size_t)l * n, &c->ninv[l], &c->ninvs[l]);
    return c;
}

void mfo_ctx_destroy(mfo_ctx *c) {
    if (!c) return;
    free(c->wexp); free(c->wntt); free(c->wntt_inv); free(c->wdft); free(c->wdft_inv);
    free(c->xyV); free(c->xyVT); free(c->xyVinv); free(c->xyVinvT);
    free(c->tw); free(c->tws); free(c->itw); free(c->itws);
    free(c);
}
const uint64_t *mfo_ctx_wntt(const mfo_ctx *c, int inverse) { return inverse ? c->wntt_inv : c->wntt; }
const uint16_t *mfo_ctx_wexp(const mfo_ctx *c) { return c->wexp; }
uint64_t mfo_ctx_eta(const mfo_ctx *c, int l) { return c->eta[l]; }
int mfo_ctx_big_words(const mfo_ctx *c) { return c->big_words; }
const uint64_t *mfo_ctx_crt(const mfo_ctx *c, int which) {
    return which == 0 ? c->crt_Q : which == 1 ? c->crt_Qhalf : which == 2 ? c->crt_M : c->crt_inv;
}
const mfo_cx *mfo_ctx_cx(const mfo_ctx *c, int which) {
    switch (which) { case 0: return c->wdft; case 1: return c->wdft_inv; case 2: return c->xyV;
                     case 3: return c->xyVT; case 4: return c->xyVinv; default: return c->xyVinvT; }
}

/* ---- W-CRT transforms ------------------------------------------------------------------ */
/* lazy 128-bit accumulation gives the same canonical residue as the reference's
 * add_mod(acc, mul_mod(a, w)) chain (HE.cu:737-743); phi * q^2 < 2^128 for q < 2^59. */
static uint64_t dot_mod(const uint64_t *a, size_t astride, const uint64_t *w, size_t wstride, size_t cnt, uint64_t q) {
    u128 acc = 0;
    for (size_t r = 0; r < cnt; ++r) {
        acc += (u128)a[r * astride] * w[r * wstride];
        if ((r & 63) == 63) acc %= q;
    }
    return (uint64_t)(acc % q);
}

void mfo_wntt_forward_matrix(const mfo_ctx *c, const uint64_t *in, uint64_t *out) {
    const size_t n = (size_t)c->n, n2 = n * n, L = (size_t)c->L, phi = (size_t)c->phi;
#pragma omp parallel for collapse(2) schedule(static)
    for (long l = 0; l < (long)L; ++l)
        for (long w = 0; w < (long)phi; ++w) {
            const uint64_t q = c->moduli[l];
            const uint64_t *pw = c->wntt + ((size_t)l * phi + (size_t)w) * phi;
            for (size_t y = 0; y < n; ++y)
                for (size_t x = 0; x < n; ++x)
                    out[(((size_t)w * n + y) * L + (size_t)l) * n + x] =
                        dot_mod(in + (size_t)l * n2 + y * n + x, L * n2, pw, 1, phi, q);
        }
}

void mfo_wntt_inverse_matrix(const mfo_ctx *c, const uint64_t *in, uint64_t *out) {
    const size_t n = (size_t)c->n, n2 = n * n, L = (size_t)c->L, phi = (size_t)c->phi;
#pragma omp parallel for collapse(2) schedule(static)
    for (long l = 0; l < (long)L; ++l)
        for (long r = 0; r < (long)phi; ++r) {
            const uint64_t q = c->moduli[l];
            const uint64_t *pw = c->wntt_inv + (size_t)l * phi * phi + (size_t)r;   /* [w][r], stride phi */
            for (size_t y = 0; y < n; ++y)
                for (size_t x = 0; x < n; ++x)
                    out[((size_t)r * L + (size_t)l) * n2 + y * n + x] =
                        dot_mod(in + (y * L + (size_t)l) * n + x, n * L * n, pw, phi, phi, q);
        }
}

void mfo_wntt_forward_vector(const mfo_ctx *c, const uint64_t *in, uint64_t *out) {
    const size_t n = (size_t)c->n, L = (size_t)c->L, phi = (size_t)c->phi;
#pragma omp parallel for collapse(2) schedule(static)
    for (long w = 0; w < (long)phi; ++w)
        for (long l = 0; l < (long)L; ++l) {
            const uint64_t *pw = c->wntt + ((size_t)l * phi + (size_t)w) * phi;
            for (size_t x = 0; x < n; ++x)
                out[((size_t)w * L + (size_t)l) * n + x] = dot_mod(in + (size_t)l * n + x, L * n, pw, 1, phi, c->moduli[l]);
        }
}

void mfo_matrix_to_poly(const mfo_ctx *c, const uint64_t *in, uint64_t *out) {
    const size_t n = (size_t)c->n, L = (size_t)c->L, phi = (size_t)c->phi;
    for (size_t w = 0; w < phi; ++w) for (size_t l = 0; l < L; ++l) for (size_t y = 0; y < n; ++y)
        memcpy(out + ((w * n + y) * L + l) * n, in + ((w * L + l) * n + y) * n, n * 8);
}
void mfo_poly_to_matrix(const mfo_ctx *c, const uint64_t *in, uint64_t *out) {
    const size_t n = (size_t)c->n, L = (size_t)c->L, phi = (size_t)c->phi;
    for (size_t w = 0; w < phi; ++w) for (size_t l = 0; l < L; ++l) for (size_t y = 0; y < n; ++y)
        memcpy(out + ((w * L + l) * n + y) * n, in + ((w * n + y) * L + l) * n, n * 8);
}

/* big-int helpers shared with scheme_oracle.c (encoder.cu:51-110, HE.cu:859-915) */
int mfo_big_cmp(const uint64_t *a, const uint64_t *b, int W) {
    for (int i = W - 1; i >= 0; --i) { if (a[i] > b[i]) return 1; if (a[i] < b[i]) return -1; }
    return 0;
}
void mfo_big_compose(const mfo_ctx *c, const uint64_t *res /* [limbs] */, int limbs, uint64_t *mag, int *neg) {
    const int W = c->big_words;
    uint64_t acc[MFO_MAX_BIG] = {0};
    for (int k = 0; k < limbs; ++k) {
        uint64_t t = mfo_mulmod(res[k], c->crt_inv[k], c->moduli[k]);
        const uint64_t *M = c->crt_M + (size_t)k * W;
        u128 carry = 0; uint64_t term[MFO_MAX_BIG];
        for (int i = 0; i < W; ++i) { u128 p = (u128)M[i] * t + carry; term[i] = (uint64_t)p; carry = p >> 64; }
        carry = 0;
        for (int i = 0; i < W; ++i) { u128 s = (u128)acc[i] + term[i] + carry; acc[i] = (uint64_t)s; carry = s >> 64; }
        if (mfo_big_cmp(acc, c->crt_Q, W) >= 0) {
            uint64_t br = 0;
            for (int i = 0; i < W; ++i) { uint64_t bi = c->crt_Q[i] + br; br = acc[i] < bi; acc[i] -= bi; }
        }
    }
    *neg = 0;
    if (mfo_big_cmp(acc, c->crt_Qhalf, W) > 0) {
        uint64_t br = 0;
        for (int i = 0; i < W; ++i) { uint64_t ai = acc[i] + br; br = c->crt_Q[i] < ai; mag[i] = c->crt_Q[i] - ai; }
        *neg = 1;
    } else memcpy(mag, acc, (size_t)W * 8);
}

/* HE.cu:1029-1081 */
void mfo_wntt_forward_centered(const mfo_ctx *c, const int64_t *in, int64_t *out) {
    const size_t n2 = (size_t)c->n * c->n, phi = (size_t)c->phi;
    const int W = c->big_words;
#pragma omp parallel for schedule(static)
    for (long idx = 0; idx < (long)(phi * n2); ++idx) {
        size_t pos = (size_t)idx % n2, w = (size_t)idx / n2;
        uint64_t res[MFO_MAX_LIMBS];
        for (int l = 0; l < c->L; ++l) {
            const uint64_t q = c->moduli[l];
            const uint64_t *pw = c->wntt + ((size_t)l * phi + w) * phi;
            u128 acc = 0;
            for (size_t r = 0; r < phi; ++r) {
                int64_t m = in[r * n2 + pos] % (int64_t)q;
                if (m < 0) m += (int64_t)q;
                acc += (u128)(uint64_t)m * pw[r];
                if ((r & 63) == 63) acc %= q;
            }
            res[l] = (uint64_t)(acc % q);
        }
        uint64_t mag[MFO_MAX_BIG]; int neg;
        mfo_big_compose(c, res, c->L, mag, &neg);
        int over = mag[0] > (uint64_t)INT64_MAX;             /* HE.cu:902-915 he_big_to_i64_checked */
        for (int i = 1; i < W; ++i) if (mag[i]) over = 1;
        out[idx] = over ? (neg ? INT64_MIN : INT64_MAX) : (neg ? -(int64_t)mag[0] : (int64_t)mag[0]);
    }
}

/* HE.cu:1083-1114: limb 0 only */
void mfo_wntt_inverse_centered(const mfo_ctx *c, const int64_t *in, int64_t *out) {
    const size_t n2 = (size_t)c->n * c->n, phi = (size_t)c->phi;
    const uint64_t q = c->moduli[0];
#pragma omp parallel for schedule(static)
    for (long idx = 0; idx < (long)(phi * n2); ++idx) {
        size_t pos = (size_t)idx % n2, r = (size_t)idx / n2;
        u128 acc = 0;
        for (size_t w = 0; w < phi; ++w) {
            int64_t m = in[w * n2 + pos] % (int64_t)q;
            if (m < 0) m += (int64_t)q;
            acc += (u128)(uint64_t)m * c->wntt_inv[w * phi + r];
            if ((w & 63) == 63) acc %= q;
        }
        uint64_t a = (uint64_t)(acc % q);
        out[idx] = a > (q >> 1) ? (int64_t)a - (int64_t)q : (int64_t)a;
    }
}

/* ---- complex W-DFT ----------------------------------------------------------------------- */
/* HE.cu:1147-1172 (sequential r order, separate mul/add as nvcc would contract to FMA: the
 * oracle keeps plain mul+add; parity with the GPU is tolerance-based) */
void mfo_wdft_forward_complex(const mfo_ctx *c, const mfo_cx *in, mfo_cx *out) {
    const size_t n2 = (size_t)c->n * c->n, phi = (size_t)c->phi;
#pragma omp parallel for schedule(static)
    for (long idx = 0; idx < (long)(phi * n2); ++idx) {
        size_t pos = (size_t)idx % n2, w = (size_t)idx / n2;
        double ar = 0, ai = 0;
        for (size_t r = 0; r < phi; ++r) {
            mfo_cx a = in[r * n2 + pos], v = c->wdft[w * phi + r];
            ar += a.re * v.re - a.im * v.im; ai += a.re * v.im + a.im * v.re;
        }
        out[idx].re = ar; out[idx].im = ai;
    }
}
/* batched_encoder.cu:104-123 */
void mfo_w_idft(const mfo_ctx *c, const mfo_cx *in, mfo_cx *out) {
    const size_t n2 = (size_t)c->n * c->n, phi = (size_t)c->phi;
#pragma omp parallel for schedule(static)
    for (long idx = 0; idx < (long)(phi * n2); ++idx) {
        size_t pos = (size_t)idx % n2, r = (size_t)idx / n2;
        double ar = 0, ai = 0;
        for (size_t w = 0; w < phi; ++w) {
            mfo_cx a = in[w * n2 + pos], v = c->wdft_inv[w * phi + r];
            ar += a.re * v.re - a.im * v.im; ai += a.re * v.im + a.im * v.re;
        }
        out[idx].re = ar; out[idx].im = ai;
    }
}
/* HE.cu:1116-1145 */
void mfo_wdft_forward_pair(const mfo_ctx *c, const int64_t *re, const int64_t *im, double *ore, double *oim) {
    const size_t n2 = (size_t)c->n * c->n, phi = (size_t)c->phi;
#pragma omp parallel for schedule(static)
    for (long idx = 0; idx < (long)(phi * n2); ++idx) {
        size_t pos = (size_t)idx % n2, w = (size_t)idx / n2;
        double ar = 0, ai = 0;
        for (size_t r = 0; r < phi; ++r) {
            double a = (double)re[r * n2 + pos], b = (double)im[r * n2 + pos]; mfo_cx v = c->wdft[w * phi + r];
            ar += a * v.re - b * v.im; ai += a * v.im + b * v.re;
        }
        ore[idx] = ar; oim[idx] = ai;
    }
}
/* HE.cu:1174-1202 */
void mfo_wdft_inverse_pair(const mfo_ctx *c, const double *re, const double *im, double *ore, double *oim) {
    const size_t n2 = (size_t)c->n * c->n, phi = (size_t)c->phi;
#pragma omp parallel for schedule(static)
    for (long idx = 0; idx < (long)(phi * n2); ++idx) {
        size_t pos = (size_t)idx % n2, r = (size_t)idx / n2;
        double ar = 0, ai = 0;
        for (size_t w = 0; w < phi; ++w) {
            double a = re[w * n2 + pos], b = im[w * n2 + pos]; mfo_cx v = c->wdft_inv[w * phi + r];
            ar += a * v.re - b * v.im; ai += a * v.im + b * v.re;
        }
        ore[idx] = ar; oim[idx] = ai;
    }
}
