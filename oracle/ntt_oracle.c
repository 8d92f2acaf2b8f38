/*
 * ntt_oracle.c -- CPU restatement of the NTT part of the hot path.
 * TEST INFRASTRUCTURE ONLY (see mfhe_oracle.h).  Plain C, gcc.
 */
#include "mfhe_oracle.h"
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef unsigned __int128 u128;

/* include/core/common.cuh:12-39 ------------------------------------------- */
uint64_t mfo_addmod(uint64_t a, uint64_t b, uint64_t q) { uint64_t s = a + b; return s >= q ? s - q : s; }
uint64_t mfo_submod(uint64_t a, uint64_t b, uint64_t q) { return a >= b ? a - b : a + q - b; }
uint64_t mfo_mulmod(uint64_t a, uint64_t b, uint64_t q) { return (uint64_t)((u128)a * b % q); }
uint64_t mfo_powmod(uint64_t b, uint64_t e, uint64_t q) {
    uint64_t r = 1; b %= q;
    while (e) { if (e & 1) r = mfo_mulmod(r, b, q); b = mfo_mulmod(b, b, q); e >>= 1; }
    return r;
}
uint64_t mfo_invmod(uint64_t a, uint64_t q) { return mfo_powmod(a, q - 2, q); }

/* include/backend/phantom_math.cuh:9-28 ------------------------------------ */
uint64_t mfo_shoup(uint64_t w, uint64_t q) { return (uint64_t)(((u128)w << 64) / q); }
static inline uint64_t shoup_lazy(uint64_t x, uint64_t w, uint64_t ws, uint64_t q) {
    uint64_t hi = (uint64_t)(((u128)x * ws) >> 64);
    return x * w - hi * q;                       /* [0, 2q) */
}
uint64_t mfo_mulmod_shoup(uint64_t x, uint64_t w, uint64_t ws, uint64_t q) {
    uint64_t r = shoup_lazy(x, w, ws, q);
    return r >= q ? r - q : r;
}

uint32_t mfo_bitrev(uint32_t x, int bits) {
    uint32_t r = 0;
    for (int i = 0; i < bits; ++i) { r = (r << 1) | (x & 1u); x >>= 1; }
    return r;
}

/* deterministic Miller-Rabin for 64-bit n (bases 2..37).  Phantom's
 * Modulus::is_prime (ph/src/host/modulus.cu) is the probabilistic version. */
int mfo_is_prime(uint64_t n) {
    static const uint64_t bases[12] = {2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37};
    if (n < 2) return 0;
    for (int i = 0; i < 12; ++i) { if (n == bases[i]) return 1; if (n % bases[i] == 0) return 0; }
    uint64_t d = n - 1; int r = 0;
    while (!(d & 1)) { d >>= 1; ++r; }
    for (int i = 0; i < 12; ++i) {
        uint64_t x = mfo_powmod(bases[i], d, n);
        if (x == 1 || x == n - 1) continue;
        int comp = 1;
        for (int j = 1; j < r; ++j) { x = mfo_mulmod(x, x, n); if (x == n - 1) { comp = 0; break; } }
        if (comp) return 0;
    }
    return 1;
}

/* ph/src/host/numth.cu:207-230 -------------------------------------------- */
int mfo_get_primes(uint64_t ntt_size, int bit_size, size_t count, uint64_t *out) {
    uint64_t factor = 2 * ntt_size;
    uint64_t value = ((uint64_t)1 << bit_size) - factor + 1;
    uint64_t lower = (uint64_t)1 << (bit_size - 1);
    size_t found = 0;
    while (found < count && value > lower) {
        if (mfo_is_prime(value)) out[found++] = value;
        value -= factor;
    }
    return found == count ? 0 : -1;
}

/* ph/src/host/modulus.cu:80-112, single bit size --------------------------- */
int mfo_coeff_modulus_create(uint64_t n, int bit_size, size_t count, uint64_t *out) {
    uint64_t *tmp = (uint64_t *)malloc(count * sizeof(uint64_t));
    if (!tmp) return -2;
    int rc = mfo_get_primes(n, bit_size, count, tmp);
    if (rc == 0) for (size_t i = 0; i < count; ++i) out[i] = tmp[count - 1 - i];
    free(tmp);
    return rc;
}

/* ph/src/host/numth.cu:265-332.  The reference draws random candidates and
 * then minimises over all odd powers; the minimum does not depend on the
 * starting root, so a deterministic search gives the same value.            */
uint64_t mfo_min_primitive_root(uint64_t degree, uint64_t q) {
    if ((q - 1) % degree) return 0;
    uint64_t quot = (q - 1) / degree, root = 0;
    for (uint64_t x = 2; x < 100000; ++x) {
        uint64_t r = mfo_powmod(x, quot, q);
        if (mfo_powmod(r, degree >> 1, q) == q - 1) { root = r; break; }
    }
    if (!root) return 0;
    uint64_t sq = mfo_mulmod(root, root, q), cur = root, best = root;
    for (uint64_t i = 0; i < degree; ++i) {
        if (cur < best) best = cur;
        cur = mfo_mulmod(cur, sq, q);
    }
    return best;
}

/* ph/src/host/ntt.cu:11-56 -------------------------------------------------- */
int mfo_phantom_tables(int logn, uint64_t q, uint64_t *tw, uint64_t *tws,
                       uint64_t *itw, uint64_t *itws, uint64_t *n_inv, uint64_t *n_inv_shoup) {
    size_t n = (size_t)1 << logn;
    uint64_t psi = mfo_min_primitive_root(2 * n, q);
    if (!psi) return -1;
    uint64_t ipsi = mfo_invmod(psi, q);
    uint64_t p = psi, ip = ipsi;
    for (size_t i = 1; i < n; ++i) {
        uint32_t r = mfo_bitrev((uint32_t)i, logn);
        tw[r] = p;   tws[r] = mfo_shoup(p, q);
        itw[r] = ip; itws[r] = mfo_shoup(ip, q);
        p = mfo_mulmod(p, psi, q); ip = mfo_mulmod(ip, ipsi, q);
    }
    tw[0] = 1;  tws[0] = mfo_shoup(1, q);
    itw[0] = 1; itws[0] = mfo_shoup(1, q);
    uint64_t ni = mfo_invmod((uint64_t)n % q, q);
    *n_inv = ni; *n_inv_shoup = mfo_shoup(ni, q);
    itw[1] = mfo_mulmod(itw[1], ni, q);
    itws[1] = mfo_shoup(itw[1], q);
    return 0;
}

/* ph/src/ntt/ntt_1d.cu:17-76 with ph/include/butterfly.cuh:10-22 ------------ */
void mfo_negacyclic_forward(uint64_t *a, int logn, uint64_t q,
                            const uint64_t *tw, const uint64_t *tws) {
    size_t n = (size_t)1 << logn;
    uint64_t q2 = 2 * q;
    for (size_t m = 1; m < n; m <<= 1) {
        size_t pairs = n / m / 2;
        for (size_t k = 0; k < m; ++k) {
            uint64_t w = tw[m + k], ws = tws[m + k];
            uint64_t *x = a + 2 * k * pairs, *y = x + pairs;
            for (size_t j = 0; j < pairs; ++j) {
                uint64_t t = shoup_lazy(y[j], w, ws, q);
                uint64_t u = x[j]; u = u >= q2 ? u - q2 : u;
                y[j] = u + q2 - t;
                x[j] = u + t;
            }
        }
    }
    for (size_t i = 0; i < n; ++i) {
        uint64_t v = a[i];
        v = v >= q2 ? v - q2 : v;
        a[i] = v >= q ? v - q : v;
    }
}

/* ph/src/ntt/ntt_1d.cu:194-256 with ph/include/butterfly.cuh:28-37 ---------- */
void mfo_negacyclic_inverse(uint64_t *a, int logn, uint64_t q,
                            const uint64_t *itw, const uint64_t *itws,
                            uint64_t n_inv, uint64_t n_inv_shoup) {
    size_t n = (size_t)1 << logn;
    uint64_t q2 = 2 * q;
    for (size_t m = n / 2; m >= 1; m >>= 1) {
        size_t pairs = n / m / 2;
        for (size_t k = 0; k < m; ++k) {
            uint64_t w = itw[m + k], ws = itws[m + k];
            uint64_t *x = a + 2 * k * pairs, *y = x + pairs;
            for (size_t j = 0; j < pairs; ++j) {
                uint64_t t = x[j] + q2 - y[j];
                uint64_t s = x[j] + y[j];
                s = s >= q2 ? s - q2 : s;
                x[j] = s;
                y[j] = shoup_lazy(t, w, ws, q);
            }
        }
    }
    for (size_t i = 0; i < n; ++i) {
        uint64_t v = a[i];
        v = v >= q ? v - q : v;
        if (i < n / 2) v = mfo_mulmod_shoup(v, n_inv, n_inv_shoup, q);
        a[i] = v;
    }
}

void mfo_negacyclic_forward_def(const uint64_t *a, uint64_t *out, int logn,
                                uint64_t q, uint64_t psi) {
    size_t n = (size_t)1 << logn;
    for (size_t k = 0; k < n; ++k) {
        uint64_t root = mfo_powmod(psi, 2 * k + 1, q), p = 1, acc = 0;
        for (size_t j = 0; j < n; ++j) {
            acc = mfo_addmod(acc, mfo_mulmod(a[j] % q, p, q), q);
            p = mfo_mulmod(p, root, q);
        }
        out[mfo_bitrev((uint32_t)k, logn)] = acc;
    }
}

/* test/test_custom_ntt_roundtrip.cu:263-280 --------------------------------- */
void mfo_schoolbook_mul(const uint64_t *a, const uint64_t *b, uint64_t *c,
                        size_t n, uint64_t q, uint64_t wrap) {
    memset(c, 0, n * sizeof(uint64_t));
    for (size_t j = 0; j < n; ++j)
        for (size_t k = 0; k < n; ++k) {
            uint64_t prod = mfo_mulmod(a[j], b[k], q);
            size_t idx = j + k;
            if (idx < n) c[idx] = mfo_addmod(c[idx], prod, q);
            else c[idx - n] = mfo_addmod(c[idx - n], mfo_mulmod(prod, wrap, q), q);
        }
}

void mfo_negacyclic_forward_batch(uint64_t *data, size_t batch, size_t L, int logn,
                                  const uint64_t *moduli, const uint64_t *tw, const uint64_t *tws) {
    size_t n = (size_t)1 << logn;
    long total = (long)(batch * L);
#pragma omp parallel for schedule(dynamic, 1)
    for (long i = 0; i < total; ++i) {
        size_t l = (size_t)i % L;
        mfo_negacyclic_forward(data + (size_t)i * n, logn, moduli[l], tw + l * n, tws + l * n);
    }
}

void mfo_negacyclic_inverse_batch(uint64_t *data, size_t batch, size_t L, int logn,
                                  const uint64_t *moduli, const uint64_t *itw, const uint64_t *itws,
                                  const uint64_t *n_inv, const uint64_t *n_inv_shoup) {
    size_t n = (size_t)1 << logn;
    long total = (long)(batch * L);
#pragma omp parallel for schedule(dynamic, 1)
    for (long i = 0; i < total; ++i) {
        size_t l = (size_t)i % L;
        mfo_negacyclic_inverse(data + (size_t)i * n, logn, moduli[l], itw + l * n, itws + l * n,
                               n_inv[l], n_inv_shoup[l]);
    }
}

int mfo_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* bench.py: torch.distributed.run exports OMP_NUM_THREADS=1; the CPU arm states and sets its thread count itself */
void mfo_set_num_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

/* ---- reference gl convention ------------------------------------------------ */
/* src/core/ntt_core.cu:49-70 */
uint64_t mfo_gl_get_psi(uint64_t q, int n) {
    uint64_t order = 4 * (uint64_t)n;
    if ((q - 1) % order) return 0;
    for (uint64_t root = 2; root <= 100000; ++root) {
        uint64_t g = mfo_powmod(root, (q - 1) / order, q);
        if (mfo_powmod(g, 2 * (uint64_t)n, q) == q - 1) return g;
    }
    return 0;
}

static int ilog2(int n) { int l = 0; while ((1 << l) < n) ++l; return l; }

/* src/core/ntt_core.cu:215-240 (bit reverse) + :271-303 (CT stage) */
static void ct_natural(uint64_t *a, int n, uint64_t q, uint64_t omega) {
    int logn = ilog2(n);
    for (int i = 0; i < n; ++i) {
        int r = (int)mfo_bitrev((uint32_t)i, logn);
        if (r > i) { uint64_t t = a[i]; a[i] = a[r]; a[r] = t; }
    }
    uint64_t *pw = (uint64_t *)malloc((size_t)n * sizeof(uint64_t));
    pw[0] = 1;
    for (int i = 1; i < n; ++i) pw[i] = mfo_mulmod(pw[i - 1], omega, q);
    for (int m = 2; m <= n; m <<= 1) {
        int step = n / m;
        for (int k = 0; k < n / m; ++k)
            for (int j = 0; j < m / 2; ++j) {
                int i = k * m + j, it = i + m / 2;
                uint64_t vw = mfo_mulmod(a[it], pw[j * step], q), u = a[i];
                a[i] = mfo_addmod(u, vw, q);
                a[it] = mfo_submod(u, vw, q);
            }
    }
    free(pw);
}

void mfo_cyclic_forward(uint64_t *a, int n, uint64_t q) {
    uint64_t psi = mfo_gl_get_psi(q, n);
    ct_natural(a, n, q, mfo_powmod(psi, 4, q));
}
void mfo_cyclic_backward(uint64_t *a, int n, uint64_t q) {
    uint64_t psi = mfo_gl_get_psi(q, n);
    ct_natural(a, n, q, mfo_invmod(mfo_powmod(psi, 4, q), q));
    uint64_t ni = mfo_invmod((uint64_t)n, q);
    for (int i = 0; i < n; ++i) a[i] = mfo_mulmod(a[i], ni, q);
}
/* src/core/ntt_core.cu:462-481 */
void mfo_gl_forward(uint64_t *a, int n, uint64_t q) {
    uint64_t psi = mfo_gl_get_psi(q, n), p = 1;
    for (int i = 0; i < n; ++i) { a[i] = mfo_mulmod(a[i], p, q); p = mfo_mulmod(p, psi, q); }
    mfo_cyclic_forward(a, n, q);
}
void mfo_gl_backward(uint64_t *a, int n, uint64_t q) {
    mfo_cyclic_backward(a, n, q);
    uint64_t ipsi = mfo_invmod(mfo_gl_get_psi(q, n), q), p = 1;
    for (int i = 0; i < n; ++i) { a[i] = mfo_mulmod(a[i], p, q); p = mfo_mulmod(p, ipsi, q); }
}
/* src/core/ntt_core.cu:150-173 */
void mfo_gl_perm(int n, uint32_t *perm, uint32_t *inv_perm) {
    int logn = ilog2(n);
    uint32_t m = 4u * (uint32_t)n, e = 1 % m;
    for (int j = 0; j < n; ++j) {
        uint32_t idx = (e - 1) / 4, t = mfo_bitrev(idx, logn);
        perm[j] = t; inv_perm[t] = (uint32_t)j;
        e = (uint32_t)((uint64_t)e * 5u % m);
    }
}

/* ---- FP64-pipe arithmetic of matrix_fhe_gpu_b200/csrc/modmath.cuh (q < 2^50), restated for the CPU ---------------
 * IEEE-754 fma / add / mul are bit-identical on the host and on the GPU, so the exactness claims and the compile-time
 * bound schedule of the kernels (fp_fwd_bound_before / fp_inv_bound_before) are checked here, without a GPU: the
 * emulation below runs the same expressions in the same order of stages, records the largest |v| / q that enters
 * every stage and fails when one exceeds the schedule.  Algorithm being replaced: the butterflies of
 * ph/include/butterfly.cuh:10-37 (ct_butterfly / gs_butterfly) inside ph/src/ntt/ntt_1d.cu:17-76, 194-256. */
#include <math.h>
static const double kRint = 6755399441055744.0;           /* 1.5 * 2^52 */
static double fp_mul(double y, double w, double wq, double q) {
    double h = y * w, l = fma(y, w, -h);
    double k = fma(y, wq, kRint) - kRint;
    return fma(-k, q, h) + l;
}
static double fp_fold(double x, double q, double qinv) {
    double k = fma(x, qinv, kRint) - kRint;
    return fma(-k, q, x);
}
static int fp_t(int B) { return 8 + (B + 15) / 16; }
static double centred(uint64_t w, uint64_t q) { return w > (q >> 1) ? (double)((int64_t)w - (int64_t)q) : (double)w; }
static uint64_t fp_canon(double r, uint64_t q) { int64_t i = (int64_t)r; return (uint64_t)(i < 0 ? i + (int64_t)q : i); }

/* products: count random (y, w) pairs with |y| <= bound16/16 * q plus the corner cases; returns the number of results
 * that are not the exact integer y*w - k*q, *max_ratio = largest |t| / q seen */
long mfo_fp64_modmul_check(uint64_t q, long count, uint64_t seed, int bound16, double *max_ratio) {
    const double qd = (double)q;
    const int64_t lim = (int64_t)((double)bound16 / 16.0 * qd);
    long bad = 0; double mx = 0;
    uint64_t s = seed * 0x9E3779B97F4A7C15ull + 1;
    for (long it = 0; it < count; ++it) {
        s ^= s << 13; s ^= s >> 7; s ^= s << 17; uint64_t r1 = s;
        s ^= s << 13; s ^= s >> 7; s ^= s << 17; uint64_t r2 = s;
        int64_t w = (int64_t)(r1 % q); if (w > (int64_t)(q >> 1)) w -= (int64_t)q;
        int64_t y = (int64_t)(r2 % (2 * (uint64_t)lim + 1)) - lim;
        switch (it & 15) {
            case 0: y = lim; break; case 1: y = -lim; break; case 2: w = (int64_t)(q >> 1); break;
            case 3: w = -(int64_t)(q >> 1); break; case 4: y = lim; w = (int64_t)(q >> 1); break;
            case 5: y = -lim; w = (int64_t)(q >> 1); break; case 6: y = 0; break; case 7: w = 0; break; default: break;
        }
        const double t = fp_mul((double)y, (double)w, (double)w / qd, qd);
        if (t != (double)(int64_t)t) { ++bad; continue; }
        __int128 d = (__int128)y * w - (__int128)(int64_t)t;
        if (d % (__int128)q != 0) ++bad;
        if (fabs(t) / qd > mx) mx = fabs(t) / qd;
    }
    if (max_ratio) *max_ratio = mx;
    return bad;
}

/* whole transforms; tw / itw are Phantom-order integer tables (mfo_phantom_tables: itw[1] carries n^-1), inputs below
 * 2q.  stage_max[s] = largest |v| / q entering stage s.  Returns 0, or -(s+1) when stage s sees a value above the
 * schedule. */
int mfo_fp64_negacyclic_forward(uint64_t *a, int logn, uint64_t q, const uint64_t *tw, double *stage_max) {
    const size_t n = (size_t)1 << logn;
    const double qd = (double)q, qinv = 1.0 / qd;
    double *v = (double *)malloc(n * sizeof(double));
    for (size_t i = 0; i < n; ++i) v[i] = (double)a[i];
    int B = 32, rc = 0, s = 0;
    for (size_t m = 1; m < n; m <<= 1, ++s) {
        const size_t pairs = n / m / 2;
        const int fold = (B + fp_t(B) > 64) && (s != logn - 1);
        double mx = 0;
        for (size_t i = 0; i < n; ++i) if (fabs(v[i]) > mx) mx = fabs(v[i]);
        if (stage_max) stage_max[s] = mx / qd;
        if (mx > (double)B / 16.0 * qd && !rc) rc = -(s + 1);
        for (size_t k = 0; k < m; ++k) {
            const double w = centred(tw[m + k], q), wq = w / qd;
            double *x = v + 2 * k * pairs, *y = x + pairs;
            for (size_t j = 0; j < pairs; ++j) {
                double X = x[j];
                if (fold) X = fp_fold(X, qd, qinv);
                const double t = fp_mul(y[j], w, wq, qd);
                y[j] = X - t; x[j] = X + t;
            }
        }
        B = (fold ? 9 : B) + fp_t(B);
    }
    for (size_t i = 0; i < n; ++i) a[i] = fp_canon(fp_fold(v[i], qd, qinv), q);
    free(v);
    return rc;
}
int mfo_fp64_negacyclic_inverse(uint64_t *a, int logn, uint64_t q, const uint64_t *itw, uint64_t n_inv, double *stage_max) {
    const size_t n = (size_t)1 << logn;
    const double qd = (double)q, qinv = 1.0 / qd;
    double *v = (double *)malloc(n * sizeof(double));
    for (size_t i = 0; i < n; ++i) v[i] = (double)a[i];
    int B = 32, rc = 0, s = 0;
    for (size_t m = n / 2; m >= 1; m >>= 1, ++s) {
        const size_t pairs = n / m / 2;
        const int fold = 2 * B > 32;
        double mx = 0;
        for (size_t i = 0; i < n; ++i) if (fabs(v[i]) > mx) mx = fabs(v[i]);
        if (stage_max) stage_max[s] = mx / qd;
        if (mx > (double)B / 16.0 * qd && !rc) rc = -(s + 1);
        for (size_t k = 0; k < m; ++k) {
            const double w = centred(itw[m + k], q), wq = w / qd;
            double *x = v + 2 * k * pairs, *y = x + pairs;
            for (size_t j = 0; j < pairs; ++j) {
                const double X = x[j], Y = y[j];
                if (m == 1) {                               /* last stage: n^-1 on the sums, itw[1] = w * n^-1 on the differences */
                    const double ni = centred(n_inv, q);
                    x[j] = fp_mul(X + Y, ni, ni / qd, qd);
                    y[j] = fp_mul(X - Y, w, wq, qd);
                } else {
                    double sum = X + Y;
                    if (fold) sum = fp_fold(sum, qd, qinv);
                    x[j] = sum;
                    y[j] = fp_mul(X - Y, w, wq, qd);
                }
            }
        }
        const int xb = fold ? 9 : 2 * B, yb = fp_t(2 * B);
        B = xb > yb ? xb : yb;
    }
    for (size_t i = 0; i < n; ++i) a[i] = fp_canon(v[i], q);
    free(v);
    return rc;
}
