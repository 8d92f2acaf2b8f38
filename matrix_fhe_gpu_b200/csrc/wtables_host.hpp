// wtables_host.hpp -- host-side construction of the matrix pipeline's constant tables.
//
// Product code (not the oracle).  Builds tables whose VALUES the reference defines:
//   src/core/HE.cu:72-133, 237-273     k_wntt_exp, h_find_eta, V_l[w][r] = eta^(exp[w] r), V_l^-1
//   src/core/HE.cu:275-310             complex V[w][r] = e^(2 pi i exp[w] r / p) and its inverse
//   src/core/encoder.cu:341-444        CRT constants (Q, Q/2, M_k, inv_k), XY Vandermonde V, V^-1
// but not the way the reference computes them: the inverses come from the closed form
//   V^-1[:, w] = ( p^-1 * sum_j eta^(-exp[w] j) W^j )  mod Phi_p(W)
// (inverse DFT of the indicator of unit exp[w], reduced modulo the cyclotomic polynomial; Phi_{3f}
// has coefficients in {-1,0,1}) instead of O(phi^3) Gauss-Jordan with 128-bit divisions -- the
// matrix is unique, so the entries are identical mod q, and in fp64 the closed form evaluated in
// long double is the correctly rounded inverse rather than one carrying elimination round-off.
#pragma once
#include <algorithm>
#include <cmath>
#include <complex>
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include "numth_host.hpp"

namespace mfhe { namespace host {

// table construction is a one-off burst: use the machine's cores even when the launcher exported OMP_NUM_THREADS=1
// (torch.distributed.run does), capped so that eight ranks on one box do not oversubscribe badly
inline int table_threads() {
    const unsigned hw = std::thread::hardware_concurrency();
    return (int)std::max(1u, std::min(hw ? hw : 1u, 16u));
}

inline u64 addmod(u64 a, u64 b, u64 q) { u64 s = a + b; return s >= q ? s - q : s; }
inline u64 submod(u64 a, u64 b, u64 q) { return a >= b ? a - b : a + q - b; }

// exact order-p element for p = 3 f (f an odd prime): smallest generator base, as the reference
inline u64 find_eta(u64 q, int f) {
    const u64 p = 3ull * (u64)f;
    if ((q - 1) % p) return 0;
    const u64 e = (q - 1) / p;
    for (u64 g = 2; g < q; ++g) {
        const u64 eta = powmod(g, e, q);
        if (eta == 1 || powmod(eta, p, q) != 1) continue;
        if (powmod(eta, p / 3, q) == 1 || powmod(eta, p / (u64)f, q) == 1) continue;
        return eta;
    }
    return 0;
}

// units of Z_p in the reference's lane order: (a f + 3 b) mod p, a in {1,2} outer, b in 1..f-1
inline std::vector<uint16_t> unit_exponents(int f) {
    std::vector<uint16_t> e;
    for (int a = 1; a <= 2; ++a)
        for (int b = 1; b < f; ++b) e.push_back((uint16_t)((a * f + 3 * b) % (3 * f)));
    return e;
}

// Phi_{3f}(W) = (W^p - 1)(W - 1) / ((W^3 - 1)(W^f - 1)), integer coefficients, degree 2(f-1)
inline std::vector<int> cyclotomic_3f(int f) {
    const int p = 3 * f;
    std::vector<long> a((size_t)p + 3, 0), t((size_t)p + 3, 0);
    a[p] = 1; a[0] = -1;
    int deg = p;
    for (int d = deg; d >= f; --d) { long c = a[d]; if (!c) continue; t[d - f] += c; a[d] -= c; a[d - f] += c; }
    a = t; deg -= f;
    for (int d = deg + 1; d >= 1; --d) a[d] = a[d - 1] - a[d];
    a[0] = -a[0]; deg += 1;
    std::fill(t.begin(), t.end(), 0);
    for (int d = deg; d >= 3; --d) { long c = a[d]; if (!c) continue; t[d - 3] += c; a[d] -= c; a[d - 3] += c; }
    deg -= 3;
    std::vector<int> phi((size_t)deg + 1);
    for (int i = 0; i <= deg; ++i) {
        if (t[i] < -1 || t[i] > 1) throw std::runtime_error("Phi_{3f} coefficient outside {-1,0,1}");
        phi[i] = (int)t[i];
    }
    return phi;
}

// V[w][r] (row-major [phi][phi]) and Vinv[r][w] (row-major [phi][phi], true inverse) for one limb
inline void build_wcrt_limb(u64 q, int f, const std::vector<uint16_t>& wexp, const std::vector<int>& Phi,
                            u64 eta, std::vector<u64>& V, std::vector<u64>& Vinv) {
    const int p = 3 * f;
    const size_t phi = wexp.size();
    V.assign(phi * phi, 0); Vinv.assign(phi * phi, 0);
    const u64 pinv = invmod((u64)p % q, q), ieta = invmod(eta, q);
#pragma omp parallel for schedule(dynamic, 8) num_threads(table_threads())
    for (long w = 0; w < (long)phi; ++w) {
        u64 root = powmod(eta, wexp[w], q), cur = 1;
        for (size_t r = 0; r < phi; ++r) { V[(size_t)w * phi + r] = cur; cur = mulmod(cur, root, q); }
        std::vector<u64> g((size_t)p);
        root = powmod(ieta, wexp[w], q); cur = pinv;
        for (int j = 0; j < p; ++j) { g[j] = cur; cur = mulmod(cur, root, q); }
        for (int d = p - 1; d >= (int)phi; --d) {
            const u64 co = g[d];
            if (!co) continue;
            for (size_t i = 0; i < phi; ++i) {
                if (!Phi[i]) continue;
                const size_t k = (size_t)d - phi + i;
                g[k] = Phi[i] > 0 ? submod(g[k], co, q) : addmod(g[k], co, q);
            }
        }
        for (size_t r = 0; r < phi; ++r) Vinv[r * phi + (size_t)w] = g[r];
    }
}

typedef std::complex<double> cplx;

// complex V[w][r] and its exact inverse Vinv[r][w]
inline void build_wdft(int f, const std::vector<uint16_t>& wexp, const std::vector<int>& Phi,
                       std::vector<cplx>& V, std::vector<cplx>& Vinv) {
    const int p = 3 * f;
    const size_t phi = wexp.size();
    const long double two_pi = 6.283185307179586476925286766559L;
    V.assign(phi * phi, cplx(0, 0)); Vinv.assign(phi * phi, cplx(0, 0));
#pragma omp parallel for schedule(dynamic, 8) num_threads(table_threads())
    for (long w = 0; w < (long)phi; ++w) {
        std::vector<long double> gr((size_t)p), gi((size_t)p);
        for (size_t r = 0; r < phi; ++r) {
            const long k = ((long)wexp[w] * (long)r) % p;       // exact angle index: no accumulated product
            const long double ang = two_pi * (long double)k / (long double)p;
            V[(size_t)w * phi + r] = cplx((double)cosl(ang), (double)sinl(ang));
        }
        for (int j = 0; j < p; ++j) {
            const long k = ((long)wexp[w] * j) % p;
            const long double ang = -two_pi * (long double)k / (long double)p;
            gr[j] = cosl(ang) / p; gi[j] = sinl(ang) / p;
        }
        for (int d = p - 1; d >= (int)phi; --d)
            for (size_t i = 0; i < phi; ++i) {
                if (!Phi[i]) continue;
                const size_t k = (size_t)d - phi + i;
                gr[k] -= Phi[i] * gr[d]; gi[k] -= Phi[i] * gi[d];
            }
        for (size_t r = 0; r < phi; ++r) Vinv[r * phi + (size_t)w] = cplx((double)gr[r], (double)gi[r]);
    }
}

// XY canonical embedding for X^n = i: V[j][k] = zeta^(5^j k), zeta = e^(2 pi i / 4n); Vinv = conj(V)^T / n
inline void build_xy(int n, std::vector<cplx>& V, std::vector<cplx>& Vinv) {
    const long double two_pi = 6.283185307179586476925286766559L;
    V.assign((size_t)n * n, cplx(0, 0)); Vinv.assign((size_t)n * n, cplx(0, 0));
    u64 e = 1;
    for (int j = 0; j < n; ++j) {
        for (int k = 0; k < n; ++k) {
            const u64 idx = (e * (u64)k) % (u64)(4 * n);
            const long double ang = two_pi * (long double)idx / (long double)(4 * n);
            V[(size_t)j * n + k] = cplx((double)cosl(ang), (double)sinl(ang));
            Vinv[(size_t)k * n + j] = cplx((double)(cosl(ang) / n), (double)(-sinl(ang) / n));
        }
        e = (e * 5) % (u64)(4 * n);
    }
}

struct CrtTables {
    int words = 0;                       // 64-bit words of Q (reference: BIGINT_LIMBS = 7 at L = 11)
    std::vector<u64> Q, Qhalf, M, inv;   // M is [L][words]
};
inline CrtTables build_crt(const std::vector<u64>& moduli) {
    CrtTables t;
    int bits = 0;
    for (u64 q : moduli) { int b = 0; while ((q >> b) != 0) ++b; bits += b; }
    const int W = t.words = std::max(1, (bits + 63) / 64);
    const size_t L = moduli.size();
    t.Q.assign(W, 0); t.Qhalf.assign(W, 0); t.M.assign(L * W, 0); t.inv.assign(L, 0);
    t.Q[0] = 1;
    for (u64 q : moduli) {
        u128 carry = 0;
        for (int i = 0; i < W; ++i) { u128 pr = (u128)t.Q[i] * q + carry; t.Q[i] = (u64)pr; carry = pr >> 64; }
        if (carry) throw std::runtime_error("CRT modulus overflow");
    }
    u64 cy = 0;
    for (int i = W - 1; i >= 0; --i) { t.Qhalf[i] = (t.Q[i] >> 1) | (cy << 63); cy = t.Q[i] & 1; }
    for (size_t l = 0; l < L; ++l) {
        const u64 q = moduli[l];
        u128 rem = 0;
        for (int i = W - 1; i >= 0; --i) { u128 cur = (rem << 64) | t.Q[i]; t.M[l * W + i] = (u64)(cur / q); rem = cur % q; }
        if (rem) throw std::runtime_error("moduli are not pairwise distinct primes");
        rem = 0;
        for (int i = W - 1; i >= 0; --i) { u128 cur = (rem << 64) | t.M[l * W + i]; rem = cur % q; }
        t.inv[l] = invmod((u64)rem, q);
    }
    return t;
}

}}  // namespace mfhe::host
