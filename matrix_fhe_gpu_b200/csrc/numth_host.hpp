// numth_host.hpp -- host-side number theory and table construction for the plans.
//
// Product code (not the oracle): builds exactly the tables whose *values* the
// reference's host code defines, so that device results are bit-identical:
//   ph/src/host/numth.cu:265-332   minimal primitive 2N-th root
//   ph/src/host/ntt.cu:11-56       psi powers at bit-reversed indices (+ Shoup words)
//   src/core/ntt_core.cu:49-70     get_psi: smallest base giving an order-4n element
#pragma once
#include <cstdint>
#include <stdexcept>
#include <vector>

namespace mfhe { namespace host {

typedef unsigned __int128 u128;
typedef uint64_t u64;

inline u64 mulmod(u64 a, u64 b, u64 q) { return (u64)((u128)a * b % q); }
inline u64 powmod(u64 b, u64 e, u64 q) {
    u64 r = 1; b %= q;
    while (e) { if (e & 1) r = mulmod(r, b, q); b = mulmod(b, b, q); e >>= 1; }
    return r;
}
inline u64 invmod(u64 a, u64 q) { return powmod(a, q - 2, q); }
inline u64 shoup(u64 w, u64 q) { return (u64)(((u128)w << 64) / q); }
inline uint32_t bitrev(uint32_t x, int bits) {
    uint32_t r = 0;
    for (int i = 0; i < bits; ++i) { r = (r << 1) | (x & 1u); x >>= 1; }
    return r;
}
inline bool is_prime(u64 n) {
    static const u64 bases[12] = {2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37};
    if (n < 2) return false;
    for (u64 b : bases) { if (n == b) return true; if (n % b == 0) return false; }
    u64 d = n - 1; int r = 0;
    while (!(d & 1)) { d >>= 1; ++r; }
    for (u64 b : bases) {
        u64 x = powmod(b, d, n);
        if (x == 1 || x == n - 1) continue;
        bool comp = true;
        for (int j = 1; j < r && comp; ++j) { x = mulmod(x, x, n); if (x == n - 1) comp = false; }
        if (comp) return false;
    }
    return true;
}

// smallest primitive degree-th root of unity mod q (degree a power of two)
inline u64 min_primitive_root(u64 degree, u64 q) {
    if ((q - 1) % degree) return 0;
    u64 root = 0;
    for (u64 x = 2; x < 100000 && !root; ++x) {
        u64 r = powmod(x, (q - 1) / degree, q);
        if (powmod(r, degree >> 1, q) == q - 1) root = r;
    }
    if (!root) return 0;
    u64 sq = mulmod(root, root, q), cur = root, best = root;
    for (u64 i = 0; i < degree / 2; ++i) { if (cur < best) best = cur; cur = mulmod(cur, sq, q); }
    return best;
}

// reference "gl" root: g = base^((q-1)/4n) for the smallest base >= 2 with g^(2n) = -1
inline u64 gl_psi(u64 q, int n) {
    u64 order = 4 * (u64)n;
    if ((q - 1) % order) return 0;
    for (u64 base = 2; base <= 100000; ++base) {
        u64 g = powmod(base, (q - 1) / order, q);
        if (powmod(g, 2 * (u64)n, q) == q - 1) return g;
    }
    return 0;
}

struct TwPair { u64 w, ws; };

// Twiddles of the merged product-tree network for X^n - psi^c0 where psi has order M:
// node h (heap index, h = m + i) carries constant psi^cexp[h]; its twiddle is
// psi^(cexp[h]/2); children get cexp/2 and cexp/2 + M/2.  Inverse twiddles are the
// modular inverses.  For X^n + 1 with M = 2n, c0 = n this reproduces Phantom's
// tw[bitrev(i)] = psi^i layout exactly.
struct TreeTables {
    std::vector<TwPair> tw, itw;   // n entries each (index 0 unused = {1, shoup(1)})
    TwPair ninv, w1n;
};
inline TreeTables build_tree_tables(int logn, u64 q, u64 psi, u64 M, u64 c0) {
    const size_t n = (size_t)1 << logn;
    TreeTables t;
    t.tw.assign(n, TwPair{1, shoup(1, q)});
    t.itw.assign(n, TwPair{1, shoup(1, q)});
    std::vector<u64> cexp(2 * n, 0);
    cexp[1] = c0 % M;
    std::vector<u64> pw((size_t)M);          // pw[i] = psi^i, i in [0, M)
    pw[0] = 1;
    for (size_t i = 1; i < (size_t)M; ++i) pw[i] = mulmod(pw[i - 1], psi, q);
    for (size_t h = 1; h < n; ++h) {
        if (cexp[h] & 1) throw std::runtime_error("tree tables: odd exponent");
        const u64 e = cexp[h] / 2;          // even at every internal node by construction
        const u64 w = pw[e], iw = pw[(M - e) % M];
        t.tw[h] = TwPair{w, shoup(w, q)};
        t.itw[h] = TwPair{iw, shoup(iw, q)};
        cexp[2 * h] = e;
        cexp[2 * h + 1] = (e + M / 2) % M;
    }
    const u64 ni = invmod((u64)n % q, q);
    t.ninv = TwPair{ni, shoup(ni, q)};
    const u64 w1 = mulmod(t.itw[1].w, ni, q);
    t.w1n = TwPair{w1, shoup(w1, q)};
    return t;
}

}}  // namespace mfhe::host
