// ref_host_wrap.cpp -- thin extern "C" view of the REAL reference host code.
//
// TEST INFRASTRUCTURE ONLY.  This file contains no reference code: it calls
// Phantom's own host classes (phantom::arith::NTT, get_primes,
// CoeffModulus::Create), which oracle/Makefile compiles straight from
// /root/reference/extern/phantom-fhe/src/host/*.cu into oracle/_ref/.  It is
// used (a) to pin oracle/ntt_oracle.c against the reference's real table
// builder and (b) by tests/golden/make_golden.py to generate the committed
// fixtures.  /root/reference does not exist on the GPU box; nothing at run
// time there depends on this library.
#include <cstdint>
#include <cstring>
#include <vector>

#include "host/modulus.h"
#include "host/ntt.h"
#include "host/numth.h"

using namespace phantom::arith;

extern "C" {

// ph/src/host/numth.cu:207-230
int ref_get_primes(uint64_t ntt_size, int bit_size, size_t count, uint64_t *out) {
    try {
        auto v = get_primes(ntt_size, bit_size, count);
        for (size_t i = 0; i < v.size(); ++i) out[i] = v[i].value();
        return 0;
    } catch (...) { return -1; }
}

// ph/src/host/modulus.cu:80-112
int ref_coeff_modulus_create(uint64_t n, int bit_size, size_t count, uint64_t *out) {
    try {
        std::vector<int> bits(count, bit_size);
        auto v = CoeffModulus::Create(n, bits);
        for (size_t i = 0; i < v.size(); ++i) out[i] = v[i].value();
        return 0;
    } catch (...) { return -1; }
}

// ph/src/host/ntt.cu:11-56
int ref_ntt_tables(int logn, uint64_t q, uint64_t *root, uint64_t *tw, uint64_t *tws,
                   uint64_t *itw, uint64_t *itws, uint64_t *n_inv, uint64_t *n_inv_shoup) {
    try {
        Modulus m(q);
        NTT t(logn, m);
        size_t n = t.coeff_count();
        *root = t.get_root();
        std::memcpy(tw, t.get_from_root_powers().data(), n * 8);
        std::memcpy(tws, t.get_from_root_powers_shoup().data(), n * 8);
        std::memcpy(itw, t.get_from_inv_root_powers().data(), n * 8);
        std::memcpy(itws, t.get_from_inv_root_powers_shoup().data(), n * 8);
        *n_inv = t.inv_degree_modulo();
        *n_inv_shoup = t.inv_degree_modulo_shoup();
        return 0;
    } catch (...) { return -1; }
}

int ref_is_prime(uint64_t q) {
    try { Modulus m(q); return m.is_prime() ? 1 : 0; } catch (...) { return -1; }
}

}  // extern "C"
