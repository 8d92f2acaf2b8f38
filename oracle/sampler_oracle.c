/* sampler_oracle.c -- TEST INFRASTRUCTURE: CPU restatement of the sound sampler option.
 * Not a restatement of reference code -- the reference has no sound sampler (HE.cu:564-627, 690-713 are index hashes;
 * Phantom's ph/src/prng.cu is blake2-based) -- but of the published algorithm the product uses: Philox4x32-10
 * (Salmon, Moraes, Dror, Shaw: "Parallel random numbers: as easy as 1, 2, 3", SC'11; the Random123 known-answer vectors
 * pin it in tests/test_sampler.py) and inversion sampling of the discrete gaussian from a 64-bit cumulative table. */
#include <math.h>
#include <stdint.h>
#include <stddef.h>
#include "mfhe_oracle.h"

void mfo_philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c[0], p1 = (uint64_t)0xCD9E8D57u * c[2];
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0, n1 = (uint32_t)p1, n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1, n3 = (uint32_t)p0;
        c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
}
static void draw128(uint64_t seed0, uint64_t seed1, uint64_t draw, uint32_t tag, uint64_t index, uint32_t out[4]) {
    out[0] = (uint32_t)index ^ (uint32_t)seed1; out[1] = (uint32_t)(index >> 32) ^ (uint32_t)(seed1 >> 32);
    out[2] = (uint32_t)draw; out[3] = (tag << 24) | ((uint32_t)(draw >> 32) & 0xFFFFFFu);
    mfo_philox4x32_10(out, (uint32_t)seed0, (uint32_t)(seed0 >> 32));
}
/* c[i] = floor(2^64 P(Z <= i - 41)), Z ~ D_{Z, sigma = 3.2} cut at |z| <= 41 */
void mfo_sound_cdt(uint64_t *out) {
    long double w[83], sum = 0, acc = 0;
    for (int i = 0; i < 83; ++i) { long double z = i - 41; w[i] = expl(-z * z / (2.0L * 3.2L * 3.2L)); sum += w[i]; }
    for (int i = 0; i < 83; ++i) {
        acc += w[i] / sum;
        long double v = ldexpl(acc, 64);
        out[i] = (i == 82 || v >= 18446744073709551615.0L) ? ~0ull : (uint64_t)v;
    }
}
/* kind 0 uniform [w][l][y][x], 1 gaussian (same layout, one draw per [w][y][x]), 2 ternary [w][l][x] (one draw per [w][x]) */
void mfo_sound_sample(int kind, uint64_t *out, const uint64_t *moduli, int n, int limbs, int phi, uint64_t seed0, uint64_t seed1,
                      uint64_t draw) {
    const size_t n2 = (size_t)n * n, per = (size_t)limbs * n2;
    uint32_t r[4];
    if (kind == 2) {
        for (size_t idx = 0; idx < (size_t)phi * n; ++idx) {
            size_t w = idx / n, x = idx % n;
            draw128(seed0, seed1, draw, 3u, idx, r);
            uint32_t t = (uint32_t)(((uint64_t)r[0] + r[1] + r[2] + r[3]) % 3u);
            for (int l = 0; l < limbs; ++l) out[(w * limbs + l) * n + x] = t == 0 ? 0 : t == 1 ? 1 : moduli[l] - 1;
        }
    } else if (kind == 0) {
        for (size_t idx = 0; idx < (size_t)phi * per; ++idx) {
            draw128(seed0, seed1, draw, 1u, idx, r);
            unsigned __int128 v = ((unsigned __int128)(((uint64_t)r[3] << 32) | r[2]) << 64) | (((uint64_t)r[1] << 32) | r[0]);
            out[idx] = (uint64_t)(v % moduli[(idx % per) / n2]);
        }
    } else {
        uint64_t cdt[83];
        mfo_sound_cdt(cdt);
        for (size_t idx = 0; idx < (size_t)phi * n2; ++idx) {
            size_t w = idx / n2, pos = idx - w * n2;
            draw128(seed0, seed1, draw, 2u, idx, r);
            uint64_t u = ((uint64_t)r[1] << 32) | r[0];
            int i = 0;
            while (i < 82 && u >= cdt[i]) ++i;          /* smallest i with u < cdt[i] */
            int z = i - 41;
            for (int l = 0; l < limbs; ++l) out[w * per + (size_t)l * n2 + pos] = z >= 0 ? (uint64_t)z : moduli[l] - (uint64_t)(-z);
        }
    }
}
