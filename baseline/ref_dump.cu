// Test infrastructure (NOT product): runs the UNMODIFIED reference (src/core/*.cu, built -rdc for sm_100a by
// baseline/build_ref.py) and Phantom's kernels on fixed inputs and writes, per output array, its SHA-256, its leading
// words and 64 sampled words as one JSON object per line.  tests/test_refcuda_golden_gpu.py feeds the same inputs to
// libmfhe_b200.so and compares word for word ("the -rdc reference build on the B200" as second pin).
//
// Inputs are closed-form, so both sides can build them without sharing files:
//   synth(seed, layout)[idx] = splitmix64(seed + idx) mod q_limb(idx)        limb(idx) by layout:
//     MATRIX [w][l][y][x] : (idx / n^2) % L      POLY [(w n + y)][l][x] : (idx / n) % L
// Reference entry points called: HE.cuh:11-112, ntt_core.cuh:19-53, batched_trace.cuh:18-49, trace.cuh:11-25,
// batched_encoder.cuh:9-49; kernels uniform_random_kernel / gaussian_noise_kernel / ternary_secret_kernel
// (HE.cu:564-627, 690-713) by their external symbols.  Phantom: ntt.cuh (nwt_2d_radix8_*), rns_bconv.cuh (bConv_BEHZ).
#include "ref_common.h"

#include "core/config.h"
#include "core/HE.cuh"
#include "core/ntt_core.cuh"
#include "core/encoder.cuh"
#include "core/batched_encoder.cuh"
#include "core/trace.cuh"
#include "core/batched_trace.cuh"
#include "phantom.h"

#include <cuComplex.h>
#include <cmath>
#include <string>

// Two flavours (baseline/build_ref.py): REF_RDC=1 links the reference built with -rdc=true (one instance of every
// cross-TU __constant__ table: needed for the centred W-CRT and the trace kernels, SURVEY fact 8, and for launching
// the reference's sampler kernels from here); REF_RDC=0 links the STOCK build (CMakeLists.txt has no separable
// compilation) and dumps the fp64 encode / decode pipeline only.
#ifndef REF_RDC
#define REF_RDC 1
#endif
#if REF_RDC
namespace matrix_fhe {
__global__ void uniform_random_kernel(uint64_t *data, size_t total_len, int limbs, int n);
__global__ void gaussian_noise_kernel(uint64_t *data, size_t total_len, int limbs, int n);
__global__ void ternary_secret_kernel(uint64_t *s, size_t total_len, int limbs, int n);
extern __device__ __constant__ uint64_t d_he_moduli[RNS_NUM_LIMBS];
extern __device__ __constant__ uint64_t d_rns_moduli[RNS_NUM_LIMBS];
extern __device__ __constant__ uint64_t d_crt_M[RNS_NUM_LIMBS * 7];
extern __device__ __constant__ uint64_t d_crt_Q[7];
extern __device__ __constant__ uint64_t d_crt_inv[RNS_NUM_LIMBS];
__global__ void probe_constants_kernel(uint64_t *out) {
    out[0] = d_he_moduli[0]; out[1] = d_rns_moduli[0]; out[2] = d_crt_M[0]; out[3] = d_crt_Q[0]; out[4] = d_crt_Q[6];
    out[5] = d_crt_inv[0]; out[6] = d_he_moduli[10]; out[7] = d_rns_moduli[10];
}
}
#endif
using namespace matrix_fhe;
#define FLAVOUR (REF_RDC ? "rdc" : "stock")

static FILE *g_out = nullptr;
static const int NS = 64;

template <typename T> static void emit_words(const char *name_, const T *h, size_t count, const char *fmt_note = "") {
    std::string nm = std::string(REF_RDC ? "" : "stock_") + name_; const char *name = nm.c_str();
    Sha256 s; s.update(h, count * sizeof(T));
    fprintf(g_out, "{\"name\": \"%s\", \"count\": %zu, \"bytes_per_word\": %zu, \"sha256\": \"%s\", \"head\": [",
            name, count, sizeof(T), s.hex().c_str());
    for (size_t i = 0; i < 16 && i < count; i++) fprintf(g_out, "%s%llu", i ? ", " : "", (unsigned long long)(uint64_t)h[i]);
    fprintf(g_out, "], \"samples\": [");
    for (int k = 0; k < NS; k++) {
        size_t idx = (size_t)(splitmix64_h(0xABCDEFULL + k) % count);
        fprintf(g_out, "%s[%zu, %llu]", k ? ", " : "", idx, (unsigned long long)(uint64_t)h[idx]);
    }
    fprintf(g_out, "]%s}\n", fmt_note);
    fflush(g_out);
    fprintf(stderr, "  dumped %-34s %12zu words  %s\n", name, count, s.hex().substr(0, 16).c_str());
}

template <typename T> static std::vector<T> d2h(const T *d, size_t count) {
    std::vector<T> h(count);
    CK(cudaMemcpy(h.data(), d, count * sizeof(T), cudaMemcpyDeviceToHost));
    return h;
}
// The reference never checks CUDA return codes, and some of its calls leave a non-sticky error behind (e.g. an
// "invalid argument" from init_he_backend); a kernel fault would surface as a sticky error at the synchronise.
static void soft_check(const char *name) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) fprintf(stderr, "  note: non-sticky CUDA error before %s: %s (ignored like the reference does)\n", name, cudaGetErrorString(e));
}
template <typename T> static void dump_dev(const char *name, const T *d, size_t count) {
    CK(cudaDeviceSynchronize()); soft_check(name);
    auto h = d2h(d, count);
    emit_words(name, h.data(), count);
}
static void dump_cx(const char *name_, const cuDoubleComplex *d, size_t count, const std::vector<cuDoubleComplex> *want) {
    std::string nm = std::string(REF_RDC ? "" : "stock_") + name_; const char *name = nm.c_str();
    CK(cudaDeviceSynchronize()); soft_check(name);
    auto h = d2h(d, count);
    double max_err = -1; size_t worst = 0;
    if (want) for (size_t i = 0; i < count; i++) {
        double e = hypot(h[i].x - (*want)[i].x, h[i].y - (*want)[i].y);
        if (e > max_err) { max_err = e; worst = i; }
    }
    fprintf(g_out, "{\"name\": \"%s\", \"count\": %zu, \"kind\": \"complex128\", \"max_err\": %.17g, \"worst\": %zu, \"samples\": [",
            name, count, max_err, worst);
    for (int k = 0; k < NS; k++) {
        size_t idx = (size_t)(splitmix64_h(0xABCDEFULL + k) % count);
        fprintf(g_out, "%s[%zu, %.17g, %.17g]", k ? ", " : "", idx, h[idx].x, h[idx].y);
    }
    fprintf(g_out, "]}\n"); fflush(g_out);
    fprintf(stderr, "  dumped %-34s %12zu complex   max_err %.3e\n", name, count, max_err);
}

enum { LAY_MATRIX = 0, LAY_POLY = 1 };
static std::vector<uint64_t> synth(uint64_t seed, size_t total, int layout, int n, int L, const uint64_t *q) {
    std::vector<uint64_t> v(total);
    size_t n2 = (size_t)n * n;
    for (size_t i = 0; i < total; i++) {
        int l = layout == LAY_MATRIX ? (int)((i / n2) % L) : (int)((i / n) % L);
        v[i] = splitmix64_h(seed + i) % q[l];
    }
    return v;
}
static uint64_t *h2d(const std::vector<uint64_t> &h) {
    uint64_t *d; CK(cudaMalloc(&d, h.size() * 8)); CK(cudaMemcpy(d, h.data(), h.size() * 8, cudaMemcpyHostToDevice));
    return d;
}

// ---------------------------------------------------------------- the reference's own pipeline, native shape ----
static void dump_reference() {
    const int n = MATRIX_N, phi = BATCH_SIZE, L = RNS_NUM_LIMBS;
    const size_t n2 = (size_t)n * n, T = (size_t)phi * L * n2, TV = (size_t)phi * L * n;
    init_he_backend();
    CK(cudaDeviceSynchronize()); soft_check("init_he_backend");
    uint64_t q[RNS_NUM_LIMBS]; copy_device_moduli(q, L);
    fprintf(g_out, "{\"name\": \"moduli\", \"values\": [");
    for (int l = 0; l < L; l++) fprintf(g_out, "%s%llu", l ? ", " : "", (unsigned long long)q[l]);
    fprintf(g_out, "]}\n");
    const int thr = 256; const int blk = (int)((T + thr - 1) / thr), blkv = (int)((TV + thr - 1) / thr);
    (void)blk; (void)blkv;

    uint64_t *d_a, *d_b, *d_c;
    CK(cudaMalloc(&d_a, T * 8)); CK(cudaMalloc(&d_b, T * 8)); CK(cudaMalloc(&d_c, T * 8));
    SecretKey sk; generate_secret_key(sk, L);                    dump_dev("secret_key", sk.data, TV);
    RLWECiphertext ct_re, ct_im, ct_sum;
    allocate_ciphertext(ct_re, L); allocate_ciphertext(ct_im, L); allocate_ciphertext(ct_sum, L);
#if REF_RDC
    {   // are the cross-TU constant tables one instance, and filled?
        uint64_t *dp; CK(cudaMalloc(&dp, 64)); probe_constants_kernel<<<1, 1>>>(dp); CK(cudaDeviceSynchronize());
        uint64_t hp[8]; CK(cudaMemcpy(hp, dp, 64, cudaMemcpyDeviceToHost)); cudaFree(dp);
        fprintf(g_out, "{\"name\": \"probe_constants\", \"d_he_moduli0\": %llu, \"d_rns_moduli0\": %llu, \"d_crt_M0\": %llu, "
                "\"d_crt_Q0\": %llu, \"d_crt_Q6\": %llu, \"d_crt_inv0\": %llu, \"d_he_moduli10\": %llu, \"d_rns_moduli10\": %llu}\n",
                (unsigned long long)hp[0], (unsigned long long)hp[1], (unsigned long long)hp[2], (unsigned long long)hp[3],
                (unsigned long long)hp[4], (unsigned long long)hp[5], (unsigned long long)hp[6], (unsigned long long)hp[7]);
    }
    // samplers (HE.cu:564-627, 690-713)
    uniform_random_kernel<<<blk, thr>>>(d_a, T, L, n);           dump_dev("uniform_random", d_a, T);
    gaussian_noise_kernel<<<blk, thr>>>(d_b, T, L, n);           dump_dev("gaussian_noise", d_b, T);
    ternary_secret_kernel<<<blkv, thr>>>(d_c, TV, L, n);         dump_dev("ternary_secret", d_c, TV);
    // W-CRT forward of the uniform stream (matrix-major in, poly-major out; HE.cu:716-747), inverse of that (HE.cu:751-781)
    wntt_forward_matrix(d_a, d_b, n, L, phi);                    dump_dev("wcrt_forward_uniform", d_b, T);
    wntt_inverse_matrix(d_b, d_c, n, L, phi);                    dump_dev("wcrt_inverse_of_forward", d_c, T);
    // W-CRT inverse of an independent poly-major array
    { auto h = synth(5005, T, LAY_POLY, n, L, q); CK(cudaMemcpy(d_a, h.data(), T * 8, cudaMemcpyHostToDevice)); }
    wntt_inverse_matrix(d_a, d_c, n, L, phi);                    dump_dev("wcrt_inverse_synth5005", d_c, T);

    // X transforms (ntt_core.cuh): Phantom convention and X^n - i ("gl") on 1024 polys x L limbs
    {
        const int cnt = 1024; size_t tot = (size_t)cnt * L * n;
        auto h = synth(6006, tot, LAY_POLY, n, L, q);
        uint64_t *d = h2d(h), *tmp; CK(cudaMalloc(&tmp, tot * 8));
        xy_ntt_forward_phantom(d, L, cnt, n);                    dump_dev("xy_forward_phantom", d, tot);
        xy_ntt_backward_phantom(d, L, cnt, n);                   dump_dev("xy_backward_phantom", d, tot);
        xy_ntt_forward_gl(d, tmp, L, cnt, n);                    dump_dev("xy_forward_gl", d, tot);
        xy_ntt_backward_gl(d, tmp, L, cnt, n);                   dump_dev("xy_backward_gl", d, tot);
        apply_gl_perm(d, tmp, L, cnt, n, false);                 dump_dev("gl_perm", tmp, tot);
        apply_gl_perm(d, tmp, L, cnt, n, true);                  dump_dev("gl_perm_inverse", tmp, tot);
        cudaFree(d); cudaFree(tmp);
    }

    // encrypt_pair on fixed integer messages (HE.cu:1455-1552), decrypt_to_eval_matrix (HE.cu:1603-1617)
    {
        auto hre = synth(1001, T, LAY_MATRIX, n, L, q), him = synth(2002, T, LAY_MATRIX, n, L, q);
        CK(cudaMemcpy(d_a, hre.data(), T * 8, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(d_b, him.data(), T * 8, cudaMemcpyHostToDevice));
        encrypt_pair(d_a, d_b, sk, ct_re, ct_im);
        dump_dev("encrypt_pair_ct_re", ct_re.data, 2 * T);
        dump_dev("encrypt_pair_ct_im", ct_im.data, 2 * T);
        decrypt_to_eval_matrix(ct_re, sk, d_c);                  dump_dev("decrypt_to_eval_matrix_re", d_c, T);
        RLWECiphertext ct1; allocate_ciphertext(ct1, L);
        encrypt(d_a, sk, ct1);                                   dump_dev("encrypt_single_ct", ct1.data, 2 * T);
        free_ciphertext(ct1);
    }
    // ct add / tensor product (HE.cu:1710-1740)
    add_ciphertexts(ct_re, ct_im, ct_sum);                       dump_dev("add_ciphertexts", ct_sum.data, 2 * T);
    multiply_ciphertexts_raw(ct_re, ct_im, d_a, d_b, d_c);
    dump_dev("mul_raw_d0", d_a, T); dump_dev("mul_raw_d1", d_b, T); dump_dev("mul_raw_d2", d_c, T);

    // centred W-CRT on the driver's pattern (test_wcrt_roundtrip.cu:34-72)
    {
        size_t tot = (size_t)phi * n2;
        std::vector<int64_t> hc(tot);
        for (int w = 0; w < phi; w++) for (int y = 0; y < n; y++) for (int x = 0; x < n; x++)
            hc[(size_t)w * n2 + (size_t)y * n + x] = (int64_t)((w + x + y) % 17) - 8;
        int64_t *dc, *de, *dr;
        CK(cudaMalloc(&dc, tot * 8)); CK(cudaMalloc(&de, tot * 8)); CK(cudaMalloc(&dr, tot * 8));
        CK(cudaMemcpy(dc, hc.data(), tot * 8, cudaMemcpyHostToDevice));
        wntt_forward_centered(dc, de, n, phi, 0);                dump_dev("wcrt_forward_centered", (uint64_t *)de, tot);
        wntt_inverse_centered(de, dr, n, phi, 0);                dump_dev("wcrt_inverse_centered", (uint64_t *)dr, tot);
        auto hr = d2h(dr, tot);
        int64_t max_err = 0; size_t worst = 0;
        for (size_t i = 0; i < tot; i++) { int64_t d = hc[i] - hr[i]; if (d < 0) d = -d; if (d > max_err) { max_err = d; worst = i; } }
        fprintf(g_out, "{\"name\": \"wcrt_roundtrip_driver\", \"max_err\": %lld, \"worst_idx\": %zu}\n", (long long)max_err, worst);
        cudaFree(dc); cudaFree(de); cudaFree(dr);
    }

    // trace step, batched over 16 lanes and single lane (batched_trace.cu:37-197, trace.cu:30-161)
    {
        const int batch = 16; size_t tot = (size_t)batch * L * n2;
        uint64_t *m[4];
        for (int k = 0; k < 4; k++) m[k] = h2d(synth(3001 + k, tot, LAY_MATRIX, n, L, q));
        uint64_t *bpr, *bpi, *cr, *ci;
        CK(cudaMalloc(&bpr, tot * 8)); CK(cudaMalloc(&bpi, tot * 8)); CK(cudaMalloc(&cr, tot * 8)); CK(cudaMalloc(&ci, tot * 8));
        map_B_to_Bprime_batched(m[2], m[3], bpr, bpi, n, L, batch);
        dump_dev("trace_map_batched_re", bpr, tot); dump_dev("trace_map_batched_im", bpi, tot);
        trace_gemm_batched(m[0], m[1], bpr, bpi, cr, ci, n, L, batch);
        dump_dev("trace_gemm_batched_re", cr, tot); dump_dev("trace_gemm_batched_im", ci, tot);
        rescale_by_delta_batched(cr, ci, n, L, batch, 123456789ULL, 987654321ULL, 555555555ULL);
        dump_dev("trace_rescale_batched_re", cr, tot); dump_dev("trace_rescale_batched_im", ci, tot);
        size_t one = (size_t)L * n2;
        map_B_to_Bprime_Xinv_twist(m[2], m[3], bpr, bpi, n, L);
        trace_gemm_ABpT_rns(m[0], m[1], bpr, bpi, cr, ci, n, L);
        dump_dev("trace_gemm_single_re", cr, one); dump_dev("trace_gemm_single_im", ci, one);
        for (int k = 0; k < 4; k++) cudaFree(m[k]);
        cudaFree(bpr); cudaFree(bpi); cudaFree(cr); cudaFree(ci);
    }

#endif   // REF_RDC
    // fp64 pipeline: encode -> encrypt_pair -> decrypt_and_decode on the e2e driver's message
    // (test_encode_encrypt_decrypt_decode_wcrt.cu:45-51); tolerance-compared, not bit-compared
    {
        size_t tot = (size_t)phi * n2;
        std::vector<cuDoubleComplex> hm(tot);
        for (int ell = 0; ell < phi; ell++) for (size_t i = 0; i < n2; i++) {
            double v = (double)ell + 0.001 * (double)i; hm[(size_t)ell * n2 + i] = make_cuDoubleComplex(v, -v);
        }
        cuDoubleComplex *dm, *dout;
        CK(cudaMalloc(&dm, tot * 16)); CK(cudaMalloc(&dout, tot * 16));
        CK(cudaMemcpy(dm, hm.data(), tot * 16, cudaMemcpyHostToDevice));
        BatchedEncoder enc(n);
        enc.encode_to_wntt_eval(dm, d_a, d_b);
        dump_dev("encode_e2e_re", d_a, T); dump_dev("encode_e2e_im", d_b, T);
        encrypt_pair(d_a, d_b, sk, ct_re, ct_im);
        decrypt_and_decode(ct_re, ct_im, sk, dout);
        dump_cx("decrypt_and_decode_e2e", dout, tot, &hm);
        // decode of an exactly-known integer ciphertext: b = encode output, a = 0, zero key (test_encode_decode_wcrt.cu)
        CK(cudaMemset(ct_re.data, 0, 2 * T * 8)); CK(cudaMemset(ct_im.data, 0, 2 * T * 8));
        CK(cudaMemcpy(ct_re.data, d_a, T * 8, cudaMemcpyDeviceToDevice));
        CK(cudaMemcpy(ct_im.data, d_b, T * 8, cudaMemcpyDeviceToDevice));
        SecretKey zk; zk.num_limbs = L; CK(cudaMalloc(&zk.data, TV * 8)); CK(cudaMemset(zk.data, 0, TV * 8));
        decrypt_and_decode(ct_re, ct_im, zk, dout);
        dump_cx("decode_zero_key_e2e", dout, tot, &hm);
        // decode of a synthetic small-integer eval pair: W-CRT forward of centred ints |v| < 2^40 -> decode is exact up to fp64
        cudaFree(dm); cudaFree(dout); cudaFree(zk.data);
    }
    free_ciphertext(ct_re); free_ciphertext(ct_im); free_ciphertext(ct_sum);
    cudaFree(d_a); cudaFree(d_b); cudaFree(d_c);
}

// ---------------------------------------------------------------- Phantom: configs 2 and 5 ----
static void make_tables(DNTTTable &tab, size_t logn, const std::vector<phantom::arith::Modulus> &mods) {
    const auto &s = cudaStreamPerThread;
    size_t N = (size_t)1 << logn;
    tab.init(N, mods.size(), s);
    for (size_t i = 0; i < mods.size(); i++) {             // same calls as ph/src/context.cu:170-183
        phantom::arith::NTT t((int)logn, mods[i]);
        DModulus dm(mods[i].value(), mods[i].const_ratio()[0], mods[i].const_ratio()[1]);
        tab.set(&dm, t.get_from_root_powers().data(), t.get_from_root_powers_shoup().data(),
                t.get_from_inv_root_powers().data(), t.get_from_inv_root_powers_shoup().data(),
                t.inv_degree_modulo(), t.inv_degree_modulo_shoup(), i, s);
        CK(cudaStreamSynchronize(s));
    }
}

static void dump_phantom() {
    const auto &s = cudaStreamPerThread;
    {   // config 2: N = 2^16, 32 x 60-bit, polys 0 and 63 of x[b][l][j] = splitmix64(42 + flat) mod q_l
        const size_t logn = 16, N = 1 << 16, L = 32;
        auto mods = phantom::arith::CoeffModulus::Create(N, std::vector<int>(L, 60));
        fprintf(g_out, "{\"name\": \"phantom_cfg2_moduli\", \"values\": [");
        for (size_t l = 0; l < L; l++) fprintf(g_out, "%s%llu", l ? ", " : "", (unsigned long long)mods[l].value());
        fprintf(g_out, "]}\n");
        DNTTTable tab; make_tables(tab, logn, mods);
        for (size_t b : {(size_t)0, (size_t)63}) {
            std::vector<uint64_t> h(L * N);
            for (size_t l = 0; l < L; l++) for (size_t j = 0; j < N; j++)
                h[l * N + j] = splitmix64_h(42 + (b * L + l) * N + j) % mods[l].value();
            uint64_t *d = h2d(h);
            nwt_2d_radix8_forward_inplace(d, tab, L, 0, s); CK(cudaStreamSynchronize(s));
            dump_dev(b == 0 ? "phantom_cfg2_forward_poly0" : "phantom_cfg2_forward_poly63", d, L * N);
            nwt_2d_radix8_backward_inplace(d, tab, L, 0, s); CK(cudaStreamSynchronize(s));
            dump_dev(b == 0 ? "phantom_cfg2_roundtrip_poly0" : "phantom_cfg2_roundtrip_poly63", d, L * N);
            cudaFree(d);
        }
    }
    {   // config 5: N = 2^17, get_primes(2N?, 55, 64): SURVEY 8(d) names get_primes(131072, 55, 64); 48 in, 16 out
        const size_t logn = 17, N = 1 << 17, LI = 48, LO = 16;
        auto mods = phantom::arith::get_primes(131072, 55, LI + LO);
        fprintf(g_out, "{\"name\": \"phantom_cfg5_moduli\", \"values\": [");
        for (size_t l = 0; l < LI + LO; l++) fprintf(g_out, "%s%llu", l ? ", " : "", (unsigned long long)mods[l].value());
        fprintf(g_out, "]}\n");
        bool ok2n = true;
        for (auto &m : mods) ok2n = ok2n && ((m.value() - 1) % (2 * N) == 0);
        fprintf(g_out, "{\"name\": \"phantom_cfg5_primes_support_2N\", \"value\": %s}\n", ok2n ? "true" : "false");
        std::vector<phantom::arith::Modulus> qi(mods.begin(), mods.begin() + LI), pj(mods.begin() + LI, mods.end());
        phantom::arith::RNSBase ibase(qi), obase(pj);
        phantom::arith::BaseConverter conv(ibase, obase);
        DBaseConverter dconv(conv, s); CK(cudaStreamSynchronize(s));
        std::vector<uint64_t> h(LI * N);
        for (size_t l = 0; l < LI; l++) for (size_t j = 0; j < N; j++) h[l * N + j] = splitmix64_h(42 + l * N + j) % qi[l].value();
        uint64_t *d = h2d(h), *o; CK(cudaMalloc(&o, LO * N * 8));
        dconv.bConv_BEHZ(o, d, N, s); CK(cudaStreamSynchronize(s));
        dump_dev("phantom_cfg5_bconv_poly0", o, LO * N);
        if (ok2n) {
            DNTTTable tab; make_tables(tab, logn, mods);
            nwt_2d_radix8_backward_inplace(d, tab, LI, 0, s); CK(cudaStreamSynchronize(s));
            dump_dev("phantom_cfg5_intt_poly0", d, LI * N);
            dconv.bConv_BEHZ(o, d, N, s); CK(cudaStreamSynchronize(s));
            dump_dev("phantom_cfg5_intt_bconv_poly0", o, LO * N);
            nwt_2d_radix8_forward_inplace(o - LI * N, tab, LO, LI, s); CK(cudaStreamSynchronize(s));   // the kernel indexes inout + (start + i) * N
            dump_dev("phantom_cfg5_step_poly0", o, LO * N);
        }
        cudaFree(d); cudaFree(o);
    }
}

int main(int argc, char **argv) {
    const char *path = argc > 1 ? argv[1] : "refcuda_golden.jsonl";
    g_out = fopen(path, "w");
    if (!g_out) { perror(path); return 1; }
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    fprintf(g_out, "{\"name\": \"meta\", \"device\": \"%s\", \"sm\": %d, \"flavour\": \"%s\", \"what\": \"outputs of the unmodified reference built for sm_100a\"}\n",
            p.name, p.major * 10 + p.minor, FLAVOUR);
    bool do_ref = true, do_ph = true;
    for (int i = 2; i < argc; i++) { if (!strcmp(argv[i], "--no-ref")) do_ref = false; if (!strcmp(argv[i], "--no-phantom")) do_ph = false; }
    if (do_ph && REF_RDC) dump_phantom();
    if (do_ref) dump_reference();
    fclose(g_out);
    fprintf(stderr, "wrote %s\n", path);
    return 0;
}
