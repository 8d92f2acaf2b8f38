// Test/bench infrastructure (NOT product): times the UNMODIFIED reference (src/core/*.cu, -rdc, sm_100a) and Phantom's
// kernels on the B200 the bench runs on -- the two GPU baselines BASELINE.json's north_star names.  bench.py runs
// this binary (when baseline/_ref/ref_time exists) and copies its JSON into `baselines.{reference_cuda,phantom}`.
// Timing: host wall clock around each public call with a device synchronise on both sides (the reference's calls
// are host loops of thousands of launches plus cudaMalloc/cudaFree, so launch cost is part of what a user pays);
// best of `reps` after one warm-up.
//   reference: HE.cuh init_he_backend / generate_secret_key / encrypt_pair / decrypt_and_decode,
//              batched_encoder.cuh encode_to_wntt_eval, wntt_forward/inverse_matrix, xy_ntt_*_phantom / _gl, trace_gemm_batched
//   Phantom:   nwt_2d_radix8_forward/backward_inplace (ph/src/ntt/fntt_2d.cu:620-653, intt_2d.cu:724-757), bConv_BEHZ
//              (ph/src/rns_bconv.cu:212-229), PhantomCKKSEncoder (ph/src/ckks.cu:68-190), encrypt_symmetric / decrypt
//              (ph/src/secretkey.cu:232-295, 533-569)
#include "ref_common.h"

#include "core/config.h"
#include "core/HE.cuh"
#include "core/ntt_core.cuh"
#include "core/encoder.cuh"
#include "core/batched_encoder.cuh"
#include "core/batched_trace.cuh"
#include "phantom.h"

#include <chrono>
#include <cuComplex.h>
#include <functional>

#ifndef REF_RDC
#define REF_RDC 1
#endif
using namespace matrix_fhe;
using clk = std::chrono::steady_clock;

static double time_ms(const std::function<void()> &f, int reps, int warm = 1) {
    (void)cudaGetLastError();
    for (int i = 0; i < warm; i++) f();
    CK(cudaDeviceSynchronize());
    double best = 1e300;
    for (int i = 0; i < reps; i++) {
        CK(cudaDeviceSynchronize());
        auto t0 = clk::now();
        f();
        CK(cudaDeviceSynchronize());
        double ms = std::chrono::duration<double, std::milli>(clk::now() - t0).count();
        if (ms < best) best = ms;
    }
    (void)cudaGetLastError();      // the reference leaves non-sticky errors behind and never looks at them
    return best;
}

static FILE *g_out;
static bool g_first = true;
static void kv(const char *k, double v) { fprintf(g_out, "%s\"%s\": %.6g", g_first ? "" : ", ", k, v); g_first = false; fflush(g_out); }

static void time_reference(int reps) {
    const int n = MATRIX_N, phi = BATCH_SIZE, L = RNS_NUM_LIMBS;
    const size_t n2 = (size_t)n * n, T = (size_t)phi * L * n2, tot = (size_t)phi * n2;
    fprintf(g_out, "\"reference_cuda\": {\"shape\": \"n=64 phi=512 L=11 (config.h; the only shape the reference builds for)\", "
            "\"build\": \"%s\", ", REF_RDC ? "-rdc=true (separable compilation)" : "stock (CMakeLists.txt flags, arch 100a)");
    g_first = true;
    auto t0 = clk::now();
    init_he_backend(); CK(cudaDeviceSynchronize());
    kv("init_he_backend_ms", std::chrono::duration<double, std::milli>(clk::now() - t0).count());
    SecretKey sk;
    t0 = clk::now(); generate_secret_key(sk, L); CK(cudaDeviceSynchronize());
    kv("generate_secret_key_ms", std::chrono::duration<double, std::milli>(clk::now() - t0).count());

    std::vector<cuDoubleComplex> hm(tot);
    for (int ell = 0; ell < phi; ell++) for (size_t i = 0; i < n2; i++) {
        double v = (double)ell + 0.001 * (double)i; hm[(size_t)ell * n2 + i] = make_cuDoubleComplex(v, -v);
    }
    cuDoubleComplex *dm, *dout; uint64_t *d_re, *d_im, *d_t;
    CK(cudaMalloc(&dm, tot * 16)); CK(cudaMalloc(&dout, tot * 16));
    CK(cudaMalloc(&d_re, T * 8)); CK(cudaMalloc(&d_im, T * 8)); CK(cudaMalloc(&d_t, T * 8));
    CK(cudaMemcpy(dm, hm.data(), tot * 16, cudaMemcpyHostToDevice));
    RLWECiphertext ct_re, ct_im; allocate_ciphertext(ct_re, L); allocate_ciphertext(ct_im, L);
    BatchedEncoder enc(n);
    double e = time_ms([&] { enc.encode_to_wntt_eval(dm, d_re, d_im); }, reps);
    double c = time_ms([&] { encrypt_pair(d_re, d_im, sk, ct_re, ct_im); }, reps);
    double d = time_ms([&] { decrypt_and_decode(ct_re, ct_im, sk, dout); }, reps);
    kv("encode_to_wntt_eval_ms", e); kv("encrypt_pair_ms", c); kv("decrypt_and_decode_ms", d);
    kv("pipeline_pair_ms", e + c + d);
    std::vector<cuDoubleComplex> ho(tot);
    CK(cudaMemcpy(ho.data(), dout, tot * 16, cudaMemcpyDeviceToHost));
    double me = 0; for (size_t i = 0; i < tot; i++) me = fmax(me, hypot(ho[i].x - hm[i].x, ho[i].y - hm[i].y));
    kv("pipeline_max_err", me);
    // encode -> decode loop (test_encode_decode_wcrt.cu: b = encoded message, a = 0, zero key)
    {
        SecretKey zk; zk.num_limbs = L; size_t TV = (size_t)phi * L * n;
        CK(cudaMalloc(&zk.data, TV * 8)); CK(cudaMemset(zk.data, 0, TV * 8));
        CK(cudaMemset(ct_re.data, 0, 2 * T * 8)); CK(cudaMemset(ct_im.data, 0, 2 * T * 8));
        CK(cudaMemcpy(ct_re.data, d_re, T * 8, cudaMemcpyDeviceToDevice));
        CK(cudaMemcpy(ct_im.data, d_im, T * 8, cudaMemcpyDeviceToDevice));
        double dd = time_ms([&] { decrypt_and_decode(ct_re, ct_im, zk, dout); }, reps);
        kv("encode_decode_loop_ms", e + dd);
        cudaFree(zk.data);
    }
    kv("wntt_forward_matrix_ms", time_ms([&] { wntt_forward_matrix(d_re, d_t, n, L, phi); }, reps));
    kv("wntt_inverse_matrix_ms", time_ms([&] { wntt_inverse_matrix(d_t, d_im, n, L, phi); }, reps));
    kv("xy_ntt_forward_phantom_32768_ms", time_ms([&] { xy_ntt_forward_phantom(d_re, L, phi * n, n); }, reps));
    kv("xy_ntt_backward_phantom_32768_ms", time_ms([&] { xy_ntt_backward_phantom(d_re, L, phi * n, n); }, reps));
    kv("xy_ntt_forward_gl_32768_ms", time_ms([&] { xy_ntt_forward_gl(d_re, d_t, L, phi * n, n); }, reps));
#if REF_RDC     // the trace kernels read d_he_moduli across TUs: zero (a division by zero) in the stock build
    {
        uint64_t *cr, *ci; CK(cudaMalloc(&cr, T * 8)); CK(cudaMalloc(&ci, T * 8));
        kv("trace_gemm_batched_512_ms", time_ms([&] { trace_gemm_batched(d_re, d_im, d_t, d_re, cr, ci, n, L, phi); }, reps));
        cudaFree(cr); cudaFree(ci);
    }
#endif
    fprintf(g_out, "}");
    free_ciphertext(ct_re); free_ciphertext(ct_im);
    cudaFree(dm); cudaFree(dout); cudaFree(d_re); cudaFree(d_im); cudaFree(d_t);
}

static void make_tables(DNTTTable &tab, size_t logn, const std::vector<phantom::arith::Modulus> &mods) {
    const auto &s = cudaStreamPerThread;
    tab.init((size_t)1 << logn, mods.size(), s);
    for (size_t i = 0; i < mods.size(); i++) {
        phantom::arith::NTT t((int)logn, mods[i]);
        DModulus dm(mods[i].value(), mods[i].const_ratio()[0], mods[i].const_ratio()[1]);
        tab.set(&dm, t.get_from_root_powers().data(), t.get_from_root_powers_shoup().data(),
                t.get_from_inv_root_powers().data(), t.get_from_inv_root_powers_shoup().data(),
                t.inv_degree_modulo(), t.inv_degree_modulo_shoup(), i, s);
        CK(cudaStreamSynchronize(s));
    }
}

static void fill_device(uint64_t *d, size_t count, const std::vector<phantom::arith::Modulus> &mods, size_t N) {
    std::vector<uint64_t> h(count);
    size_t L = mods.size();
    for (size_t i = 0; i < count; i++) h[i] = splitmix64_h(42 + i) % mods[(i / N) % L].value();
    CK(cudaMemcpy(d, h.data(), count * 8, cudaMemcpyHostToDevice));
}

static double cx_max_err(const std::vector<std::vector<cuDoubleComplex>> &a, const std::vector<std::vector<cuDoubleComplex>> &b) {
    double me = 0;
    for (size_t p = 0; p < a.size(); p++) for (size_t i = 0; i < a[p].size(); i++) {
        double e = hypot(a[p][i].x - b[p][i].x, a[p][i].y - b[p][i].y);
        if (!(e <= me)) me = (e == e) ? e : 1e300;          // NaN counts as a failure, not as zero
    }
    return me;
}

static void time_phantom_pipeline(const char *tag, size_t L, int first_bits, int bits_rest, double scale, int reps) {
    // Phantom's CKKS at N = 2^16 with L data primes (+1 special prime, the library's default): 64 plaintexts of 2^15
    // slots carry the 2^21 complex values of one matrix-FHE (re, im) pair.
    using namespace phantom;
    const size_t N = 65536, P = 64;
    EncryptionParameters parms(scheme_type::ckks);
    parms.set_poly_modulus_degree(N);
    std::vector<int> bits; bits.push_back(first_bits); for (size_t i = 1; i < L; i++) bits.push_back(bits_rest); bits.push_back(first_bits);
    parms.set_coeff_modulus(arith::CoeffModulus::Create(N, bits));
    PhantomContext context(parms);
    PhantomSecretKey sk(context);
    PhantomCKKSEncoder encoder(context);
    size_t slots = encoder.slot_count();
    std::vector<std::vector<cuDoubleComplex>> in(P, std::vector<cuDoubleComplex>(slots));
    for (size_t p = 0; p < P; p++) for (size_t i = 0; i < slots; i++) {
        double v = (double)(p * 8) + 0.001 * (double)(i % 4096); in[p][i] = make_cuDoubleComplex(v, -v);
    }
    std::vector<PhantomPlaintext> pt(P), pt2(P);
    std::vector<PhantomCiphertext> ct(P);
    std::vector<std::vector<cuDoubleComplex>> out(P);
    double enc = time_ms([&] { for (size_t p = 0; p < P; p++) encoder.encode(context, in[p], scale, pt[p]); }, reps);
    double dec = time_ms([&] { for (size_t p = 0; p < P; p++) encoder.decode(context, pt[p], out[p]); }, reps);
    double me0 = cx_max_err(out, in);
    double ecr = time_ms([&] { for (size_t p = 0; p < P; p++) sk.encrypt_symmetric(context, pt[p], ct[p]); }, reps);
    double dcr = time_ms([&] { for (size_t p = 0; p < P; p++) sk.decrypt(context, ct[p], pt2[p]); }, reps);
    for (size_t p = 0; p < P; p++) encoder.decode(context, pt2[p], out[p]);
    double me = cx_max_err(out, in);
    char k[96];
    snprintf(k, sizeof k, "ckks_%s_encode_64x32768_ms", tag); kv(k, enc);
    snprintf(k, sizeof k, "ckks_%s_decode_64x32768_ms", tag); kv(k, dec);
    snprintf(k, sizeof k, "ckks_%s_encode_decode_max_err", tag); kv(k, me0);
    snprintf(k, sizeof k, "ckks_%s_encrypt_symmetric_64_ms", tag); kv(k, ecr);
    snprintf(k, sizeof k, "ckks_%s_decrypt_64_ms", tag); kv(k, dcr);
    snprintf(k, sizeof k, "ckks_%s_pipeline_pair_ms", tag); kv(k, enc + ecr + dcr + dec);
    snprintf(k, sizeof k, "ckks_%s_pipeline_max_err", tag); kv(k, me);
}

static void time_phantom(int reps) {
    const auto &s = cudaStreamPerThread;
    fprintf(g_out, "\"phantom\": {"); g_first = true;
    {   // config 2: [64][32][65536], 60-bit primes, one call per poly (the API transforms one L-limb poly per call)
        const size_t logn = 16, N = 1 << 16, L = 32, B = 64;
        for (int bits : {60, 50}) {
            auto mods = phantom::arith::CoeffModulus::Create(N, std::vector<int>(L, bits));
            DNTTTable tab; make_tables(tab, logn, mods);
            uint64_t *d; CK(cudaMalloc(&d, B * L * N * 8)); fill_device(d, B * L * N, mods, N);
            double f = time_ms([&] { for (size_t b = 0; b < B; b++) nwt_2d_radix8_forward_inplace(d + b * L * N, tab, L, 0, s); }, reps);
            double i = time_ms([&] { for (size_t b = 0; b < B; b++) nwt_2d_radix8_backward_inplace(d + b * L * N, tab, L, 0, s); }, reps);
            char k[96];
            snprintf(k, sizeof k, "cfg2_%dbit_forward_ms", bits); kv(k, f);
            snprintf(k, sizeof k, "cfg2_%dbit_inverse_ms", bits); kv(k, i);
            snprintf(k, sizeof k, "cfg2_%dbit_limb_poly_pairs_per_s", bits); kv(k, (double)(B * L) / ((f + i) * 1e-3));
            cudaFree(d);
        }
    }
    {   // config 5 on one GPU: per poly INTT over 48 limbs -> bConv_BEHZ 48 -> 16 -> NTT over 16 limbs; batch 16
        const size_t logn = 17, N = 1 << 17, LI = 48, LO = 16, B = 16;
        auto mods = phantom::arith::get_primes(N, 55, LI + LO)   /* SURVEY 8(d) config 5; all 64 happen to be 1 mod 2N */;
        std::vector<phantom::arith::Modulus> qi(mods.begin(), mods.begin() + LI), pj(mods.begin() + LI, mods.end());
        phantom::arith::RNSBase ibase(qi), obase(pj);
        phantom::arith::BaseConverter conv(ibase, obase);
        DBaseConverter dconv(conv, s);
        DNTTTable tab; make_tables(tab, logn, mods);
        uint64_t *d, *o; CK(cudaMalloc(&d, B * LI * N * 8)); CK(cudaMalloc(&o, B * LO * N * 8));
        fill_device(d, B * LI * N, qi, N);
        double bc = time_ms([&] { for (size_t b = 0; b < B; b++) dconv.bConv_BEHZ(o + b * LO * N, d + b * LI * N, N, s); }, reps);
        double st = time_ms([&] { for (size_t b = 0; b < B; b++) {
            nwt_2d_radix8_backward_inplace(d + b * LI * N, tab, LI, 0, s);
            dconv.bConv_BEHZ(o + b * LO * N, d + b * LI * N, N, s);
            nwt_2d_radix8_forward_inplace(o + b * LO * N - LI * N, tab, LO, LI, s); } }, reps);     // kernel indexes inout + (start + i) * N
        kv("cfg5_bconv_batch16_ms", bc); kv("cfg5_step_batch16_ms", st); kv("cfg5_polys_per_s", (double)B / (st * 1e-3));
        cudaFree(d); cudaFree(o);
    }
    // the reference's own parameter style (45-bit first prime, 35-bit rest, scale 2^35: config.h:25-44) and Phantom's
    // (60-bit first / special prime, 40-bit rest, scale 2^40: ph/examples/3_ckks.cu)
    time_phantom_pipeline("L11", 11, 45, 35, 34359738368.0, reps);
    time_phantom_pipeline("L24", 24, 45, 35, 34359738368.0, reps);
    time_phantom_pipeline("L11_60_40", 11, 60, 40, 1099511627776.0, reps);
    fprintf(g_out, "}");
}

int main(int argc, char **argv) {
    const char *path = argc > 1 ? argv[1] : "-";
    int reps = argc > 2 ? atoi(argv[2]) : 3;
    bool do_ref = true, do_ph = true;
    for (int i = 3; i < argc; i++) { if (!strcmp(argv[i], "--no-ref")) do_ref = false; if (!strcmp(argv[i], "--no-phantom")) do_ph = false; }
    g_out = strcmp(path, "-") ? fopen(path, "w") : stdout;
    if (!g_out) { perror(path); return 1; }
    // the reference prints banners on stdout; keep our JSON clean by writing it to a file whenever a path is given
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    fprintf(g_out, "{\"device\": \"%s\", \"timing\": \"host wall clock, device-synchronised, best of %d after 1 warm-up\", ", p.name, reps);
    if (do_ph) { time_phantom(reps); if (do_ref) fprintf(g_out, ", "); }
    if (do_ref) time_reference(reps);
    fprintf(g_out, "}\n");
    if (g_out != stdout) fclose(g_out);
    return 0;
}
