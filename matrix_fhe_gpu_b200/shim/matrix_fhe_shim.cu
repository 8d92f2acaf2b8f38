// matrix_fhe_shim.cu -- implementation of the reference-compatible C++ API (include/core/*.cuh,
// namespace matrix_fhe) on top of the C ABI of libmfhe_b200.so.  This is the whole "binding" a
// maintainer of the reference needs: link the drivers against this object + libmfhe_b200.so
// instead of src/core/*.cu + Phantom.  See INTEGRATION.md.
//
// Behaviour kept from the reference: process-global state created lazily (every entry point
// initialises the backend, so the two CMake-built drivers that forget init_he_backend() work),
// void functions, fatal errors print to stderr and exit(1).  Behaviour NOT kept: per-call
// cudaMalloc/cudaFree of scratch and the device-wide syncs they imply; per-lane host loops.
#include <cstdio>
#include <cstdlib>
#include <map>
#include <utility>

#include "../../include/core/HE.cuh"
#include "../../include/core/batched_encoder.cuh"
#include "../../include/core/batched_trace.cuh"
#include "../../include/core/encoder.cuh"
#include "../../include/core/ntt_core.cuh"
#include "../../include/core/trace.cuh"
#include "../../include/mfhe.h"

class DNTTTable {                       // opaque handle type behind get_xy_ntt_table()
public:
    const mfhe_ntt_plan* plan = nullptr;
};

namespace matrix_fhe {
namespace {

mfhe_ctx* g_ctx = nullptr;
std::map<std::pair<int, int>, mfhe_ntt_plan*> g_plans;       // (n, convention) -> plan over RNS_MODULI
DNTTTable g_xy_table;
uint32_t *g_gl_perm = nullptr, *g_gl_inv_perm = nullptr;

[[noreturn]] void die(const char* where) {
    std::fprintf(stderr, "Error: %s: %s\n", where, mfhe_last_error());
    std::exit(1);
}
inline void ok(int rc, const char* where) { if (rc != MFHE_OK) die(where); }

mfhe_ctx* ctx() {
    if (!g_ctx) {
        mfhe_params p;
        p.n = MATRIX_N; p.f = BATCH_PRIME_P / 3; p.num_limbs = RNS_NUM_LIMBS; p.moduli = RNS_MODULI;
        p.delta = SCALING_FACTOR; p.arena_bytes = 0;
        ok(mfhe_ctx_create(&p, &g_ctx), "init_he_backend");
        g_xy_table.plan = mfhe_ctx_xy_plan(g_ctx);
    }
    return g_ctx;
}
void need_native(int n, int limbs, int phi, const char* where) {
    if (n != MATRIX_N || limbs != RNS_NUM_LIMBS || phi != BATCH_SIZE) {
        std::fprintf(stderr, "Error: %s: (n, limbs, phi) = (%d, %d, %d) differs from the compiled parameter set\n", where, n, limbs, phi);
        std::exit(1);
    }
}
const mfhe_ntt_plan* plan_for(int n, int convention) {
    if (n == MATRIX_N && convention == MFHE_NTT_PHANTOM_NEGACYCLIC) return mfhe_ctx_xy_plan(ctx());
    auto key = std::make_pair(n, convention);
    auto it = g_plans.find(key);
    if (it != g_plans.end()) return it->second;
    int logn = 0;
    while ((1 << logn) < n) ++logn;
    if ((1 << logn) != n) { std::fprintf(stderr, "Error: NTT size must be a power of two\n"); std::exit(1); }
    mfhe_ntt_plan* p = nullptr;
    ok(mfhe_ntt_plan_create(logn, RNS_NUM_LIMBS, RNS_MODULI, convention, &p), "ntt plan");
    return g_plans[key] = p;
}
void small_ntt(bool fwd, int convention, uint64_t* data, int limbs, int batch_count, int n, cudaStream_t st, const char* where) {
    if (limbs < 1 || limbs > RNS_NUM_LIMBS) { std::fprintf(stderr, "Error: %s: bad limb count\n", where); std::exit(1); }
    const mfhe_ntt_plan* p = plan_for(n, convention);
    ok(fwd ? mfhe_ntt_forward_limbs(p, data, (size_t)batch_count, 0, limbs, st)
           : mfhe_ntt_inverse_limbs(p, data, (size_t)batch_count, 0, limbs, st), where);
}

}  // namespace

// ---- HE.cuh --------------------------------------------------------------------------------------------
void init_he_backend() { (void)ctx(); }
void copy_device_moduli(uint64_t* h_out, int count) {
    if (count <= 0 || count > RNS_NUM_LIMBS) { std::fprintf(stderr, "Error: invalid moduli count\n"); std::exit(1); }
    ok(mfhe_ctx_get_moduli(ctx(), h_out, count), "copy_device_moduli");
}
const DNTTTable& get_ntt_table() {
    std::fprintf(stderr, "Error: get_ntt_table(): the scheme moduli admit no %d-point negacyclic NTT\n", POLY_N);
    std::exit(1);
}
const DNTTTable& get_xy_ntt_table() { (void)ctx(); return g_xy_table; }

void wntt_forward_matrix(const uint64_t* in, uint64_t* out, int n, int limbs, int phi, cudaStream_t s) {
    need_native(n, limbs, phi, "wntt_forward_matrix");
    ok(mfhe_wcrt_forward(ctx(), in, out, MFHE_LAYOUT_MATRIX, MFHE_LAYOUT_POLY, s), "wntt_forward_matrix");
}
void wntt_inverse_matrix(const uint64_t* in, uint64_t* out, int n, int limbs, int phi, cudaStream_t s) {
    need_native(n, limbs, phi, "wntt_inverse_matrix");
    ok(mfhe_wcrt_inverse(ctx(), in, out, MFHE_LAYOUT_POLY, MFHE_LAYOUT_MATRIX, s), "wntt_inverse_matrix");
}
void wntt_forward_centered(const int64_t* in, int64_t* out, int n, int phi, cudaStream_t s) {
    need_native(n, RNS_NUM_LIMBS, phi, "wntt_forward_centered");
    ok(mfhe_wcrt_forward_centered(ctx(), in, out, s), "wntt_forward_centered");
}
void wntt_inverse_centered(const int64_t* in, int64_t* out, int n, int phi, cudaStream_t s) {
    need_native(n, RNS_NUM_LIMBS, phi, "wntt_inverse_centered");
    ok(mfhe_wcrt_inverse_centered(ctx(), in, out, s), "wntt_inverse_centered");
}
void wdft_forward_centered_pair(const int64_t* re, const int64_t* im, double* ore, double* oim, int n, int phi, cudaStream_t s) {
    need_native(n, RNS_NUM_LIMBS, phi, "wdft_forward_centered_pair");
    ok(mfhe_wdft_forward_centered_pair(ctx(), re, im, ore, oim, s), "wdft_forward_centered_pair");
}
void wdft_inverse_pair(const double* re, const double* im, double* ore, double* oim, int n, int phi, cudaStream_t s) {
    need_native(n, RNS_NUM_LIMBS, phi, "wdft_inverse_pair");
    ok(mfhe_wdft_inverse_pair(ctx(), re, im, ore, oim, s), "wdft_inverse_pair");
}

void allocate_ciphertext(RLWECiphertext& ct, int limbs) {
    ct.num_limbs = limbs; ct.is_ntt = false;
    const size_t bytes = 2 * (size_t)BATCH_SIZE * MATRIX_N * limbs * MATRIX_N * sizeof(uint64_t);
    if (cudaMalloc(&ct.data, bytes) != cudaSuccess) { std::fprintf(stderr, "Error: allocate_ciphertext: out of memory\n"); std::exit(1); }
    cudaMemset(ct.data, 0, bytes);
}
void free_ciphertext(RLWECiphertext& ct) {
    if (ct.data) { cudaFree(ct.data); ct.data = nullptr; }
}
void generate_secret_key(SecretKey& sk, int limbs) {
    need_native(MATRIX_N, limbs, BATCH_SIZE, "generate_secret_key");
    sk.num_limbs = limbs;
    if (cudaMalloc(&sk.data, (size_t)BATCH_SIZE * limbs * MATRIX_N * sizeof(uint64_t)) != cudaSuccess) {
        std::fprintf(stderr, "Error: generate_secret_key: out of memory\n"); std::exit(1);
    }
    ok(mfhe_keygen(ctx(), sk.data, nullptr), "generate_secret_key");
}
void encrypt(const uint64_t* m, const SecretKey& sk, RLWECiphertext& ct) {
    need_native(MATRIX_N, ct.num_limbs, BATCH_SIZE, "encrypt");
    ok(mfhe_encrypt(ctx(), m, sk.data, ct.data, nullptr), "encrypt");
}
void encrypt_pair(const uint64_t* re, const uint64_t* im, const SecretKey& sk, RLWECiphertext& ct_re, RLWECiphertext& ct_im) {
    need_native(MATRIX_N, ct_re.num_limbs, BATCH_SIZE, "encrypt_pair");
    ok(mfhe_encrypt_pair(ctx(), re, im, sk.data, ct_re.data, ct_im.data, nullptr), "encrypt_pair");
}
void decrypt_and_decode(const RLWECiphertext& ct_re, const RLWECiphertext& ct_im, const SecretKey& sk, cuDoubleComplex* out) {
    need_native(MATRIX_N, ct_re.num_limbs, BATCH_SIZE, "decrypt_and_decode");
    ok(mfhe_decrypt_decode(ctx(), ct_re.data, ct_im.data, sk.data, (double*)out, nullptr), "decrypt_and_decode");
}
void decrypt_to_eval_matrix(const RLWECiphertext& ct, const SecretKey& sk, uint64_t* out) {
    need_native(MATRIX_N, ct.num_limbs, BATCH_SIZE, "decrypt_to_eval_matrix");
    ok(mfhe_decrypt_to_eval(ctx(), ct.data, sk.data, out, MFHE_LAYOUT_MATRIX, nullptr), "decrypt_to_eval_matrix");
}
void add_ciphertexts(const RLWECiphertext& c1, const RLWECiphertext& c2, RLWECiphertext& res) {
    need_native(MATRIX_N, c1.num_limbs, BATCH_SIZE, "add_ciphertexts");
    ok(mfhe_ct_add(ctx(), c1.data, c2.data, res.data, /*reference_indexing=*/1, nullptr), "add_ciphertexts");
}
void multiply_ciphertexts_raw(const RLWECiphertext& c1, const RLWECiphertext& c2, uint64_t* d0, uint64_t* d1, uint64_t* d2) {
    need_native(MATRIX_N, c1.num_limbs, BATCH_SIZE, "multiply_ciphertexts_raw");
    ok(mfhe_ct_mul_raw(ctx(), c1.data, c2.data, d0, d1, d2, /*reference_indexing=*/1, nullptr), "multiply_ciphertexts_raw");
}

// ---- ntt_core.cuh ----------------------------------------------------------------------------------------
void init_ntt_moduli_manual(const uint64_t*) { (void)ctx(); }              // moduli are part of the context
void init_ntt_tables_manual(int n, int) { (void)plan_for(n, MFHE_NTT_CYCLIC_NATURAL); }
void init_gl_twist_tables(int n, int) { (void)plan_for(n, MFHE_NTT_GL_XN_MINUS_I); }
void init_gl_perm_tables(int n) {
    if (g_gl_perm) return;
    if (n & (n - 1)) { std::fprintf(stderr, "Error: GL perm requires power-of-two n\n"); std::exit(1); }
    std::vector<uint32_t> p(n), ip(n);
    ok(mfhe_gl_perm_tables(n, p.data(), ip.data()), "init_gl_perm_tables");
    cudaMalloc(&g_gl_perm, n * sizeof(uint32_t)); cudaMalloc(&g_gl_inv_perm, n * sizeof(uint32_t));
    cudaMemcpy(g_gl_perm, p.data(), n * sizeof(uint32_t), cudaMemcpyHostToDevice);
    cudaMemcpy(g_gl_inv_perm, ip.data(), n * sizeof(uint32_t), cudaMemcpyHostToDevice);
}
const NTTTable& get_manual_ntt_table() {
    static NTTTable t = {nullptr, nullptr, nullptr, nullptr, nullptr, RNS_NUM_LIMBS, MATRIX_N};
    return t;
}
const uint32_t* get_gl_perm() { return g_gl_perm; }
const uint32_t* get_gl_inv_perm() { return g_gl_inv_perm; }
void apply_gl_perm(const uint64_t* in, uint64_t* out, int limbs, int batch_count, int n, bool inverse, cudaStream_t s) {
    ok(mfhe_apply_gl_perm(in, out, limbs, batch_count, n, inverse ? 1 : 0, s), "apply_gl_perm");
}
void xy_ntt_forward_phantom(uint64_t* d, int limbs, int bc, int n, cudaStream_t s) { small_ntt(true, MFHE_NTT_PHANTOM_NEGACYCLIC, d, limbs, bc, n, s, "xy_ntt_forward_phantom"); }
void xy_ntt_backward_phantom(uint64_t* d, int limbs, int bc, int n, cudaStream_t s) { small_ntt(false, MFHE_NTT_PHANTOM_NEGACYCLIC, d, limbs, bc, n, s, "xy_ntt_backward_phantom"); }
void xy_ntt_forward_gl(uint64_t* d, uint64_t*, int limbs, int bc, int n, cudaStream_t s) { small_ntt(true, MFHE_NTT_GL_XN_MINUS_I, d, limbs, bc, n, s, "xy_ntt_forward_gl"); }
void xy_ntt_backward_gl(uint64_t* d, uint64_t*, int limbs, int bc, int n, cudaStream_t s) { small_ntt(false, MFHE_NTT_GL_XN_MINUS_I, d, limbs, bc, n, s, "xy_ntt_backward_gl"); }
void custom_ntt_forward(uint64_t* d, int limbs, int bc, int n, cudaStream_t s) { small_ntt(true, MFHE_NTT_CYCLIC_NATURAL, d, limbs, bc, n, s, "custom_ntt_forward"); }
void custom_ntt_backward(uint64_t* d, int limbs, int bc, int n, cudaStream_t s) { small_ntt(false, MFHE_NTT_CYCLIC_NATURAL, d, limbs, bc, n, s, "custom_ntt_backward"); }

// ---- encoder.cuh ---------------------------------------------------------------------------------------------
// The three kernels of the public header stay launchable with the reference's <<<grid, block>>> shapes:
// they are thin per-thread forms; the library's own path uses the batched kernels instead.
__global__ void mat_mul_kernel_complex(const cuDoubleComplex* A, const cuDoubleComplex* B, cuDoubleComplex* C, int n) {
    const int row = blockIdx.y * blockDim.y + threadIdx.y, col = blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= n || col >= n) return;
    double sr = 0.0, si = 0.0;
    for (int k = 0; k < n; ++k) {
        const cuDoubleComplex a = A[row * n + k], b = B[k * n + col];
        sr += a.x * b.x - a.y * b.y; si += a.x * b.y + a.y * b.x;
    }
    C[row * n + col] = make_cuDoubleComplex(sr, si);
}
namespace {
constexpr int kW = 7;                                    // 64-bit words of prod(RNS_MODULI): 385 bits
__device__ uint64_t d_sh_Q[kW], d_sh_Qhalf[kW], d_sh_M[RNS_NUM_LIMBS * kW], d_sh_inv[RNS_NUM_LIMBS], d_sh_q[RNS_NUM_LIMBS];
bool g_crt_up = false;
void upload_crt() {
    if (g_crt_up) return;
    uint64_t Q[kW], Qh[kW], M[RNS_NUM_LIMBS * kW], inv[RNS_NUM_LIMBS];
    ok(mfhe_ctx_get_crt(ctx(), 0, Q), "crt"); ok(mfhe_ctx_get_crt(ctx(), 1, Qh), "crt");
    ok(mfhe_ctx_get_crt(ctx(), 2, M), "crt"); ok(mfhe_ctx_get_crt(ctx(), 3, inv), "crt");
    cudaMemcpyToSymbol(d_sh_Q, Q, sizeof(Q)); cudaMemcpyToSymbol(d_sh_Qhalf, Qh, sizeof(Qh));
    cudaMemcpyToSymbol(d_sh_M, M, sizeof(M)); cudaMemcpyToSymbol(d_sh_inv, inv, sizeof(inv));
    cudaMemcpyToSymbol(d_sh_q, RNS_MODULI, sizeof(RNS_MODULI));
    g_crt_up = true;
}
// centred CRT lift of one coefficient: magnitude words + sign
__device__ bool lift(const uint64_t* data, int stride, int limbs, uint64_t (&mag)[kW]) {
    uint64_t acc[kW] = {0};
    for (int k = 0; k < limbs; ++k) {
        const uint64_t t = (uint64_t)(((unsigned __int128)data[(size_t)k * stride] * d_sh_inv[k]) % d_sh_q[k]);
        unsigned __int128 mc = 0, ac = 0;
        for (int i = 0; i < kW; ++i) {
            mc += (unsigned __int128)d_sh_M[k * kW + i] * t;
            ac += (unsigned __int128)acc[i] + (uint64_t)mc;
            acc[i] = (uint64_t)ac; ac >>= 64; mc >>= 64;
        }
        int c = 0;
        for (int i = kW - 1; i >= 0 && !c; --i) c = acc[i] > d_sh_Q[i] ? 1 : acc[i] < d_sh_Q[i] ? -1 : 0;
        if (c >= 0) { uint64_t br = 0; for (int i = 0; i < kW; ++i) { const uint64_t b = d_sh_Q[i] + br; br = acc[i] < b; acc[i] -= b; } }
    }
    int c = 0;
    for (int i = kW - 1; i >= 0 && !c; --i) c = acc[i] > d_sh_Qhalf[i] ? 1 : acc[i] < d_sh_Qhalf[i] ? -1 : 0;
    if (c > 0) { uint64_t br = 0; for (int i = 0; i < kW; ++i) { const uint64_t a = acc[i] + br; br = d_sh_Q[i] < a; mag[i] = d_sh_Q[i] - a; } return true; }
    for (int i = 0; i < kW; ++i) mag[i] = acc[i];
    return false;
}
}  // namespace
__global__ void crt_compose_centerlift_kernel(const uint64_t* d_in_rns, int64_t* d_out_centered, int n2, int limbs) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n2) return;
    uint64_t mag[kW];
    const bool neg = lift(d_in_rns + idx, n2, limbs, mag);
    const int64_t v = (int64_t)mag[0];                   // low word only, like the reference
    d_out_centered[idx] = neg ? -v : v;
}
__global__ void dequantize_exact_kernel(const uint64_t* d_real_rns, const uint64_t* d_imag_rns, cuDoubleComplex* d_out, int n2,
                                        double scaling_factor, int limbs) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n2) return;
    double part[2];
    for (int c = 0; c < 2; ++c) {
        uint64_t mag[kW];
        const bool neg = lift((c ? d_imag_rns : d_real_rns) + idx, n2, limbs, mag);
        double v = 0.0;
        for (int i = kW - 1; i >= 0; --i) v = v * 18446744073709551616.0 + (double)mag[i];
        v /= scaling_factor;
        part[c] = neg ? -v : v;
    }
    d_out[idx] = make_cuDoubleComplex(part[0], part[1]);
}
void crt_compose_centerlift_big(const uint64_t* in, uint64_t* mag, uint8_t* neg, int n2, int limbs, cudaStream_t s) {
    ok(mfhe_crt_compose_big(ctx(), in, mag, neg, n2, limbs, 1, s), "crt_compose_centerlift_big");
}

Encoder::Encoder(int n_dim) : n(n_dim), d_V_cx(nullptr), d_V_cx_T(nullptr), d_V_inv_cx(nullptr), d_V_inv_cx_T(nullptr) {
    if (n != MATRIX_N) { std::fprintf(stderr, "Error: Encoder: n must be %d\n", MATRIX_N); std::exit(1); }
    const size_t bytes = (size_t)n * n * sizeof(cuDoubleComplex);
    cudaMalloc(&d_V_cx, bytes); cudaMalloc(&d_V_cx_T, bytes); cudaMalloc(&d_V_inv_cx, bytes); cudaMalloc(&d_V_inv_cx_T, bytes);
    init_complex_matrices();
    upload_crt();
}
Encoder::~Encoder() { cudaFree(d_V_cx); cudaFree(d_V_cx_T); cudaFree(d_V_inv_cx); cudaFree(d_V_inv_cx_T); }
void Encoder::init_complex_matrices() {
    cuDoubleComplex* dst[4] = {d_V_cx, d_V_cx_T, d_V_inv_cx, d_V_inv_cx_T};
    for (int k = 0; k < 4; ++k) ok(mfhe_ctx_copy_xy_matrix(ctx(), k, (double*)dst[k], nullptr), "Encoder tables");
}
void Encoder::idft2(const cuDoubleComplex* in, cuDoubleComplex* out) { ok(mfhe_xy_idft2(ctx(), (const double*)in, (double*)out, 1, nullptr), "Encoder::idft2"); }
void Encoder::decode_from_eval_complex(const cuDoubleComplex* in, cuDoubleComplex* out) {
    ok(mfhe_xy_dft2(ctx(), (const double*)in, (double*)out, 1, nullptr), "Encoder::decode_from_eval_complex");
}
void Encoder::encode(const cuDoubleComplex* msg, uint64_t* re, uint64_t* im) {
    ok(mfhe_encode_lane(ctx(), (const double*)msg, re, im, nullptr), "Encoder::encode");
    cudaDeviceSynchronize();                       // the reference's encode ends with a device sync
}
void Encoder::decode_lane_from_rns_eval(const uint64_t* re, const uint64_t* im, cuDoubleComplex* msg) {
    ok(mfhe_decode_lane_from_rns_eval(ctx(), re, im, (double*)msg, nullptr), "Encoder::decode_lane_from_rns_eval");
    cudaDeviceSynchronize();
}

// ---- batched_encoder.cuh --------------------------------------------------------------------------------------
BatchedEncoder::BatchedEncoder(int n) : n_(n), n2_(n * n) {
    if (n != MATRIX_N) { std::fprintf(stderr, "Error: BatchedEncoder: n must be %d\n", MATRIX_N); std::exit(1); }
    init_tables_p17();
}
void BatchedEncoder::init_tables_p17() { (void)ctx(); }
void BatchedEncoder::encode_to_wntt_eval(const cuDoubleComplex* msg, uint64_t* out_re, uint64_t* out_im, cudaStream_t s) {
    ok(mfhe_encode(ctx(), (const double*)msg, out_re, out_im, s), "BatchedEncoder::encode_to_wntt_eval");
}
void BatchedEncoder::unpack_eval_p17(const uint64_t* in_re, const uint64_t* in_im, uint64_t* out_re, uint64_t* out_im, cudaStream_t s) {
    const size_t bytes = (size_t)BATCH_SIZE * RNS_NUM_LIMBS * n2_ * sizeof(uint64_t);
    cudaMemcpyAsync(out_re, in_re, bytes, cudaMemcpyDeviceToDevice, s);
    cudaMemcpyAsync(out_im, in_im, bytes, cudaMemcpyDeviceToDevice, s);
}

// ---- trace.cuh / batched_trace.cuh --------------------------------------------------------------------------------
void map_B_to_Bprime_Xinv_twist(const uint64_t* Br, const uint64_t* Bi, uint64_t* Bpr, uint64_t* Bpi, int n, int) {
    need_native(n, RNS_NUM_LIMBS, BATCH_SIZE, "map_B_to_Bprime_Xinv_twist");
    ok(mfhe_trace_map(ctx(), Br, Bi, Bpr, Bpi, 1, nullptr), "map_B_to_Bprime_Xinv_twist");
}
void trace_gemm_ABpT_rns(const uint64_t* Ar, const uint64_t* Ai, const uint64_t* Bpr, const uint64_t* Bpi, uint64_t* Cr, uint64_t* Ci, int n, int) {
    need_native(n, RNS_NUM_LIMBS, BATCH_SIZE, "trace_gemm_ABpT_rns");
    ok(mfhe_trace_gemm(ctx(), Ar, Ai, Bpr, Bpi, Cr, Ci, 1, nullptr), "trace_gemm_ABpT_rns");
}
void rescale_by_delta_rns(uint64_t* Cr, uint64_t* Ci, int n, int, uint64_t inv0, uint64_t inv1, uint64_t inv2) {
    need_native(n, RNS_NUM_LIMBS, BATCH_SIZE, "rescale_by_delta_rns");
    ok(mfhe_trace_rescale(ctx(), Cr, Ci, 1, inv0, inv1, inv2, nullptr), "rescale_by_delta_rns");
}
void map_B_to_Bprime_batched(const uint64_t* Br, const uint64_t* Bi, uint64_t* Bpr, uint64_t* Bpi, int n, int rns_limbs, int batch) {
    need_native(n, rns_limbs, BATCH_SIZE, "map_B_to_Bprime_batched");
    ok(mfhe_trace_map(ctx(), Br, Bi, Bpr, Bpi, batch, nullptr), "map_B_to_Bprime_batched");
}
void trace_gemm_batched(const uint64_t* Ar, const uint64_t* Ai, const uint64_t* Bpr, const uint64_t* Bpi, uint64_t* Cr, uint64_t* Ci, int n,
                        int rns_limbs, int batch) {
    need_native(n, rns_limbs, BATCH_SIZE, "trace_gemm_batched");
    ok(mfhe_trace_gemm(ctx(), Ar, Ai, Bpr, Bpi, Cr, Ci, batch, nullptr), "trace_gemm_batched");
}
void rescale_by_delta_batched(uint64_t* Cr, uint64_t* Ci, int n, int, int batch, uint64_t inv0, uint64_t inv1, uint64_t inv2) {
    need_native(n, RNS_NUM_LIMBS, BATCH_SIZE, "rescale_by_delta_batched");
    ok(mfhe_trace_rescale(ctx(), Cr, Ci, batch, inv0, inv1, inv2, nullptr), "rescale_by_delta_batched");
}

}  // namespace matrix_fhe
