// ctx.h -- internal definition of mfhe_ctx (the matrix pipeline's context) and the launchers
// the C-ABI layer (capi_scheme.cu) calls.  Everything here is device-pointer plumbing.
#pragma once
#include <cstddef>
#include <cstdint>
#include <mutex>
#include <vector>
#include <cuda_runtime.h>

#include "../../include/mfhe.h"
#include "modmath.cuh"

struct mfhe_ctx {
    int n = 0, logn = 0, f = 0, p = 0, phi = 0, L = 0, W = 0;     // W = 64-bit words of Q
    double delta = 0;
    std::vector<uint64_t> moduli, eta;
    std::vector<uint16_t> wexp;
    std::vector<uint64_t> hQ, hQhalf, hM, hInv;
    // device tables
    mfhe::u64* d_moduli = nullptr;            // [L]
    ulonglong2* d_mu = nullptr;               // [L] floor(2^128/q) {lo, hi}
    mfhe::u64* d_wtab_fwd = nullptr;          // [L][i = W-coeff r][o = lane w]   (transposed for the GEMM)
    mfhe::u64* d_wtab_inv = nullptr;          // [L][i = lane w][o = W-coeff r]
    int wsplit = 0;                           // 1: tables stored as {lo26, hi} pairs (all q < 2^52)
    unsigned char* d_wdig_fwd = nullptr;      // [L][8 digits][o][i] bytes: the same tables cut into 8-bit digits,
    unsigned char* d_wdig_inv = nullptr;      //   K-major, for the tensor-core W-CRT (wcrt_tc.cu)
    unsigned char* d_wdigt_fwd = nullptr;     // the same digits pre-tiled into UMMA core-matrix blocks
    unsigned char* d_wdigt_inv = nullptr;     //   [L][8][o/128][i/64][8 KiB] for the tcgen05 W-CRT (wcrt_umma.cu)
    std::vector<int> ndig;                    // [L] ceil(bits(q_l) / 8)
    double2* d_wdft_fwd = nullptr;            // [i = r][o = w]
    double2* d_wdft_inv = nullptr;            // [i = w][o = r]
    // conjugate-pair form of the same tables (fp64ops.cu): lanes (w, w') with exponents (e, p - e)
    int2* d_wpairs = nullptr;                 // [phi/2] {w, w'}
    double2* d_wdfth_fwd = nullptr;           // [i = r][pair]      = d_wdft_fwd[r][w]
    double2* d_wdfth_inv = nullptr;           // [pair][o = r]      = d_wdft_inv[w][r]
    int crt_fast = 0; mfhe::u64 crt_ginv = 0, crt_Q2lo = 0, crt_Q2hi = 0;     // see CrtDev
    int wdft_generic = 0;                     // 1: MFHE_WDFT_GENERIC was set at creation, use the full complex GEMM
    double2 *d_xyV = nullptr, *d_xyVT = nullptr, *d_xyVinv = nullptr, *d_xyVinvT = nullptr;   // [n][n]
    mfhe::u64 *d_crt_Q = nullptr, *d_crt_Qhalf = nullptr, *d_crt_M = nullptr, *d_crt_inv = nullptr;
    mfhe_ntt_plan* xy_plan = nullptr;
    // sampler option (mfhe_ctx_set_sampler): 0 = the reference's index hashes (parity mode, default), 1 = Philox4x32-10
    // keyed by `seed`; `draw` numbers the sampled arrays so that no two draws share randomness
    int sampler_mode = 0;
    uint64_t seed[2] = {0, 0};
    uint64_t draw = 0;
    char* arena = nullptr;
    size_t arena_bytes = 0, arena_used = 0;
    // the arena is one scratch area shared by all calls on this context: calls are serialised on the host by arena_mu
    // and ordered on the device by arena_ev (recorded when a call has enqueued its last arena user; a later call on a
    // DIFFERENT stream waits for it before touching the arena) -- see ArenaScope
    std::mutex arena_mu;
    cudaEvent_t arena_ev = nullptr;
    cudaStream_t arena_stream = nullptr;
    bool arena_busy = false;

    size_t n2() const { return (size_t)n * n; }
    size_t T() const { return (size_t)phi * L * n * n; }
    // bump allocator over the arena reserved at creation (reset at the start of every API call by ArenaScope)
    void* take(size_t bytes) {
        size_t off = (arena_used + 255) & ~(size_t)255;
        if (off + bytes > arena_bytes) return nullptr;
        arena_used = off + bytes;
        return arena + off;
    }
};

// One per API call that takes arena scratch.  Host threads calling into the same context are serialised for the
// duration of the enqueue; work enqueued on another stream than the previous arena user first waits (on the device,
// cudaStreamWaitEvent) until that user has finished, so two streams never share live scratch.
struct ArenaScope {
    mfhe_ctx* c; cudaStream_t st; std::unique_lock<std::mutex> lk;
    ArenaScope(mfhe_ctx* c_, cudaStream_t st_) : c(c_), st(st_), lk(c_->arena_mu) {
        c->arena_used = 0;
        if (c->arena_busy && c->arena_stream != st && c->arena_ev) cudaStreamWaitEvent(st, c->arena_ev, 0);
    }
    ~ArenaScope() {
        if (c->arena_ev && cudaEventRecord(c->arena_ev, st) == cudaSuccess) { c->arena_stream = st; c->arena_busy = true; }
    }
    ArenaScope(const ArenaScope&) = delete;
    ArenaScope& operator=(const ArenaScope&) = delete;
};

namespace mfhe {
// wtables_gpu.cu: V_l, V_l^-1 in all device layouts, built on the device (closed form); buffers must be allocated
cudaError_t wtables_build_device(mfhe_ctx* c, const u64* h_eta, const u64* h_ieta, const u64* h_pinv, const int* h_Phi,
                                 cudaStream_t st);

// ---- W-CRT as a per-limb modular GEMM  out[o][col] = sum_i tab[limb][i][o] * in[i][col] ----------
enum WcrtSource { kSrcMem = 0, kSrcI64 = 1, kSrcQuantRe = 2, kSrcQuantIm = 3, kSrcUniform = 4, kSrcGauss = 5 };
struct WcrtArgs {
    const void* in = nullptr;      // u64 / int64 / double2 depending on the source
    u64* out = nullptr;
    u64* out2 = nullptr;           // optional duplicate of the output (encrypt_pair stores a twice)
    const u64* tab = nullptr;      // [limbs][phi][phi]
    const u64* moduli = nullptr;
    const ulonglong2* mu = nullptr;
    int n = 0, cols = 0;           // cols = n*n (matrix form) or n (vector form); rows of n along x
    int limbs = 0, phi = 0;
    int limb0 = 0;                 // first limb of `tab`/`moduli` used (kSrcI64 inverse: limb 0 only)
    int in_layout = 0, out_layout = 0;
    int source = kSrcMem;
    int split = 0;
    double delta = 0;
    const unsigned char* tabd = nullptr;   // digit tables [limbs][8][phi][phi] (tensor-core path), may be null
    const unsigned char* tabt = nullptr;   // digit tables pre-tiled for UMMA (tcgen05 path), may be null
    const int* h_ndig = nullptr;           // HOST array [limbs]: 8-bit digits per modulus
    int lz0 = 0, lzn = 0;                  // relative limb window of one launch (set by wcrt_launch)
    // optional [phi][limbs][cols] words: on the tcgen05 path a computing source (int64, quantiser, samplers) is first
    // written out as canonical residues here and the GEMM then runs on the plain-memory loader (the fused loaders
    // cannot feed 128 x 48 tiles fast enough; the extra pass is 2 x S of HBM traffic, ~0.06 ms at L = 11)
    u64* scratch = nullptr;
};
cudaError_t wcrt_launch(const WcrtArgs& a, cudaStream_t st);
// kSrcGauss input: int8 noise value of every logical coefficient [phi][cols]
cudaError_t gauss_table_launch(signed char* out, int phi, int cols, cudaStream_t st);
cudaError_t llround_noise_planes_launch(const double2* in, long long* re, long long* im, size_t total, double delta,
                                        cudaStream_t st);

// ---- fp64 complex kernels (fp64ops.cu) --------------------------------------------------------------
// out[o][col] = sum_i tab[i][o] * in[i][col], phi x phi table, cols columns
// mode 1: conjugate pairs along the outputs (forward table), 2: along the inputs (inverse), 0: full table only
struct WdftTab { const double2* full; const double2* half; const int2* pairs; int mode; };
cudaError_t wdft_launch(const WdftTab& tab, const double2* in, double2* out, int phi, int cols, cudaStream_t st);
// same on split planes: in = (re, im) of type double (pair = 0) or int64 (pair = 1)
cudaError_t wdft_pair_launch(const WdftTab& tab, const void* in_re, const void* in_im, double* out_re, double* out_im,
                             int phi, int cols, int int_input, cudaStream_t st);
// (re, im)[i] = llround(in[i] * delta) as int64 planes (total even or odd, 16-byte aligned buffers)
cudaError_t llround_planes_launch(const double2* in, long long* re, long long* im, size_t total, double delta, cudaStream_t st);
// out = A * in * B per lane (A, B n x n), lanes matrices
cudaError_t xy_sandwich_launch(const double2* A, const double2* B, const double2* in, double2* out, int n, int lanes,
                               cudaStream_t st);
cudaError_t matmul_complex_launch(const double2* A, const double2* B, double2* C, int n, cudaStream_t st);
cudaError_t quantize_launch(const double2* in, u64* re, u64* im, const u64* moduli, const ulonglong2* mu, int n2,
                            int limbs, int lanes, double delta, cudaStream_t st);

// ---- CRT compose (crt.cu) ----------------------------------------------------------------------------
struct CrtDev {
    const u64 *Q, *Qhalf, *M, *inv, *moduli; const ulonglong2* mu; int W;
    // two-limb fast path of crt_to_complex (crt.cu: crt_small): q0^-1 mod q1, q0 q1 and floor(q0 q1 / 2)
    int fast, L; u64 ginv, Q2lo, Q2hi, H2lo, H2hi;     // fast path only when all L limbs are composed
};
cudaError_t crt_compose_big_launch(const CrtDev& c, const u64* in, u64* mag, uint8_t* neg, int n2, int limbs, int lanes,
                                   int lane_stride_limbs, cudaStream_t st);
cudaError_t crt_compose_i64_launch(const CrtDev& c, const u64* in, long long* out, int n2, int limbs, int lanes,
                                   int saturate, cudaStream_t st);
cudaError_t center_i64_launch(const u64* in, long long* out, u64 q, size_t total, cudaStream_t st);
// in_re / in_im [lane][limbs][n2] -> complex out [lane][n2] = centred value / delta
cudaError_t crt_to_complex_launch(const CrtDev& c, const u64* in_re, const u64* in_im, double2* out, int n2, int limbs,
                                  int lanes, double delta, cudaStream_t st);
cudaError_t big_pair_to_complex_launch(const u64* re_mag, const uint8_t* re_neg, const u64* im_mag, const uint8_t* im_neg,
                                       double2* out, size_t total, int W, double delta, cudaStream_t st);

// ---- scheme kernels (scheme.cu) ------------------------------------------------------------------------
struct RowArgs {                 // one launch over all phi*L*n rows of n coefficients
    const u64 *a = nullptr, *s = nullptr;          // a rows (layout a_layout), key [w][l][x]
    const u64 *a1 = nullptr;                       // optional: the second component's own a (compared with a row by row)
    const u64 *m0 = nullptr, *e0 = nullptr; u64* o0 = nullptr;   // out0 = m0 -/+ t + e0
    const u64 *m1 = nullptr, *e1 = nullptr; u64* o1 = nullptr;   // optional second component
    int a_layout = 0, m_layout = 0, e_layout = 0, o_layout = 0;
    int subtract = 1;            // 1: out = m - t + e (encrypt), 0: out = m + t (decrypt; e ignored)
    int n = 0, limbs = 0, phi = 0;
};
struct SmallTables;
cudaError_t xrow_mul_s_launch(const RowArgs& r, const SmallTables& T, const ulonglong2* mu, cudaStream_t st);
// fused decryption of the 8 poly sets of mfhe_ct_matmul (n = 64): out = evals [w][l][z][x]
cudaError_t matmul_decrypt_launch(const u64* D, size_t T, const u64* sk, u64* out_re, u64* out_im, const SmallTables& Tb,
                                  const ulonglong2* mu, int phi, int limbs, cudaStream_t st);
cudaError_t sample_launch(int kind, u64* out, const u64* moduli, const ulonglong2* mu, int n, int limbs, int phi, cudaStream_t st);
// sampler_sound.cu: kind 0 uniform [w][l][y][x], 1 gaussian (same), 2 ternary [w][l][x]
cudaError_t sound_sample_launch(int kind, u64* out, const u64* moduli, const ulonglong2* mu, int n, int limbs, int phi,
                                u64 seed0, u64 seed1, u64 draw, cudaStream_t st);
void sound_cdt_copy(u64* h_out);            // the 83-entry cumulative table (tests)
cudaError_t layout_launch(const u64* in, u64* out, int n, int limbs, int phi, int in_layout, int out_layout, cudaStream_t st);
cudaError_t ct_add_launch(const u64* c1, const u64* c2, u64* res, const u64* moduli, size_t T, int n, int limbs,
                          int ref_index, cudaStream_t st);
cudaError_t ct_mul_launch(const u64* c1, const u64* c2, u64* d0, u64* d1, u64* d2, const u64* moduli, const ulonglong2* mu,
                          size_t T, int n, int limbs, int ref_index, cudaStream_t st);
cudaError_t trace_map_launch(const u64* Br, const u64* Bi, u64* Bpr, u64* Bpi, const u64* moduli, int n, int limbs,
                             int batch, cudaStream_t st, int axis = 0);
cudaError_t transpose_launch(const u64* in, u64* out, int n, size_t count, cudaStream_t st);
struct RescaleConsts { u64 inv[64]; };
cudaError_t rescale_combine_launch(const u64* in, u64* out, const u64* moduli, const ulonglong2* mu, int cols, int limbs, int phi,
                                   const RescaleConsts& k, cudaStream_t st);
cudaError_t trace_rescale_limbs_launch(u64* Cr, u64* Ci, const u64* moduli, const ulonglong2* mu, int n, int limbs, int batch,
                                       const RescaleConsts& k, cudaStream_t st);
// the four GEMMs of mfhe_ct_matmul in one tcgen05 kernel (ct_matmul_umma.cu)
bool ct_matmul_umma_eligible(int n, int limbs, const int* h_ndig);
cudaError_t ct_matmul_umma_launch(const u64* Ure, const u64* Uim, const u64* Vre, const u64* Vim, u64* out, size_t T,
                                  const u64* moduli, const ulonglong2* mu, const int* h_ndig, int limbs, int phi, cudaStream_t st);
bool trace_gemm_tc_eligible(int n, int limbs, const int* h_ndig);
cudaError_t trace_gemm_tc_launch(const u64* Ar, const u64* Ai, const u64* Bpr, const u64* Bpi, u64* Cr, u64* Ci,
                                 const u64* moduli, const ulonglong2* mu, const int* h_ndig, int limbs, int batch,
                                 cudaStream_t st, int transposed = 0);
cudaError_t trace_gemm_launch(const u64* Ar, const u64* Ai, const u64* Bpr, const u64* Bpi, u64* Cr, u64* Ci,
                              const u64* moduli, const ulonglong2* mu, int n, int limbs, int batch, cudaStream_t st,
                              int transposed = 0);
cudaError_t trace_rescale_launch(u64* Cr, u64* Ci, const u64* moduli, const ulonglong2* mu, int n, int limbs, int batch,
                                 u64 inv0, u64 inv1, u64 inv2, cudaStream_t st);
cudaError_t gl_perm_launch(const u64* in, u64* out, const uint32_t* perm, int limbs, int batch, int n, cudaStream_t st);

}  // namespace mfhe
