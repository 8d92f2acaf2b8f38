// wtables_gpu.cu -- W-CRT tables built ON THE DEVICE.
//
// The reference builds V_l[w][r] = eta_l^(exp[w] r) and V_l^-1 on the host, the inverse by O(phi^3) Gauss-Jordan with a
// 128-bit division per step (src/core/HE.cu:135-185, 237-273): 12.5 s of init_he_backend on the bench box
// (profiles/r2_01_ref_time_stock.json).  Here one CTA per (limb, lane w) forms
//   row w of V          : the 512 powers of eta^exp[w]
//   column w of V^-1    : p^-1 * sum_j eta^(-exp[w] j) W^j  reduced modulo Phi_p(W)   (inverse DFT of the indicator of the
//                         unit exp[w], restricted to the units; Phi_{3f} has coefficients in {-1, 0, 1})
// -- the same closed form wtables_host.hpp evaluates (kept as the cross-check path, MFHE_TABLES_HOST=1) -- and writes
// the four layouts the kernels consume straight from shared memory: the split-26 GEMM tables, the 8-bit digit planes
// (K-major) and the digit planes pre-tiled into UMMA core matrices.  The matrix is unique, so every entry equals the
// reference's Gauss-Jordan result mod q (tests/test_scheme_native_gpu.py::test_native_tables against the oracle, which is
// pinned to the literal Gauss-Jordan in tests/test_oracle_scheme.py).
#include "ctx.h"

namespace mfhe {

namespace {
__device__ __forceinline__ u64 mulm(u64 a, u64 b, u64 q, ulonglong2 mu) { return mul_barrett(a, b, q, mu.x, mu.y); }
__device__ u64 powm(u64 b, u32 e, u64 q, ulonglong2 mu) {
    u64 r = 1;
    while (e) { if (e & 1) r = mulm(r, b, q, mu); b = mulm(b, b, q, mu); e >>= 1; }
    return r;
}

struct WtabArgs {
    const u64* moduli; const ulonglong2* mu;
    const u64* eta; const u64* ieta; const u64* pinv;       // [L] device
    const unsigned short* wexp;                             // [phi]
    const signed char* Phi;                                 // [phi + 1] coefficients of Phi_p, degree phi
    u64 *tab_fwd, *tab_inv;                                 // [L][i][o]
    unsigned char *dig_fwd, *dig_inv, *digt_fwd, *digt_inv; // may be null
    int phi, p, split;
};

__global__ void __launch_bounds__(256)
wtab_kernel(WtabArgs a) {
    extern __shared__ __align__(16) u64 sm_wt[];
    u64* g = sm_wt;                       // [p]
    u64* V = sm_wt + a.p;                 // [phi]
    const int w = blockIdx.x, l = blockIdx.y, phi = a.phi, p = a.p;
    const u64 q = a.moduli[l]; const ulonglong2 mu = a.mu[l];
    const u32 e = a.wexp[w];
    const u64 root = powm(a.eta[l], e, q, mu), iroot = powm(a.ieta[l], e, q, mu), pinv = a.pinv[l];
    for (int r = threadIdx.x; r < phi; r += blockDim.x) V[r] = powm(root, (u32)r, q, mu);
    for (int j = threadIdx.x; j < p; j += blockDim.x) g[j] = mulm(pinv, powm(iroot, (u32)j, q, mu), q, mu);
    __syncthreads();
    // reduce modulo Phi_p: Phi is monic of degree phi, so W^d = -sum_{i<phi} Phi[i] W^(d-phi+i)
    for (int d = p - 1; d >= phi; --d) {
        const u64 co = g[d];
        for (int i = threadIdx.x; i < phi; i += blockDim.x) {
            const int c = a.Phi[i];
            if (c) { u64& t = g[d - phi + i]; t = c > 0 ? sub_mod(t, co, q) : add_mod(t, co, q); }
        }
        __syncthreads();
    }
    const size_t pp = (size_t)phi * phi;
    auto pack = [&](u64 v) { return a.split ? ((v & ((1ull << 26) - 1)) | ((v >> 26) << 32)) : v; };
    for (int r = threadIdx.x; r < phi; r += blockDim.x) {
        const u64 vf = V[r], vi = g[r];              // V[w][r], V^-1[r][w]
        a.tab_fwd[(size_t)l * pp + (size_t)r * phi + w] = pack(vf);        // [i = r][o = w]
        a.tab_inv[(size_t)l * pp + (size_t)w * phi + r] = pack(vi);        // [i = w][o = r]
        if (a.dig_fwd) {
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                a.dig_fwd[((size_t)l * 8 + u) * pp + (size_t)w * phi + r] = (unsigned char)(vf >> (8 * u));   // V[w][r]
                a.dig_inv[((size_t)l * 8 + u) * pp + (size_t)r * phi + w] = (unsigned char)(vi >> (8 * u));   // Vinv[r][w]
            }
        }
        if (a.digt_fwd) {
            const size_t kch = (size_t)phi / 64;
            auto tidx = [&](size_t o, size_t i) {
                const size_t rr = o % 128, kk = i % 64;
                return ((o / 128) * kch + i / 64) * 8192 + (rr / 8) * 512 + (kk / 16) * 128 + (rr % 8) * 16 + kk % 16;
            };
            const size_t xf = tidx((size_t)w, (size_t)r), xi = tidx((size_t)r, (size_t)w);
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                a.digt_fwd[((size_t)l * 8 + u) * pp + xf] = (unsigned char)(vf >> (8 * u));
                a.digt_inv[((size_t)l * 8 + u) * pp + xi] = (unsigned char)(vi >> (8 * u));
            }
        }
    }
}
}  // namespace

// device buffers of the context must be allocated; h_* are tiny host arrays
cudaError_t wtables_build_device(mfhe_ctx* c, const u64* h_eta, const u64* h_ieta, const u64* h_pinv, const int* h_Phi,
                                 cudaStream_t st) {
    const int L = c->L, phi = c->phi;
    u64* d_small = nullptr; unsigned short* d_exp = nullptr; signed char* d_Phi = nullptr;
    std::vector<u64> small((size_t)3 * L);
    for (int l = 0; l < L; ++l) { small[l] = h_eta[l]; small[L + l] = h_ieta[l]; small[2 * L + l] = h_pinv[l]; }
    std::vector<signed char> Phi8((size_t)phi + 1);
    for (int i = 0; i <= phi; ++i) Phi8[i] = (signed char)h_Phi[i];
    cudaError_t e = cudaMalloc((void**)&d_small, small.size() * 8);
    if (e == cudaSuccess) e = cudaMalloc((void**)&d_exp, (size_t)phi * 2);
    if (e == cudaSuccess) e = cudaMalloc((void**)&d_Phi, Phi8.size());
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_small, small.data(), small.size() * 8, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_exp, c->wexp.data(), (size_t)phi * 2, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_Phi, Phi8.data(), Phi8.size(), cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) {
        WtabArgs a{c->d_moduli, c->d_mu, d_small, d_small + L, d_small + 2 * L, d_exp, d_Phi, c->d_wtab_fwd, c->d_wtab_inv,
                   c->d_wdig_fwd, c->d_wdig_inv, c->d_wdigt_fwd, c->d_wdigt_inv, phi, c->p, c->wsplit};
        const size_t smem = (size_t)(c->p + phi) * sizeof(u64);
        wtab_kernel<<<dim3((unsigned)phi, (unsigned)L), 256, smem, st>>>(a);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);      // the small staging arrays die with this call
    cudaFree(d_small); cudaFree(d_exp); cudaFree(d_Phi);
    return e;
}

}  // namespace mfhe
