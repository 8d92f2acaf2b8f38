// capi_scheme.cu -- C ABI (include/mfhe.h) of the matrix pipeline: context, W-CRT, W-DFT, XY
// embedding, CRT, encode / decode, samplers, keygen, encrypt / decrypt, ct ops, trace, GL perm.
// Host-side orchestration only; the kernels live in wcrt.cu, fp64ops.cu, crt.cu, scheme.cu.
#include <cstdlib>
#include <algorithm>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "capi_common.h"
#include "ctx.h"
#include "plan.h"
#include "wtables_host.hpp"

using namespace mfhe;

namespace {

template <class Tp>
cudaError_t upload(Tp*& dst, const void* src, size_t bytes) {
    cudaError_t e = cudaMalloc((void**)&dst, bytes);
    if (e != cudaSuccess) return e;
    return cudaMemcpy(dst, src, bytes, cudaMemcpyHostToDevice);
}

inline cudaStream_t S(void* s) { return (cudaStream_t)s; }
inline int done(cudaError_t e, const char* where) { return e == cudaSuccess ? MFHE_OK : cuda_fail(e, where); }

CrtDev crt_dev(const mfhe_ctx* c) {
    CrtDev d{c->d_crt_Q, c->d_crt_Qhalf, c->d_crt_M, c->d_crt_inv, c->d_moduli, c->d_mu, c->W, c->crt_fast, c->L, c->crt_ginv,
             c->crt_Q2lo, c->crt_Q2hi, 0, 0};
    d.H2lo = (c->crt_Q2lo >> 1) | (c->crt_Q2hi << 63); d.H2hi = c->crt_Q2hi >> 1;
    return d;
}

WcrtArgs wcrt_args(const mfhe_ctx* c, bool inverse) {
    WcrtArgs a;
    a.tab = inverse ? c->d_wtab_inv : c->d_wtab_fwd;
    a.moduli = c->d_moduli; a.mu = c->d_mu;
    a.n = c->n; a.cols = c->n * c->n; a.limbs = c->L; a.phi = c->phi;
    a.split = c->wsplit; a.delta = c->delta;
    a.tabd = inverse ? c->d_wdig_inv : c->d_wdig_fwd; a.h_ndig = c->ndig.data();
    a.tabt = inverse ? c->d_wdigt_inv : c->d_wdigt_fwd;
    return a;
}
WdftTab wdft_tab(const mfhe_ctx* c, bool inverse) {
    WdftTab w;
    w.full = inverse ? c->d_wdft_inv : c->d_wdft_fwd;
    w.half = inverse ? c->d_wdfth_inv : c->d_wdfth_fwd;
    w.pairs = c->d_wpairs;
    w.mode = c->wdft_generic ? 0 : (inverse ? 2 : 1);
    return w;
}

}  // namespace

extern "C" void mfhe_ctx_destroy(mfhe_ctx* c) {
    if (!c) return;
    cudaFree(c->d_moduli); cudaFree(c->d_mu); cudaFree(c->d_wtab_fwd); cudaFree(c->d_wtab_inv); cudaFree(c->d_wdig_fwd); cudaFree(c->d_wdig_inv); cudaFree(c->d_wdigt_fwd); cudaFree(c->d_wdigt_inv);
    cudaFree(c->d_wdft_fwd); cudaFree(c->d_wdft_inv); cudaFree(c->d_wpairs); cudaFree(c->d_wdfth_fwd); cudaFree(c->d_wdfth_inv);
    cudaFree(c->d_xyV); cudaFree(c->d_xyVT); cudaFree(c->d_xyVinv); cudaFree(c->d_xyVinvT);
    cudaFree(c->d_crt_Q); cudaFree(c->d_crt_Qhalf); cudaFree(c->d_crt_M); cudaFree(c->d_crt_inv);
    cudaFree(c->arena);
    if (c->arena_ev) cudaEventDestroy(c->arena_ev);
    if (c->xy_plan) mfhe_ntt_plan_destroy(c->xy_plan);
    delete c;
}

extern "C" int mfhe_ctx_create(const mfhe_params* p, mfhe_ctx** out) {
    if (!out) return fail(MFHE_EINVAL, "ctx_create: out is NULL");
    *out = nullptr;
    if (!p || !p->moduli) return fail(MFHE_EINVAL, "ctx_create: no parameters");
    const int n = p->n, f = p->f, L = p->num_limbs;
    if (n < 16 || n > 64 || (n & (n - 1))) return fail(MFHE_EUNSUPPORTED, "ctx_create: n must be 16, 32 or 64");
    if (f < 3 || f > 683 || f % 2 == 0) return fail(MFHE_EINVAL, "ctx_create: f must be an odd prime <= 683");
    for (int d = 3; d * d <= f; d += 2) if (f % d == 0) return fail(MFHE_EINVAL, "ctx_create: f is not prime");
    if (f == 3) return fail(MFHE_EINVAL, "ctx_create: f = 3 gives p = 9 (not square-free)");
    if (L < 1 || L > 64) return fail(MFHE_EINVAL, "ctx_create: num_limbs out of range");
    if (!(p->delta > 0)) return fail(MFHE_EINVAL, "ctx_create: delta must be positive");
    mfhe_ctx* c = new (std::nothrow) mfhe_ctx;
    if (!c) return fail(MFHE_ENOMEM, "ctx_create: host allocation failed");
    c->n = n; c->f = f; c->p = 3 * f; c->phi = 2 * (f - 1); c->L = L; c->delta = p->delta;
    while ((1 << c->logn) < n) ++c->logn;
    c->moduli.assign(p->moduli, p->moduli + L);
    const size_t phi = (size_t)c->phi, pp = phi * phi;
    // W-CRT tables: built on the device (wtables_gpu.cu) unless MFHE_TABLES_HOST is set (the host closed form below,
    // kept as the cross-check path: tests compare the two word for word)
    const bool host_tables = std::getenv("MFHE_TABLES_HOST") != nullptr;
    std::vector<uint64_t> tab_fwd(host_tables ? (size_t)L * pp : 0), tab_inv(tab_fwd.size());
    std::vector<host::TwPair> mu(L);
    const bool want_digits = (phi % 64 == 0);               // the tensor-core W-CRT tiling (wcrt_tc.cu)
    std::vector<unsigned char> dig_fwd(want_digits && host_tables ? (size_t)L * 8 * pp : 0), dig_inv(dig_fwd.size());
    const bool want_tiles = (phi % 128 == 0);                // tcgen05 tiling (wcrt_umma.cu)
    std::vector<unsigned char> digt_fwd(want_tiles && host_tables ? (size_t)L * 8 * pp : 0), digt_inv(digt_fwd.size());
    std::vector<uint64_t> h_ieta(L), h_pinv(L);
    std::vector<int> Phi;
    c->ndig.assign(L, 8);
    std::vector<host::cplx> wV, wVinv, xV, xVinv;
    host::CrtTables crt;
    try {
        bool small = true;
        for (int l = 0; l < L; ++l) {
            const uint64_t q = c->moduli[l];
            if (q >= (1ull << 59) || q < 3) throw std::runtime_error("modulus must be below 2^59");
            if (!host::is_prime(q)) throw std::runtime_error("modulus " + std::to_string(q) + " is not prime");
            if ((q - 1) % (uint64_t)c->p || (q - 1) % (uint64_t)(2 * n))
                throw std::runtime_error("modulus " + std::to_string(q) + " is not 1 mod lcm(2n, p)");
            if (q >= (1ull << 52)) small = false;
            for (int k = 0; k < l; ++k) if (c->moduli[k] == q) throw std::runtime_error("moduli must be distinct");
        }
        c->wsplit = small ? 1 : 0;
        c->wexp = host::unit_exponents(f);
        Phi = host::cyclotomic_3f(f);
        for (int l = 0; l < L; ++l) {
            const uint64_t q = c->moduli[l];
            const uint64_t eta = host::find_eta(q, f);
            if (!eta) throw std::runtime_error("no element of order p mod " + std::to_string(q));
            c->eta.push_back(eta);
            h_ieta[l] = host::invmod(eta, q); h_pinv[l] = host::invmod((uint64_t)c->p % q, q);
            { int bits = 0; while (bits < 64 && (q >> bits)) ++bits; c->ndig[l] = (bits + 7) / 8; }
            const unsigned __int128 m128 = ~(unsigned __int128)0 / q;
            mu[l] = host::TwPair{(uint64_t)m128, (uint64_t)(m128 >> 64)};
            if (!host_tables) continue;
            std::vector<uint64_t> V, Vinv;
            host::build_wcrt_limb(q, f, c->wexp, Phi, eta, V, Vinv);
            // GEMM tables [i][o]: forward i = r (W-coeff), o = w  -> V[w][r];  inverse i = w, o = r -> Vinv[r][w]
            auto pack = [&](uint64_t v) { return small ? ((v & ((1ull << 26) - 1)) | ((v >> 26) << 32)) : v; };
            for (size_t i = 0; i < phi; ++i)
                for (size_t o = 0; o < phi; ++o) {
                    tab_fwd[(size_t)l * pp + i * phi + o] = pack(V[o * phi + i]);
                    tab_inv[(size_t)l * pp + i * phi + o] = pack(Vinv[o * phi + i]);
                }
            if (want_digits)
                for (int u = 0; u < 8; ++u)
                    for (size_t k = 0; k < pp; ++k) {
                        dig_fwd[((size_t)l * 8 + u) * pp + k] = (unsigned char)(V[k] >> (8 * u));
                        dig_inv[((size_t)l * 8 + u) * pp + k] = (unsigned char)(Vinv[k] >> (8 * u));
                    }
            if (want_tiles) {
                const size_t kch = phi / 64;
                for (size_t o = 0; o < phi; ++o)
                    for (size_t i = 0; i < phi; ++i) {
                        const size_t r = o % 128, kk = i % 64;
                        const size_t idx = ((o / 128) * kch + i / 64) * 8192 + (r / 8) * 512 + (kk / 16) * 128 + (r % 8) * 16 + kk % 16;
                        for (int u = 0; u < 8; ++u) {
                            digt_fwd[((size_t)l * 8 + u) * pp + idx] = (unsigned char)(V[o * phi + i] >> (8 * u));
                            digt_inv[((size_t)l * 8 + u) * pp + idx] = (unsigned char)(Vinv[o * phi + i] >> (8 * u));
                        }
                    }
            }
        }
        host::build_wdft(f, c->wexp, Phi, wV, wVinv);
        host::build_xy(n, xV, xVinv);
        crt = host::build_crt(c->moduli);
        if (crt.words > 32) throw std::runtime_error("product of the moduli exceeds 2048 bits");
    } catch (const std::exception& e) {
        delete c;
        return fail(MFHE_EINVAL, std::string("ctx_create: ") + e.what());
    }
    if (L >= 2 && c->moduli[0] != c->moduli[1] && !std::getenv("MFHE_CRT_FULL")) {
        const unsigned __int128 q01 = (unsigned __int128)c->moduli[0] * c->moduli[1];
        if (!(q01 >> 126)) {
            c->crt_fast = 1;
            c->crt_ginv = host::invmod(c->moduli[0] % c->moduli[1], c->moduli[1]);
            c->crt_Q2lo = (uint64_t)q01; c->crt_Q2hi = (uint64_t)(q01 >> 64);
        }
    }
    c->W = crt.words; c->hQ = crt.Q; c->hQhalf = crt.Qhalf; c->hM = crt.M; c->hInv = crt.inv;

    // transposed complex tables for the GEMM form  out[o] = sum_i tab[i][o] in[i]
    std::vector<host::cplx> wf(pp), wi(pp);
    for (size_t i = 0; i < phi; ++i)
        for (size_t o = 0; o < phi; ++o) { wf[i * phi + o] = wV[o * phi + i]; wi[i * phi + o] = wVinv[o * phi + i]; }
    // conjugate pairs of lanes: exponents e and p - e (both units); representative = the smaller lane index
    std::vector<int2> wpairs;
    {
        std::vector<int> at(c->p, -1);
        for (size_t w = 0; w < phi; ++w) at[c->wexp[w]] = (int)w;
        std::vector<char> seen(phi, 0);
        for (size_t w = 0; w < phi; ++w) {
            if (seen[w]) continue;
            const int w2 = at[c->p - c->wexp[w]];
            if (w2 < 0 || (size_t)w2 == w || seen[w2]) { wpairs.clear(); break; }
            seen[w] = seen[w2] = 1;
            wpairs.push_back(make_int2((int)w, w2));
        }
        if (wpairs.size() * 2 != phi) wpairs.clear();
    }
    const size_t hp = wpairs.size();
    std::vector<host::cplx> whf(phi * hp), whi(phi * hp);
    for (size_t k = 0; k < hp; ++k)
        for (size_t r = 0; r < phi; ++r) {
            whf[r * hp + k] = wf[r * phi + wpairs[k].x];          // [i = r][pair]
            whi[k * phi + r] = wi[(size_t)wpairs[k].x * phi + r];  // [pair][o = r]
        }
    c->wdft_generic = (hp == 0 || std::getenv("MFHE_WDFT_GENERIC")) ? 1 : 0;
    const size_t nn = (size_t)n * n;
    std::vector<host::cplx> xVT(nn), xVinvT(nn);
    for (int r = 0; r < n; ++r)
        for (int k = 0; k < n; ++k) { xVT[(size_t)k * n + r] = xV[(size_t)r * n + k]; xVinvT[(size_t)k * n + r] = xVinv[(size_t)r * n + k]; }

    cudaError_t e = cudaSuccess;
#define UP(dst, vec) if (e == cudaSuccess) e = upload(dst, (vec).data(), (vec).size() * sizeof((vec)[0]))
    UP(c->d_moduli, c->moduli); UP(c->d_mu, mu);
    if (host_tables) {
        UP(c->d_wtab_fwd, tab_fwd); UP(c->d_wtab_inv, tab_inv);
        if (want_digits) { UP(c->d_wdig_fwd, dig_fwd); UP(c->d_wdig_inv, dig_inv); }
        if (want_tiles) { UP(c->d_wdigt_fwd, digt_fwd); UP(c->d_wdigt_inv, digt_inv); }
    } else {
#define AL(dst, bytes) if (e == cudaSuccess) e = cudaMalloc((void**)&(dst), (bytes))
        AL(c->d_wtab_fwd, (size_t)L * pp * 8); AL(c->d_wtab_inv, (size_t)L * pp * 8);
        if (want_digits) { AL(c->d_wdig_fwd, (size_t)L * 8 * pp); AL(c->d_wdig_inv, (size_t)L * 8 * pp); }
        if (want_tiles) { AL(c->d_wdigt_fwd, (size_t)L * 8 * pp); AL(c->d_wdigt_inv, (size_t)L * 8 * pp); }
#undef AL
        if (e == cudaSuccess) e = wtables_build_device(c, (const u64*)c->eta.data(), (const u64*)h_ieta.data(), (const u64*)h_pinv.data(), Phi.data(), nullptr);
    }
    UP(c->d_wdft_fwd, wf); UP(c->d_wdft_inv, wi);
    if (hp) { UP(c->d_wpairs, wpairs); UP(c->d_wdfth_fwd, whf); UP(c->d_wdfth_inv, whi); }
    UP(c->d_xyV, xV); UP(c->d_xyVT, xVT); UP(c->d_xyVinv, xVinv); UP(c->d_xyVinvT, xVinvT);
    UP(c->d_crt_Q, crt.Q); UP(c->d_crt_Qhalf, crt.Qhalf); UP(c->d_crt_M, crt.M); UP(c->d_crt_inv, crt.inv);
#undef UP
    if (e == cudaSuccess) {
        // scratch: 6 poly sets + 4 complex lane sets (largest consumers: decrypt_decode with 4 + 4, and encrypt_pair under
        // the sound sampler with 3 drawn arrays + 2 transformed errors + 1 materialised source)
        const size_t need = 6 * c->T() * 8 + 4 * phi * nn * 16 + (1 << 20);
        c->arena_bytes = std::max(need, p->arena_bytes);
        e = cudaMalloc((void**)&c->arena, c->arena_bytes);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->arena_ev, cudaEventDisableTiming);
    }
    if (e != cudaSuccess) { mfhe_ctx_destroy(c); return cuda_fail(e, "ctx_create"); }
    int rc = mfhe_ntt_plan_create(c->logn, L, c->moduli.data(), MFHE_NTT_PHANTOM_NEGACYCLIC, &c->xy_plan);
    if (rc != MFHE_OK) { mfhe_ctx_destroy(c); return rc; }
    *out = c;
    return MFHE_OK;
}

extern "C" int mfhe_ctx_info(const mfhe_ctx* c, int out[6]) {
    if (!c || !out) return fail(MFHE_EINVAL, "ctx_info: NULL argument");
    out[0] = c->n; out[1] = c->f; out[2] = c->p; out[3] = c->phi; out[4] = c->L; out[5] = c->W;
    return MFHE_OK;
}
extern "C" int mfhe_ctx_get_moduli(const mfhe_ctx* c, uint64_t* h, int count) {
    if (!c || !h || count < 0 || count > c->L) return fail(MFHE_EINVAL, "get_moduli: bad argument");
    // read back from the device, as copy_device_moduli does (HE.cu:410-422)
    return done(cudaMemcpy(h, c->d_moduli, (size_t)count * 8, cudaMemcpyDeviceToHost), "get_moduli");
}
extern "C" uint64_t mfhe_ctx_eta(const mfhe_ctx* c, int l) { return (c && l >= 0 && l < c->L) ? c->eta[l] : 0; }
extern "C" int mfhe_ctx_get_wcrt_table(const mfhe_ctx* c, int l, int inverse, uint64_t* h) {
    if (!c || !h || l < 0 || l >= c->L) return fail(MFHE_EINVAL, "get_wcrt_table: bad argument");
    const size_t phi = (size_t)c->phi, pp = phi * phi;
    // read back from the device GEMM table [i][o] (split-26 packing undone): forward i = r, o = w -> V[w][r];
    // inverse i = w, o = r -> the reference's inv_table[l][w][r] = V^-1[r][w] (HE.cu:263-266)
    std::vector<uint64_t> t(pp);
    cudaError_t e = cudaMemcpy(t.data(), (inverse ? c->d_wtab_inv : c->d_wtab_fwd) + (size_t)l * pp, pp * 8, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) return cuda_fail(e, "get_wcrt_table");
    auto unpack = [&](uint64_t v) { return c->wsplit ? ((v & 0xffffffffull) | ((v >> 32) << 26)) : v; };
    if (!inverse) {
        for (size_t w = 0; w < phi; ++w)
            for (size_t r = 0; r < phi; ++r) h[w * phi + r] = unpack(t[r * phi + w]);
        return MFHE_OK;
    }
    for (size_t i = 0; i < pp; ++i) h[i] = unpack(t[i]);
    return MFHE_OK;
}
extern "C" int mfhe_ctx_get_crt(const mfhe_ctx* c, int which, uint64_t* h) {
    if (!c || !h || which < 0 || which > 3) return fail(MFHE_EINVAL, "get_crt: bad argument");
    const std::vector<uint64_t>& v = which == 0 ? c->hQ : which == 1 ? c->hQhalf : which == 2 ? c->hM : c->hInv;
    std::memcpy(h, v.data(), v.size() * 8);
    return MFHE_OK;
}
extern "C" const mfhe_ntt_plan* mfhe_ctx_xy_plan(const mfhe_ctx* c) { return c ? c->xy_plan : nullptr; }
extern "C" int mfhe_ctx_copy_xy_matrix(const mfhe_ctx* c, int which, double* d_out, void* s) {
    if (!c || !d_out || which < 0 || which > 3) return fail(MFHE_EINVAL, "copy_xy_matrix: bad argument");
    const double2* src = which == 0 ? c->d_xyV : which == 1 ? c->d_xyVT : which == 2 ? c->d_xyVinv : c->d_xyVinvT;
    return done(cudaMemcpyAsync(d_out, src, c->n2() * sizeof(double2), cudaMemcpyDeviceToDevice, S(s)), "copy_xy_matrix");
}

// ---- W-CRT ------------------------------------------------------------------------------------------
static int wcrt_run(mfhe_ctx* c, bool inverse, const uint64_t* in, uint64_t* out, int il, int ol, void* s, const char* who) {
    if (!c || !in || !out) return fail(MFHE_EINVAL, std::string(who) + ": NULL argument");
    if ((il | ol) & ~1) return fail(MFHE_EINVAL, std::string(who) + ": unknown layout");
    if (in == out) return fail(MFHE_EINVAL, std::string(who) + ": in-place is not supported");
    WcrtArgs a = wcrt_args(c, inverse);
    a.in = in; a.out = (u64*)out; a.in_layout = il; a.out_layout = ol;
    return done(wcrt_launch(a, S(s)), who);
}
extern "C" int mfhe_wcrt_forward(mfhe_ctx* c, const uint64_t* in, uint64_t* out, int il, int ol, void* s) {
    return wcrt_run(c, false, in, out, il, ol, s, "wcrt_forward");
}
extern "C" int mfhe_wcrt_inverse(mfhe_ctx* c, const uint64_t* in, uint64_t* out, int il, int ol, void* s) {
    return wcrt_run(c, true, in, out, il, ol, s, "wcrt_inverse");
}
extern "C" int mfhe_wcrt_forward_vector(mfhe_ctx* c, const uint64_t* in, uint64_t* out, void* s) {
    if (!c || !in || !out || in == out) return fail(MFHE_EINVAL, "wcrt_forward_vector: bad argument");
    WcrtArgs a = wcrt_args(c, false);
    a.in = in; a.out = (u64*)out; a.cols = c->n;
    return done(wcrt_launch(a, S(s)), "wcrt_forward_vector");
}
extern "C" int mfhe_wcrt_forward_centered(mfhe_ctx* c, const int64_t* in, int64_t* out, void* s) {
    if (!c || !in || !out) return fail(MFHE_EINVAL, "wcrt_forward_centered: NULL argument");
    ArenaScope arena_scope(c, S(s));
    u64* tmp = (u64*)c->take(c->T() * 8);                      // [w][limb][n2]
    if (!tmp) return fail(MFHE_ENOMEM, "wcrt_forward_centered: arena too small");
    WcrtArgs a = wcrt_args(c, false);
    a.in = in; a.out = tmp; a.source = kSrcI64;
    cudaError_t e = wcrt_launch(a, S(s));
    if (e == cudaSuccess)
        e = crt_compose_i64_launch(crt_dev(c), tmp, (long long*)out, (int)c->n2(), c->L, c->phi, 1, S(s));
    return done(e, "wcrt_forward_centered");
}
extern "C" int mfhe_wcrt_inverse_centered(mfhe_ctx* c, const int64_t* in, int64_t* out, void* s) {
    if (!c || !in || !out) return fail(MFHE_EINVAL, "wcrt_inverse_centered: NULL argument");
    ArenaScope arena_scope(c, S(s));
    const size_t total = (size_t)c->phi * c->n2();
    u64* tmp = (u64*)c->take(total * 8);
    if (!tmp) return fail(MFHE_ENOMEM, "wcrt_inverse_centered: arena too small");
    WcrtArgs a = wcrt_args(c, true);                           // limb 0 only (HE.cu:1100-1111)
    a.in = in; a.out = tmp; a.source = kSrcI64; a.limbs = 1;
    cudaError_t e = wcrt_launch(a, S(s));
    if (e == cudaSuccess) e = center_i64_launch(tmp, (long long*)out, c->moduli[0], total, S(s));
    return done(e, "wcrt_inverse_centered");
}
extern "C" int mfhe_layout_convert(mfhe_ctx* c, const uint64_t* in, uint64_t* out, int il, int ol, void* s) {
    if (!c || !in || !out || in == out || ((il | ol) & ~1)) return fail(MFHE_EINVAL, "layout_convert: bad argument");
    return done(layout_launch((const u64*)in, (u64*)out, c->n, c->L, c->phi, il, ol, S(s)), "layout_convert");
}

// ---- complex W-DFT ------------------------------------------------------------------------------------
extern "C" int mfhe_wdft_forward_complex(mfhe_ctx* c, const double* in, double* out, void* s) {
    if (!c || !in || !out || in == out) return fail(MFHE_EINVAL, "wdft_forward_complex: bad argument");
    return done(wdft_launch(wdft_tab(c, false), (const double2*)in, (double2*)out, c->phi, (int)c->n2(), S(s)), "wdft_forward_complex");
}
extern "C" int mfhe_wdft_inverse_complex(mfhe_ctx* c, const double* in, double* out, void* s) {
    if (!c || !in || !out || in == out) return fail(MFHE_EINVAL, "wdft_inverse_complex: bad argument");
    return done(wdft_launch(wdft_tab(c, true), (const double2*)in, (double2*)out, c->phi, (int)c->n2(), S(s)), "wdft_inverse_complex");
}
extern "C" int mfhe_wdft_forward_centered_pair(mfhe_ctx* c, const int64_t* re, const int64_t* im, double* ore, double* oim, void* s) {
    if (!c || !re || !im || !ore || !oim) return fail(MFHE_EINVAL, "wdft_forward_centered_pair: NULL argument");
    return done(wdft_pair_launch(wdft_tab(c, false), re, im, ore, oim, c->phi, (int)c->n2(), 1, S(s)), "wdft_forward_centered_pair");
}
extern "C" int mfhe_wdft_inverse_pair(mfhe_ctx* c, const double* re, const double* im, double* ore, double* oim, void* s) {
    if (!c || !re || !im || !ore || !oim) return fail(MFHE_EINVAL, "wdft_inverse_pair: NULL argument");
    return done(wdft_pair_launch(wdft_tab(c, true), re, im, ore, oim, c->phi, (int)c->n2(), 0, S(s)), "wdft_inverse_pair");
}

// ---- XY embedding, quantise, CRT ------------------------------------------------------------------------
extern "C" int mfhe_xy_idft2(mfhe_ctx* c, const double* in, double* out, int lanes, void* s) {
    if (!c || !in || !out || lanes < 0) return fail(MFHE_EINVAL, "xy_idft2: bad argument");
    if (!lanes) return MFHE_OK;
    return done(xy_sandwich_launch(c->d_xyVinv, c->d_xyVinvT, (const double2*)in, (double2*)out, c->n, lanes, S(s)), "xy_idft2");
}
extern "C" int mfhe_xy_dft2(mfhe_ctx* c, const double* in, double* out, int lanes, void* s) {
    if (!c || !in || !out || lanes < 0) return fail(MFHE_EINVAL, "xy_dft2: bad argument");
    if (!lanes) return MFHE_OK;
    return done(xy_sandwich_launch(c->d_xyV, c->d_xyVT, (const double2*)in, (double2*)out, c->n, lanes, S(s)), "xy_dft2");
}
extern "C" int mfhe_matmul_complex(const double* A, const double* B, double* C, int n, void* s) {
    if (!A || !B || !C || n <= 0) return fail(MFHE_EINVAL, "matmul_complex: bad argument");
    return done(matmul_complex_launch((const double2*)A, (const double2*)B, (double2*)C, n, S(s)), "matmul_complex");
}
extern "C" int mfhe_quantize(mfhe_ctx* c, const double* in, uint64_t* re, uint64_t* im, int lanes, void* s) {
    if (!c || !in || !re || !im || lanes < 0) return fail(MFHE_EINVAL, "quantize: bad argument");
    if (!lanes) return MFHE_OK;
    return done(quantize_launch((const double2*)in, (u64*)re, (u64*)im, c->d_moduli, c->d_mu, (int)c->n2(), c->L, lanes,
                                c->delta, S(s)), "quantize");
}
extern "C" int mfhe_crt_compose_big(mfhe_ctx* c, const uint64_t* in, uint64_t* mag, uint8_t* neg, int n2, int limbs,
                                    int lanes, void* s) {
    if (!c || !in || !mag || !neg || n2 <= 0 || limbs < 1 || limbs > c->L || lanes < 1)
        return fail(MFHE_EINVAL, "crt_compose_big: bad argument");
    return done(crt_compose_big_launch(crt_dev(c), (const u64*)in, (u64*)mag, neg, n2, limbs, lanes, limbs, S(s)), "crt_compose_big");
}
extern "C" int mfhe_crt_compose_i64(mfhe_ctx* c, const uint64_t* in, int64_t* out, int n2, int limbs, void* s) {
    if (!c || !in || !out || n2 <= 0 || limbs < 1 || limbs > c->L) return fail(MFHE_EINVAL, "crt_compose_i64: bad argument");
    return done(crt_compose_i64_launch(crt_dev(c), (const u64*)in, (long long*)out, n2, limbs, 1, 0, S(s)), "crt_compose_i64");
}
extern "C" int mfhe_dequantize_exact(mfhe_ctx* c, const uint64_t* re, const uint64_t* im, double* out, int n2, int limbs, void* s) {
    if (!c || !re || !im || !out || n2 <= 0 || limbs < 1 || limbs > c->L) return fail(MFHE_EINVAL, "dequantize_exact: bad argument");
    return done(crt_to_complex_launch(crt_dev(c), (const u64*)re, (const u64*)im, (double2*)out, n2, limbs, 1, c->delta, S(s)),
                "dequantize_exact");
}
extern "C" int mfhe_big_pair_to_complex(mfhe_ctx* c, const uint64_t* rm, const uint8_t* rn, const uint64_t* im,
                                        const uint8_t* in_, double* out, size_t total, void* s) {
    if (!c || !rm || !rn || !im || !in_ || !out) return fail(MFHE_EINVAL, "big_pair_to_complex: NULL argument");
    if (!total) return MFHE_OK;
    return done(big_pair_to_complex_launch((const u64*)rm, rn, (const u64*)im, in_, (double2*)out, total, c->W, c->delta, S(s)),
                "big_pair_to_complex");
}

// ---- encode / decode --------------------------------------------------------------------------------------
extern "C" int mfhe_encode(mfhe_ctx* c, const double* msg, uint64_t* out_re, uint64_t* out_im, void* s) {
    if (!c || !msg || !out_re || !out_im) return fail(MFHE_EINVAL, "encode: NULL argument");
    ArenaScope arena_scope(c, S(s));
    const size_t lanes_bytes = (size_t)c->phi * c->n2() * sizeof(double2);
    double2* xy = (double2*)c->take(lanes_bytes);
    double2* wc = (double2*)c->take(lanes_bytes);
    if (!xy || !wc) return fail(MFHE_ENOMEM, "encode: arena too small");
    cudaStream_t st = S(s);
    // 1) per-lane XY-IDFT, all lanes in one launch   2) complex W-IDFT over the lane axis
    cudaError_t e = xy_sandwich_launch(c->d_xyVinv, c->d_xyVinvT, (const double2*)msg, xy, c->n, c->phi, st);
    if (e == cudaSuccess) e = wdft_launch(wdft_tab(c, true), xy, wc, c->phi, (int)c->n2(), st);
    // 3) llround(z * delta) once per value into two int64 planes (they reuse the xy buffer, dead by now)
    // 4) W-CRT: per-limb residues of a plane written to scratch, then the tensor-core GEMM, matrix-major result
    const size_t total = (size_t)c->phi * c->n2();
    long long* qre = (long long*)xy;
    long long* qim = qre + total;
    u64* scratch = (u64*)c->take(c->T() * 8);           // optional: without it the loader reduces on the fly
    if (e == cudaSuccess) e = llround_planes_launch(wc, qre, qim, total, c->delta, st);
    for (int part = 0; part < 2 && e == cudaSuccess; ++part) {
        WcrtArgs a = wcrt_args(c, false);
        a.in = part ? qim : qre; a.out = (u64*)(part ? out_im : out_re);
        a.source = kSrcI64; a.out_layout = MFHE_LAYOUT_MATRIX; a.scratch = scratch;
        e = wcrt_launch(a, st);
    }
    return done(e, "encode");
}
extern "C" int mfhe_encode_lane(mfhe_ctx* c, const double* msg, uint64_t* re, uint64_t* im, void* s) {
    if (!c || !msg || !re || !im) return fail(MFHE_EINVAL, "encode_lane: NULL argument");
    ArenaScope arena_scope(c, S(s));
    double2* P = (double2*)c->take(c->n2() * sizeof(double2));
    if (!P) return fail(MFHE_ENOMEM, "encode_lane: arena too small");
    cudaError_t e = xy_sandwich_launch(c->d_xyVinv, c->d_xyVinvT, (const double2*)msg, P, c->n, 1, S(s));
    if (e == cudaSuccess) e = quantize_launch(P, (u64*)re, (u64*)im, c->d_moduli, c->d_mu, (int)c->n2(), c->L, 1, c->delta, S(s));
    return done(e, "encode_lane");
}
// shared tail of decode: evals (any layout) -> complex message; uses arena from its current mark
static cudaError_t decode_tail(mfhe_ctx* c, const u64* er, const u64* ei, int layout, double2* out, cudaStream_t st) {
    const size_t T = c->T(), lanes_bytes = (size_t)c->phi * c->n2() * sizeof(double2);
    u64* cr = (u64*)c->take(T * 8);
    u64* ci = (u64*)c->take(T * 8);
    double2* cc = (double2*)c->take(lanes_bytes);
    double2* ec = (double2*)c->take(lanes_bytes);
    if (!cr || !ci || !cc || !ec) return cudaErrorMemoryAllocation;
    cudaError_t e = cudaSuccess;
    for (int part = 0; part < 2 && e == cudaSuccess; ++part) {
        WcrtArgs a = wcrt_args(c, true);
        a.in = part ? ei : er; a.out = part ? ci : cr; a.in_layout = layout; a.out_layout = MFHE_LAYOUT_MATRIX;
        e = wcrt_launch(a, st);
    }
    // CRT compose + centre + /delta for all lanes in one launch (reference: 1024 launches, HE.cu:1653-1672)
    if (e == cudaSuccess) e = crt_to_complex_launch(crt_dev(c), cr, ci, cc, (int)c->n2(), c->L, c->phi, c->delta, st);
    if (e == cudaSuccess) e = wdft_launch(wdft_tab(c, false), cc, ec, c->phi, (int)c->n2(), st);
    if (e == cudaSuccess) e = xy_sandwich_launch(c->d_xyV, c->d_xyVT, ec, out, c->n, c->phi, st);
    return e;
}
extern "C" int mfhe_decode_eval_pair(mfhe_ctx* c, const uint64_t* er, const uint64_t* ei, int layout, double* out, void* s) {
    if (!c || !er || !ei || !out || (layout & ~1)) return fail(MFHE_EINVAL, "decode_eval_pair: bad argument");
    ArenaScope arena_scope(c, S(s));
    cudaError_t e = decode_tail(c, (const u64*)er, (const u64*)ei, layout, (double2*)out, S(s));
    if (e == cudaErrorMemoryAllocation) return fail(MFHE_ENOMEM, "decode_eval_pair: arena too small");
    return done(e, "decode_eval_pair");
}
extern "C" int mfhe_decode_lane_from_rns_eval(mfhe_ctx* c, const uint64_t* re, const uint64_t* im, double* out, void* s) {
    if (!c || !re || !im || !out) return fail(MFHE_EINVAL, "decode_lane_from_rns_eval: NULL argument");
    ArenaScope arena_scope(c, S(s));
    double2* P = (double2*)c->take(c->n2() * sizeof(double2));
    if (!P) return fail(MFHE_ENOMEM, "decode_lane_from_rns_eval: arena too small");
    cudaError_t e = crt_to_complex_launch(crt_dev(c), (const u64*)re, (const u64*)im, P, (int)c->n2(), c->L, 1, c->delta, S(s));
    if (e == cudaSuccess) e = xy_sandwich_launch(c->d_xyV, c->d_xyVT, P, (double2*)out, c->n, 1, S(s));
    return done(e, "decode_lane_from_rns_eval");
}

// ---- samplers, keys -------------------------------------------------------------------------------------
// draws one array in the context's sampler mode; the sound mode consumes one draw number per array
static cudaError_t draw_array(mfhe_ctx* c, int kind, u64* out, cudaStream_t st) {
    if (c->sampler_mode == 0) return sample_launch(kind, out, c->d_moduli, c->d_mu, c->n, c->L, c->phi, st);
    return sound_sample_launch(kind, out, c->d_moduli, c->d_mu, c->n, c->L, c->phi, c->seed[0], c->seed[1], c->draw++, st);
}
static int sample(mfhe_ctx* c, int kind, uint64_t* out, void* s, const char* who) {
    if (!c || !out) return fail(MFHE_EINVAL, std::string(who) + ": NULL argument");
    return done(draw_array(c, kind, (u64*)out, S(s)), who);
}
extern "C" int mfhe_ctx_set_sampler(mfhe_ctx* c, int mode, const uint64_t* seed) {
    if (!c || (mode != 0 && mode != 1) || (mode == 1 && !seed)) return fail(MFHE_EINVAL, "ctx_set_sampler: bad argument");
    std::lock_guard<std::mutex> lk(c->arena_mu);
    c->sampler_mode = mode; c->draw = 0;
    c->seed[0] = mode ? seed[0] : 0; c->seed[1] = mode ? seed[1] : 0;
    return MFHE_OK;
}
extern "C" int mfhe_sampler_cdt(uint64_t* h_out, int count) {
    if (!h_out || count != 83) return fail(MFHE_EINVAL, "sampler_cdt: the table has 83 entries");
    sound_cdt_copy((u64*)h_out);
    return MFHE_OK;
}
extern "C" int mfhe_sample_uniform(mfhe_ctx* c, uint64_t* out, void* s) { return sample(c, 0, out, s, "sample_uniform"); }
extern "C" int mfhe_sample_gaussian(mfhe_ctx* c, uint64_t* out, void* s) { return sample(c, 1, out, s, "sample_gaussian"); }
extern "C" int mfhe_sample_ternary(mfhe_ctx* c, uint64_t* out, void* s) { return sample(c, 2, out, s, "sample_ternary"); }

extern "C" int mfhe_keygen(mfhe_ctx* c, uint64_t* sk, void* s) {
    if (!c || !sk) return fail(MFHE_EINVAL, "keygen: NULL argument");
    ArenaScope arena_scope(c, S(s));
    const size_t total = (size_t)c->phi * c->L * c->n;
    u64* sc = (u64*)c->take(total * 8);
    if (!sc) return fail(MFHE_ENOMEM, "keygen: arena too small");
    cudaError_t e = draw_array(c, 2, sc, S(s));
    if (e == cudaSuccess) {
        WcrtArgs a = wcrt_args(c, false);
        a.in = sc; a.out = (u64*)sk; a.cols = c->n;
        e = wcrt_launch(a, S(s));
    }
    if (e == cudaSuccess) e = ntt_small_launch(true, (u64*)sk, c->xy_plan->small(), c->L, (size_t)c->phi, S(s));
    return done(e, "keygen");
}

// ---- encrypt / decrypt -------------------------------------------------------------------------------------
// a_src / e_src: kSrcMem with pointers, or the fused sampler sources
static int encrypt_core(mfhe_ctx* c, const uint64_t* m_re, const uint64_t* m_im, const uint64_t* sk, int a_src,
                        const uint64_t* a_coeff, int e_src, const uint64_t* e_re, const uint64_t* e_im, uint64_t* ct_re,
                        uint64_t* ct_im, void* s, const char* who) {
    if (!c || !m_re || !sk || !ct_re || (ct_im && !m_im)) return fail(MFHE_EINVAL, std::string(who) + ": NULL argument");
    ArenaScope arena_scope(c, S(s));
    const size_t T = c->T();
    cudaStream_t st = S(s);
    if (c->sampler_mode == 1 && a_src == kSrcUniform) {
        // sound sampler: a and e are drawn (Philox, fresh draw numbers) into scratch and enter as memory sources; the
        // error of the im ciphertext is its OWN draw (the reference reuses one stream for both, HE.cu:1516-1517)
        u64* da = (u64*)c->take(T * 8);
        u64* de = (u64*)c->take(T * 8);
        u64* de2 = ct_im ? (u64*)c->take(T * 8) : nullptr;
        if (!da || !de || (ct_im && !de2)) return fail(MFHE_ENOMEM, std::string(who) + ": arena too small for the sound sampler");
        cudaError_t e0 = draw_array(c, 0, da, st);
        if (e0 == cudaSuccess) e0 = draw_array(c, 1, de, st);
        if (e0 == cudaSuccess && ct_im) e0 = draw_array(c, 1, de2, st);
        if (e0 != cudaSuccess) return done(e0, who);
        // the three arrays stay live: run the memory-source body on the rest of the arena
        a_src = kSrcMem; a_coeff = (const uint64_t*)da; e_src = kSrcMem; e_re = (const uint64_t*)de; e_im = (const uint64_t*)(ct_im ? de2 : de);
    }
    const bool two_e = ct_im && e_src == kSrcMem && e_im != e_re;
    u64* E0 = (u64*)c->take(T * 8);
    u64* E1 = two_e ? (u64*)c->take(T * 8) : E0;
    if (!E0 || !E1) return fail(MFHE_ENOMEM, std::string(who) + ": arena too small");
    // a: W-coeff -> W-eval, written straight into the ciphertexts' a halves (matrix-major)
    WcrtArgs a = wcrt_args(c, false);
    a.source = a_src; a.in = a_coeff; a.out = (u64*)ct_re + T; a.out2 = ct_im ? (u64*)ct_im + T : nullptr;
    // scratch for a sampled `a`: the b half of ct_re is free until the row kernel writes it (unless the caller
    // encrypts in place), else the arena
    a.scratch = ((const void*)m_re != (const void*)ct_re && (const void*)m_im != (const void*)ct_re) ? (u64*)ct_re : (u64*)c->take(T * 8);
    cudaError_t e = wcrt_launch(a, st);
    // e: W-coeff -> W-eval (the reference draws the same gaussian stream for e_re and e_im: HE.cu:1516-1517)
    if (e == cudaSuccess) {
        WcrtArgs g = wcrt_args(c, false);
        g.source = e_src; g.in = e_re; g.out = E0;
        if (e_src == kSrcGauss) {
            signed char* tab = (signed char*)c->take((size_t)c->phi * c->n2());
            if (!tab) return fail(MFHE_ENOMEM, std::string(who) + ": arena too small");
            e = gauss_table_launch(tab, c->phi, (int)c->n2(), st);
            g.in = tab;
            g.scratch = (u64*)c->take(T * 8);   // optional
        }
        if (e == cudaSuccess) e = wcrt_launch(g, st);
        if (e == cudaSuccess && two_e) { g.in = e_im; g.out = E1; e = wcrt_launch(g, st); }
    }
    // b = m - INTT(NTT(a) (.) s) + e, fused per row
    if (e == cudaSuccess) {
        RowArgs r;
        r.a = (const u64*)ct_re + T; r.s = (const u64*)sk;
        r.m0 = (const u64*)m_re; r.e0 = E0; r.o0 = (u64*)ct_re;
        if (ct_im) { r.m1 = (const u64*)m_im; r.e1 = E1; r.o1 = (u64*)ct_im; }
        r.subtract = 1; r.n = c->n; r.limbs = c->L; r.phi = c->phi;
        e = xrow_mul_s_launch(r, c->xy_plan->small_rows(), c->d_mu, st);
    }
    return done(e, who);
}
extern "C" int mfhe_encrypt_pair(mfhe_ctx* c, const uint64_t* m_re, const uint64_t* m_im, const uint64_t* sk,
                                 uint64_t* ct_re, uint64_t* ct_im, void* s) {
    if (!ct_im) return fail(MFHE_EINVAL, "encrypt_pair: NULL argument");
    return encrypt_core(c, m_re, m_im, sk, kSrcUniform, nullptr, kSrcGauss, nullptr, nullptr, ct_re, ct_im, s, "encrypt_pair");
}
extern "C" int mfhe_encrypt_pair_with(mfhe_ctx* c, const uint64_t* m_re, const uint64_t* m_im, const uint64_t* sk,
                                      const uint64_t* a_coeff, const uint64_t* e_re, const uint64_t* e_im,
                                      uint64_t* ct_re, uint64_t* ct_im, void* s) {
    if (!a_coeff || !e_re || !e_im || !ct_im) return fail(MFHE_EINVAL, "encrypt_pair_with: NULL argument");
    return encrypt_core(c, m_re, m_im, sk, kSrcMem, a_coeff, kSrcMem, e_re, e_im, ct_re, ct_im, s, "encrypt_pair_with");
}
// Fused BatchedEncoder::encode_to_wntt_eval + encrypt_pair (batched_encoder.cu:161-228, HE.cu:1455-1552): the same
// ciphertext words as the two calls in sequence, but the quantised message and the error are added in the W-coeff
// domain (two int64 planes) and transformed together -- 3 W-CRTs instead of 4, and the encoded message never
// makes a round trip through HBM on its own: the W-CRT writes (m + e) straight into the b halves, the row kernel
// then subtracts INTT(NTT(a) (.) s) in place.
extern "C" int mfhe_encode_encrypt_pair(mfhe_ctx* c, const double* msg, const uint64_t* sk, uint64_t* ct_re, uint64_t* ct_im,
                                        void* s) {
    if (!c || !msg || !sk || !ct_re || !ct_im) return fail(MFHE_EINVAL, "encode_encrypt_pair: NULL argument");
    if (c->sampler_mode == 1) {
        // the shared-W-CRT fusion below rests on the reference's gaussian stream; under the sound sampler the message is
        // encoded into the b halves and encrypted in place (same result as the two calls in sequence)
        int rc = mfhe_encode(c, msg, ct_re, ct_im, s);
        return rc != MFHE_OK ? rc : encrypt_core(c, ct_re, ct_im, sk, kSrcUniform, nullptr, kSrcGauss, nullptr, nullptr, ct_re, ct_im, s,
                                                 "encode_encrypt_pair");
    }
    ArenaScope arena_scope(c, S(s));
    const size_t T = c->T(), total = (size_t)c->phi * c->n2(), lanes_bytes = total * sizeof(double2);
    double2* xy = (double2*)c->take(lanes_bytes);
    double2* wc = (double2*)c->take(lanes_bytes);
    u64* scratch = (u64*)c->take(T * 8);
    if (!xy || !wc || !scratch) return fail(MFHE_ENOMEM, "encode_encrypt_pair: arena too small");
    cudaStream_t st = S(s);
    cudaError_t e = xy_sandwich_launch(c->d_xyVinv, c->d_xyVinvT, (const double2*)msg, xy, c->n, c->phi, st);
    if (e == cudaSuccess) e = wdft_launch(wdft_tab(c, true), xy, wc, c->phi, (int)c->n2(), st);
    long long* qre = (long long*)xy;
    long long* qim = qre + total;
    if (e == cudaSuccess) e = llround_noise_planes_launch(wc, qre, qim, total, c->delta, st);
    for (int part = 0; part < 2 && e == cudaSuccess; ++part) {          // (m + e) -> W-eval, into the b halves
        WcrtArgs a = wcrt_args(c, false);
        a.in = part ? qim : qre; a.out = (u64*)(part ? ct_im : ct_re);
        a.source = kSrcI64; a.out_layout = MFHE_LAYOUT_MATRIX; a.scratch = scratch;
        e = wcrt_launch(a, st);
    }
    if (e == cudaSuccess) {                                              // a: uniform W-coeff -> W-eval, both a halves
        WcrtArgs a = wcrt_args(c, false);
        a.source = kSrcUniform; a.in = nullptr; a.out = (u64*)ct_re + T; a.out2 = (u64*)ct_im + T; a.scratch = scratch;
        e = wcrt_launch(a, st);
    }
    if (e == cudaSuccess) {
        RowArgs r;
        r.a = (const u64*)ct_re + T; r.s = (const u64*)sk;
        r.m0 = (const u64*)ct_re; r.o0 = (u64*)ct_re;
        r.m1 = (const u64*)ct_im; r.o1 = (u64*)ct_im;
        r.subtract = 1; r.n = c->n; r.limbs = c->L; r.phi = c->phi;
        e = xrow_mul_s_launch(r, c->xy_plan->small_rows(), c->d_mu, st);
    }
    return done(e, "encode_encrypt_pair");
}
extern "C" int mfhe_encrypt(mfhe_ctx* c, const uint64_t* m, const uint64_t* sk, uint64_t* ct, void* s) {
    return encrypt_core(c, m, nullptr, sk, kSrcUniform, nullptr, kSrcGauss, nullptr, nullptr, ct, nullptr, s, "encrypt");
}

static cudaError_t decrypt_rows(mfhe_ctx* c, const u64* ct, const u64* sk, u64* out, int out_layout, cudaStream_t st) {
    RowArgs r;
    r.a = ct + c->T(); r.s = sk; r.m0 = ct; r.o0 = out; r.o_layout = out_layout;
    r.subtract = 0; r.n = c->n; r.limbs = c->L; r.phi = c->phi;
    return xrow_mul_s_launch(r, c->xy_plan->small_rows(), c->d_mu, st);
}
extern "C" int mfhe_decrypt_to_eval(mfhe_ctx* c, const uint64_t* ct, const uint64_t* sk, uint64_t* out, int layout, void* s) {
    if (!c || !ct || !sk || !out || (layout & ~1)) return fail(MFHE_EINVAL, "decrypt_to_eval: bad argument");
    return done(decrypt_rows(c, (const u64*)ct, (const u64*)sk, (u64*)out, layout, S(s)), "decrypt_to_eval");
}
extern "C" int mfhe_decrypt_decode(mfhe_ctx* c, const uint64_t* ct_re, const uint64_t* ct_im, const uint64_t* sk,
                                   double* out, void* s) {
    if (!c || !ct_re || !ct_im || !sk || !out) return fail(MFHE_EINVAL, "decrypt_decode: NULL argument");
    ArenaScope arena_scope(c, S(s));
    u64* mr = (u64*)c->take(c->T() * 8);
    u64* mi = (u64*)c->take(c->T() * 8);
    if (!mr || !mi) return fail(MFHE_ENOMEM, "decrypt_decode: arena too small");
    // both ciphertexts in one launch: rows whose a parts agree (always, for a pair made by encrypt_pair) share t
    RowArgs r;
    r.a = (const u64*)ct_re + c->T(); r.a1 = (const u64*)ct_im + c->T(); r.s = (const u64*)sk;
    r.m0 = (const u64*)ct_re; r.o0 = mr; r.m1 = (const u64*)ct_im; r.o1 = mi;
    r.o_layout = MFHE_LAYOUT_MATRIX; r.subtract = 0; r.n = c->n; r.limbs = c->L; r.phi = c->phi;
    cudaError_t e = xrow_mul_s_launch(r, c->xy_plan->small_rows(), c->d_mu, S(s));
    if (e == cudaSuccess) e = decode_tail(c, mr, mi, MFHE_LAYOUT_MATRIX, (double2*)out, S(s));
    if (e == cudaErrorMemoryAllocation) return fail(MFHE_ENOMEM, "decrypt_decode: arena too small");
    return done(e, "decrypt_decode");
}

extern "C" int mfhe_ct_add(mfhe_ctx* c, const uint64_t* c1, const uint64_t* c2, uint64_t* res, int ref, void* s) {
    if (!c || !c1 || !c2 || !res) return fail(MFHE_EINVAL, "ct_add: NULL argument");
    return done(ct_add_launch((const u64*)c1, (const u64*)c2, (u64*)res, c->d_moduli, c->T(), c->n, c->L, ref, S(s)), "ct_add");
}
extern "C" int mfhe_ct_mul_raw(mfhe_ctx* c, const uint64_t* c1, const uint64_t* c2, uint64_t* d0, uint64_t* d1, uint64_t* d2,
                               int ref, void* s) {
    if (!c || !c1 || !c2 || !d0 || !d1 || !d2) return fail(MFHE_EINVAL, "ct_mul_raw: NULL argument");
    return done(ct_mul_launch((const u64*)c1, (const u64*)c2, (u64*)d0, (u64*)d1, (u64*)d2, c->d_moduli, c->d_mu, c->T(), c->n,
                              c->L, ref, S(s)), "ct_mul_raw");
}

// ---- trace step -------------------------------------------------------------------------------------------
extern "C" int mfhe_trace_map(mfhe_ctx* c, const uint64_t* Br, const uint64_t* Bi, uint64_t* Bpr, uint64_t* Bpi, int batch, void* s) {
    if (!c || !Br || !Bi || !Bpr || !Bpi || batch < 1 || Br == Bpr || Bi == Bpi) return fail(MFHE_EINVAL, "trace_map: bad argument");
    return done(trace_map_launch((const u64*)Br, (const u64*)Bi, (u64*)Bpr, (u64*)Bpi, c->d_moduli, c->n, c->L, batch, S(s)), "trace_map");
}
extern "C" int mfhe_trace_gemm(mfhe_ctx* c, const uint64_t* Ar, const uint64_t* Ai, const uint64_t* Bpr, const uint64_t* Bpi,
                               uint64_t* Cr, uint64_t* Ci, int batch, void* s) {
    if (!c || !Ar || !Ai || !Bpr || !Bpi || !Cr || !Ci || batch < 1) return fail(MFHE_EINVAL, "trace_gemm: bad argument");
    // n = 64 and moduli below 2^48: int8 digit GEMMs on the tensor cores (trace_tc.cu); MFHE_TRACE_PATH=c pins the
    // CUDA-core kernel (A/B and parity tests)
    static const bool pin_cuda_core = [] { const char* e = std::getenv("MFHE_TRACE_PATH"); return e && e[0] == 'c'; }();
    if (!pin_cuda_core && trace_gemm_tc_eligible(c->n, c->L, c->ndig.data()))
        return done(trace_gemm_tc_launch((const u64*)Ar, (const u64*)Ai, (const u64*)Bpr, (const u64*)Bpi, (u64*)Cr, (u64*)Ci,
                                         c->d_moduli, c->d_mu, c->ndig.data(), c->L, batch, S(s)), "trace_gemm");
    return done(trace_gemm_launch((const u64*)Ar, (const u64*)Ai, (const u64*)Bpr, (const u64*)Bpi, (u64*)Cr, (u64*)Ci,
                                  c->d_moduli, c->d_mu, c->n, c->L, batch, S(s)), "trace_gemm");
}
extern "C" int mfhe_trace_rescale(mfhe_ctx* c, uint64_t* Cr, uint64_t* Ci, int batch, uint64_t inv0, uint64_t inv1,
                                  uint64_t inv2, void* s) {
    if (!c || !Cr || !Ci || batch < 1) return fail(MFHE_EINVAL, "trace_rescale: bad argument");
    return done(trace_rescale_launch((u64*)Cr, (u64*)Ci, c->d_moduli, c->d_mu, c->n, c->L, batch, inv0, inv1, inv2, S(s)),
                "trace_rescale");
}

// ---- encrypted matrix product -------------------------------------------------------------
// The slot matrix of a lane is M[a][b] = m(X = xi_b, Y = xi_a) with xi^n = i; the secret key depends on X only, so a
// contraction over the Y index of two ciphertexts is a plain GEMM on their component matrices -- no ring product, no
// relinearisation inside the sum.  With U = (b_U | a_U) and V' = (b_V' | a_V'), V' the X-twisted encoding of V
// (mfhe_trace_map_x, the map of trace.cu:30-62 along x):
//     D_uv[x][z] = n * sum_y  u[y][x] * v'[y][z]      u in {b_U, a_U}, v' in {b_V', a_V'}     (Gaussian integers mod q)
// and  D_bb + D_ab (*)_x s + D_ba (*)_z s + D_aa (*)_x s (*)_z s  decrypts to  n * sum_y U[y][x] V'[y][z]  (+ noise),
// the coefficient matrix in (X, Z) of  R[b][b'] = sum_a M_U[a][b] conj(M_V[a][b'])  (derivation in DESIGN.md).
extern "C" int mfhe_trace_map_x(mfhe_ctx* c, const uint64_t* Br, const uint64_t* Bi, uint64_t* Bpr, uint64_t* Bpi, int batch, void* s) {
    if (!c || !Br || !Bi || !Bpr || !Bpi || batch < 1 || Br == Bpr || Bi == Bpi) return fail(MFHE_EINVAL, "trace_map_x: bad argument");
    return done(trace_map_launch((const u64*)Br, (const u64*)Bi, (u64*)Bpr, (u64*)Bpi, c->d_moduli, c->n, c->L, batch, S(s), 1), "trace_map_x");
}
static cudaError_t gemm_ty(mfhe_ctx* c, const u64* Ur, const u64* Ui, const u64* Vr, const u64* Vi, u64* Dr, u64* Di, int batch,
                           cudaStream_t st) {
    static const bool pin_cuda_core = [] { const char* e = std::getenv("MFHE_TRACE_PATH"); return e && e[0] == 'c'; }();
    if (!pin_cuda_core && trace_gemm_tc_eligible(c->n, c->L, c->ndig.data()))
        return trace_gemm_tc_launch(Ur, Ui, Vr, Vi, Dr, Di, c->d_moduli, c->d_mu, c->ndig.data(), c->L, batch, st, 1);
    return trace_gemm_launch(Ur, Ui, Vr, Vi, Dr, Di, c->d_moduli, c->d_mu, c->n, c->L, batch, st, 1);
}
extern "C" int mfhe_trace_gemm_ty(mfhe_ctx* c, const uint64_t* Ur, const uint64_t* Ui, const uint64_t* Vr, const uint64_t* Vi,
                                  uint64_t* Dr, uint64_t* Di, int batch, void* s) {
    if (!c || !Ur || !Ui || !Vr || !Vi || !Dr || !Di || batch < 1) return fail(MFHE_EINVAL, "trace_gemm_ty: bad argument");
    return done(gemm_ty(c, (const u64*)Ur, (const u64*)Ui, (const u64*)Vr, (const u64*)Vi, (u64*)Dr, (u64*)Di, batch, S(s)), "trace_gemm_ty");
}
extern "C" int mfhe_transpose(mfhe_ctx* c, const uint64_t* in, uint64_t* out, size_t count, void* s) {
    if (!c || !in || !out || in == out) return fail(MFHE_EINVAL, "transpose: bad argument");
    return done(transpose_launch((const u64*)in, (u64*)out, c->n, count, S(s)), "transpose");
}
extern "C" int mfhe_trace_rescale_limbs(mfhe_ctx* c, uint64_t* Cr, uint64_t* Ci, int batch, const uint64_t* h_inv, void* s) {
    if (!c || !Cr || !Ci || !h_inv || batch < 1) return fail(MFHE_EINVAL, "trace_rescale_limbs: bad argument");
    RescaleConsts k{};
    for (int l = 0; l < c->L; ++l) k.inv[l] = h_inv[l];
    return done(trace_rescale_limbs_launch((u64*)Cr, (u64*)Ci, c->d_moduli, c->d_mu, c->n, c->L, batch, k, S(s)), "trace_rescale_limbs");
}
// out: 8 poly sets [D_bb re, im | D_ab re, im | D_ba re, im | D_aa re, im], each [w][l][x][z]
extern "C" int mfhe_ct_matmul(mfhe_ctx* c, const uint64_t* ctU_re, const uint64_t* ctU_im, const uint64_t* ctV_re,
                              const uint64_t* ctV_im, uint64_t* out, void* s) {
    if (!c || !ctU_re || !ctU_im || !ctV_re || !ctV_im || !out) return fail(MFHE_EINVAL, "ct_matmul: NULL argument");
    const size_t T = c->T();
    cudaError_t e = cudaSuccess;
    // MFHE_MATMUL_PATH: unset = tcgen05 kernel when the shape allows it; m = the four mma.sync GEMMs (trace_tc.cu)
    const char* path = std::getenv("MFHE_MATMUL_PATH");
    if (!(path && path[0] == 'm') && ct_matmul_umma_eligible(c->n, c->L, c->ndig.data()))
        return done(ct_matmul_umma_launch((const u64*)ctU_re, (const u64*)ctU_im, (const u64*)ctV_re, (const u64*)ctV_im, (u64*)out, T,
                                          c->d_moduli, c->d_mu, c->ndig.data(), c->L, c->phi, S(s)), "ct_matmul");
    for (int k = 0; k < 4 && e == cudaSuccess; ++k) {
        const size_t uo = (k & 1) ? T : 0, vo = (k & 2) ? T : 0;      // k: 0 bb, 1 ab, 2 ba, 3 aa
        e = gemm_ty(c, (const u64*)ctU_re + uo, (const u64*)ctU_im + uo, (const u64*)ctV_re + vo, (const u64*)ctV_im + vo,
                    (u64*)out + (size_t)(2 * k) * T, (u64*)out + (size_t)(2 * k + 1) * T, c->phi, S(s));
    }
    return done(e, "ct_matmul");
}
// prod (the 8 poly sets of mfhe_ct_matmul) is CONSUMED (used as scratch); out_re / out_im: [w][l][z][x] evals of the product
extern "C" int mfhe_ct_matmul_decrypt(mfhe_ctx* c, uint64_t* prod, const uint64_t* sk, uint64_t* out_re, uint64_t* out_im, void* s) {
    if (!c || !prod || !sk || !out_re || !out_im) return fail(MFHE_EINVAL, "ct_matmul_decrypt: NULL argument");
    const size_t T = c->T();
    u64* D = (u64*)prod;
    cudaStream_t st = S(s);
    // n = 64: one kernel per (lane, limb) matrix with R1, R2 in shared memory (scheme.cu); MFHE_MATMUL_DECRYPT_PATH=r keeps
    // the sequence of row-kernel passes and transposes below (any n)
    const char* dpath = std::getenv("MFHE_MATMUL_DECRYPT_PATH");
    if (c->n == 64 && !(dpath && dpath[0] == 'r'))
        return done(matmul_decrypt_launch(D, T, (const u64*)sk, (u64*)out_re, (u64*)out_im, c->xy_plan->small_rows(), c->d_mu,
                                          c->phi, c->L, st), "ct_matmul_decrypt");
    auto rows = [&](const u64* mr, const u64* mi, const u64* ar, const u64* ai, u64* o_r, u64* o_i) {
        RowArgs r;                                      // (o_r, o_i) = (mr, mi) + (ar, ai) (*) s along the fast index
        r.a = ar; r.a1 = ai; r.s = (const u64*)sk; r.m0 = mr; r.o0 = o_r; r.m1 = mi; r.o1 = o_i;
        r.subtract = 0; r.n = c->n; r.limbs = c->L; r.phi = c->phi;
        return xrow_mul_s_launch(r, c->xy_plan->small_rows(), c->d_mu, st);
    };
    const size_t mats = (size_t)c->phi * c->L;
    // 1. D_bb += D_ba (*)_z s      2. D_ab += D_aa (*)_z s      (in place)
    cudaError_t e = rows(D, D + T, D + 4 * T, D + 5 * T, D, D + T);
    if (e == cudaSuccess) e = rows(D + 2 * T, D + 3 * T, D + 6 * T, D + 7 * T, D + 2 * T, D + 3 * T);
    // 3. transposes into the two free pairs: [x][z] -> [z][x]
    for (int k = 0; k < 4 && e == cudaSuccess; ++k) e = transpose_launch(D + (size_t)k * T, D + (size_t)(4 + k) * T, c->n, mats, st);
    // 4. out = (D_bb)^T + (D_ab)^T (*)_x s
    if (e == cudaSuccess) e = rows(D + 4 * T, D + 5 * T, D + 6 * T, D + 7 * T, (u64*)out_re, (u64*)out_im);
    return done(e, "ct_matmul_decrypt");
}

// ---- RNS rescale of the matrix scheme ---------------------------------------------------------
// Divides a poly set by its last modulus q_last (~ delta: the reference's ten 35-bit primes next to delta = 2^35 are a
// rescale chain, config.h:25-44) and drops that limb:  X' = (X - [X]_{q_last}) / q_last  on the integer W-coefficients.
// In the W-eval domain of limb i < L-1 that is  (x_i - Wcrt_i([X]_{q_last})) * q_last^-1 mod q_i, and [X]_{q_last} (centred)
// is the inverse W-CRT of the last limb alone: 1/L of an inverse W-CRT + (L-1)/L of a forward one (int64 source).
// in: [w][L][n^2] evals (MATRIX layout), out: [w][L-1][n^2] -- a poly set of a context built on moduli[0 .. L-2].
extern "C" int mfhe_rescale_last(mfhe_ctx* c, const uint64_t* in, uint64_t* out, void* s) {
    if (!c || !in || !out || in == out) return fail(MFHE_EINVAL, "rescale_last: bad argument");
    if (c->L < 2) return fail(MFHE_EINVAL, "rescale_last: needs at least two limbs");
    ArenaScope arena_scope(c, S(s));
    cudaStream_t st = S(s);
    const int L = c->L, cols = (int)c->n2();
    const size_t lane = (size_t)c->phi * cols;
    u64* last = (u64*)c->take(lane * 8);
    u64* coef = (u64*)c->take(lane * 8);
    long long* cen = (long long*)c->take(lane * 8);
    if (!last || !coef || !cen) return fail(MFHE_ENOMEM, "rescale_last: arena too small");
    cudaError_t e = cudaMemcpy2DAsync(last, (size_t)cols * 8, (const u64*)in + (size_t)(L - 1) * cols, (size_t)L * cols * 8,
                                      (size_t)cols * 8, (size_t)c->phi, cudaMemcpyDeviceToDevice, st);
    if (e == cudaSuccess) {
        WcrtArgs a = wcrt_args(c, true);
        a.in = last; a.out = coef; a.limbs = 1; a.limb0 = L - 1;
        e = wcrt_launch(a, st);
    }
    if (e == cudaSuccess) e = center_i64_launch(coef, cen, c->moduli[L - 1], lane, st);
    if (e == cudaSuccess) {
        WcrtArgs a = wcrt_args(c, false);
        a.in = cen; a.out = (u64*)out; a.source = kSrcI64; a.limbs = L - 1; a.limb0 = 0;
        a.scratch = (u64*)c->take((size_t)(L - 1) * lane * 8);         // optional
        e = wcrt_launch(a, st);
    }
    if (e == cudaSuccess) {
        RescaleConsts k{};
        for (int l = 0; l < L - 1; ++l) k.inv[l] = host::invmod(c->moduli[L - 1] % c->moduli[l], c->moduli[l]);
        e = rescale_combine_launch((const u64*)in, (u64*)out, c->d_moduli, c->d_mu, cols, L, c->phi, k, st);
    }
    return done(e, "rescale_last");
}

// ---- GL permutation -----------------------------------------------------------------------------------------
// ntt_core.cu:150-173: perm[j] = bitrev((5^j mod 4n - 1) / 4)
extern "C" int mfhe_gl_perm_tables(int n, uint32_t* perm, uint32_t* inv_perm) {
    if (n < 2 || (n & (n - 1)) || !perm || !inv_perm) return fail(MFHE_EINVAL, "gl_perm_tables: bad argument");
    int logn = 0; while ((1 << logn) < n) ++logn;
    uint32_t m = 4u * (uint32_t)n, e = 1 % m;
    for (int j = 0; j < n; ++j) {
        const uint32_t t = host::bitrev((e - 1) / 4, logn);
        perm[j] = t; inv_perm[t] = (uint32_t)j;
        e = (uint32_t)((uint64_t)e * 5u % m);
    }
    return MFHE_OK;
}
extern "C" int mfhe_apply_gl_perm(const uint64_t* in, uint64_t* out, int limbs, int batch, int n, int inverse, void* s) {
    if (!in || !out || in == out || limbs < 1 || batch < 1 || n < 2 || n > 4096 || (n & (n - 1)))
        return fail(MFHE_EINVAL, "apply_gl_perm: bad argument");
    // tables are tiny: keep one device copy per n for the life of the process (the reference keeps
    // d_gl_perm / d_gl_inv_perm the same way, ntt_core.cu:16-18)
    static uint32_t* d_tabs[16][13][2] = {};                // per device
    static std::mutex tab_mu;
    int dev = 0; cudaGetDevice(&dev);
    if (dev < 0 || dev >= 16) return fail(MFHE_EUNSUPPORTED, "apply_gl_perm: device ordinal above 15");
    std::lock_guard<std::mutex> lk(tab_mu);
    auto& d_tab = d_tabs[dev];
    int logn = 0; while ((1 << logn) < n) ++logn;
    if (!d_tab[logn][0]) {
        std::vector<uint32_t> p(n), ip(n);
        mfhe_gl_perm_tables(n, p.data(), ip.data());
        cudaError_t e = upload(d_tab[logn][0], p.data(), (size_t)n * 4);
        if (e == cudaSuccess) e = upload(d_tab[logn][1], ip.data(), (size_t)n * 4);
        if (e != cudaSuccess) return cuda_fail(e, "apply_gl_perm tables");
    }
    return done(gl_perm_launch((const u64*)in, (u64*)out, d_tab[logn][inverse ? 1 : 0], limbs, batch, n, S(s)), "apply_gl_perm");
}
