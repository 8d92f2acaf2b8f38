// capi_ntt.cu -- C ABI (include/mfhe.h) for the NTT family: plans, transforms, pointwise ops.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "../../include/mfhe.h"
#include "capi_common.h"
#include "ntt_api.h"
#include "numth_host.hpp"
#include "plan.h"

using namespace mfhe;

namespace mfhe {
thread_local std::string g_last_error;
int fail(int code, const std::string& msg) { g_last_error = msg; return code; }
int cuda_fail(cudaError_t e, const char* where) {
    return fail(MFHE_ECUDA, std::string(where) + ": " + cudaGetErrorString(e));
}
}  // namespace mfhe

extern "C" const char* mfhe_last_error(void) { return g_last_error.c_str(); }
extern "C" const char* mfhe_version(void) { return "mfhe-b200 0.1 (sm_100a)"; }

extern "C" void mfhe_ntt_plan_destroy(mfhe_ntt_plan* p) {
    if (!p) return;
    cudaFree(p->d_moduli); cudaFree(p->d_tw); cudaFree(p->d_itw);
    cudaFree(p->d_ninv); cudaFree(p->d_w1n); cudaFree(p->d_stage); cudaFree(p->d_mu);
    cudaFree(p->d_twd); cudaFree(p->d_itwd); cudaFree(p->d_ninvd); cudaFree(p->d_w1nd);
    for (cudaStream_t st : p->host_streams) if (st) cudaStreamDestroy(st);
    delete p;
}

extern "C" int mfhe_ntt_plan_create(int logn, int L, const uint64_t* moduli, int convention,
                                    mfhe_ntt_plan** out) {
    if (!out) return fail(MFHE_EINVAL, "plan_create: out is NULL");
    *out = nullptr;
    const bool small = logn >= 4 && logn <= 8, flat = logn >= 12 && logn <= 17;
    if (!small && !flat) return fail(MFHE_EUNSUPPORTED, "plan_create: logn must be in 4..8 or 12..17");
    if (L <= 0 || !moduli) return fail(MFHE_EINVAL, "plan_create: no moduli");
    if (convention < 0 || convention > 2) return fail(MFHE_EINVAL, "plan_create: unknown convention");
    if (convention != MFHE_NTT_PHANTOM_NEGACYCLIC && !small)
        return fail(MFHE_EUNSUPPORTED, "plan_create: gl/cyclic conventions need logn <= 8");
    mfhe_ntt_plan* p = new (std::nothrow) mfhe_ntt_plan;
    if (!p) return fail(MFHE_ENOMEM, "plan_create: host allocation failed");
    p->logn = logn; p->L = L; p->convention = convention;
    p->moduli.assign(moduli, moduli + L);
    const size_t n = (size_t)1 << logn;
    p->h_tw.resize(n * L); p->h_itw.resize(n * L);
    std::vector<host::TwPair> ninv(L), w1n(L), mu(L);
    try {
        for (int l = 0; l < L; ++l) {
            const uint64_t q = moduli[l];
            if (q >= (1ull << 60) || q < 3 || !(q & 1))
                throw std::runtime_error("modulus must be odd and below 2^60");
            uint64_t psi, M, c0;
            if (convention == MFHE_NTT_PHANTOM_NEGACYCLIC) {
                M = 2 * n; c0 = n; psi = host::min_primitive_root(M, q);
            } else {
                M = 4 * n; psi = host::gl_psi(q, (int)n);
                c0 = convention == MFHE_NTT_GL_XN_MINUS_I ? n : 0;
            }
            if (!psi) throw std::runtime_error("modulus " + std::to_string(q) + " has no root of the required order");
            p->roots.push_back(psi);
            host::TreeTables t = host::build_tree_tables(logn, q, psi, M, c0);
            std::memcpy(&p->h_tw[l * n], t.tw.data(), n * sizeof(host::TwPair));
            std::memcpy(&p->h_itw[l * n], t.itw.data(), n * sizeof(host::TwPair));
            ninv[l] = t.ninv; w1n[l] = t.w1n;
            const unsigned __int128 m = ~(unsigned __int128)0 / q;   // floor(2^128/q), q odd
            mu[l] = host::TwPair{(uint64_t)m, (uint64_t)(m >> 64)};
        }
    } catch (const std::exception& e) {
        delete p;
        return fail(MFHE_EINVAL, std::string("plan_create: ") + e.what());
    }
    cudaError_t e;
#define UP(dst, src, bytes)                                                              \
    if ((e = cudaMalloc((void**)&(dst), (bytes))) != cudaSuccess ||                       \
        (e = cudaMemcpy((dst), (src), (bytes), cudaMemcpyHostToDevice)) != cudaSuccess) { \
        mfhe_ntt_plan_destroy(p); return cuda_fail(e, "plan_create upload"); }
    UP(p->d_moduli, p->moduli.data(), L * sizeof(uint64_t));
    UP(p->d_tw, p->h_tw.data(), n * L * sizeof(tw_t));
    UP(p->d_itw, p->h_itw.data(), n * L * sizeof(tw_t));
    UP(p->d_ninv, ninv.data(), L * sizeof(tw_t));
    UP(p->d_w1n, w1n.data(), L * sizeof(tw_t));
    UP(p->d_mu, mu.data(), L * sizeof(ulonglong2));
#undef UP
    // every modulus below 2^50: the transforms can run on the FP64 pipe (modmath.cuh); on by default, MFHE_NTT_ARITH=int
    // keeps the integer kernels, mfhe_ntt_plan_set_arith switches a plan
    uint64_t qmax = 0;
    for (int l = 0; l < L; ++l) qmax = std::max(qmax, moduli[l]);
    if (qmax < (1ull << 50)) {
        if ((e = cudaMalloc((void**)&p->d_twd, n * L * sizeof(tw_t))) != cudaSuccess ||
            (e = cudaMalloc((void**)&p->d_itwd, n * L * sizeof(tw_t))) != cudaSuccess ||
            (e = cudaMalloc((void**)&p->d_ninvd, L * sizeof(tw_t))) != cudaSuccess ||
            (e = cudaMalloc((void**)&p->d_w1nd, L * sizeof(tw_t))) != cudaSuccess ||
            (e = mfhe::ntt_fp_tables_launch(p->d_tw, p->d_twd, p->d_moduli, n, L, nullptr)) != cudaSuccess ||
            (e = mfhe::ntt_fp_tables_launch(p->d_itw, p->d_itwd, p->d_moduli, n, L, nullptr)) != cudaSuccess ||
            (e = mfhe::ntt_fp_tables_launch(p->d_ninv, p->d_ninvd, p->d_moduli, 1, L, nullptr)) != cudaSuccess ||
            (e = mfhe::ntt_fp_tables_launch(p->d_w1n, p->d_w1nd, p->d_moduli, 1, L, nullptr)) != cudaSuccess ||
            (e = cudaDeviceSynchronize()) != cudaSuccess) {
            mfhe_ntt_plan_destroy(p); return cuda_fail(e, "plan_create fp64 tables");
        }
        const char* env = getenv("MFHE_NTT_ARITH");
        p->arith = (env && !strcmp(env, "int")) ? 0 : 1;
    }
    *out = p;
    return MFHE_OK;
}

extern "C" int mfhe_ntt_plan_arith(const mfhe_ntt_plan* p) { return p ? p->arith : MFHE_EINVAL; }
extern "C" int mfhe_ntt_plan_set_arith(mfhe_ntt_plan* p, int arith) {
    if (!p || (arith != 0 && arith != 1)) return fail(MFHE_EINVAL, "set_arith: 0 (integer) or 1 (FP64 pipe)");
    if (arith && !p->d_twd) return fail(MFHE_EUNSUPPORTED, "set_arith: the FP64-pipe arithmetic needs every modulus of the plan below 2^50");
    p->arith = arith;
    return MFHE_OK;
}

extern "C" int mfhe_ntt_plan_logn(const mfhe_ntt_plan* p) { return p ? p->logn : MFHE_EINVAL; }
extern "C" int mfhe_ntt_plan_limbs(const mfhe_ntt_plan* p) { return p ? p->L : MFHE_EINVAL; }
extern "C" uint64_t mfhe_ntt_plan_root(const mfhe_ntt_plan* p, int limb) {
    return (p && limb >= 0 && limb < p->L) ? p->roots[limb] : 0;
}
extern "C" int mfhe_ntt_plan_get_table(const mfhe_ntt_plan* p, int limb, int inverse, uint64_t* w,
                                       uint64_t* ws) {
    if (!p || limb < 0 || limb >= p->L || !w || !ws) return fail(MFHE_EINVAL, "get_table: bad argument");
    const size_t n = (size_t)1 << p->logn;
    const host::TwPair* t = (inverse ? p->h_itw.data() : p->h_tw.data()) + limb * n;
    for (size_t i = 0; i < n; ++i) { w[i] = t[i].w; ws[i] = t[i].ws; }
    return MFHE_OK;
}
extern "C" int mfhe_ntt_plan_set_chunk(mfhe_ntt_plan* p, size_t chunk) {
    if (!p) return fail(MFHE_EINVAL, "set_chunk: plan is NULL");
    p->chunk = chunk;
    return MFHE_OK;
}

static int run(const mfhe_ntt_plan* p, bool fwd, uint64_t* d, size_t batch, int ls, int lc, void* stream) {
    if (!p || (!d && batch)) return fail(MFHE_EINVAL, "ntt: NULL plan or data");
    if (ls < 0 || lc <= 0 || ls + lc > p->L) return fail(MFHE_EINVAL, "ntt: limb window out of range");
    if (batch == 0) return MFHE_OK;
    cudaStream_t st = (cudaStream_t)stream;
    cudaError_t e;
    if (p->logn <= 8) {
        SmallTables t = p->small();                 // limb window = the same tables seen from limb `ls`
        t.moduli += ls; t.ninv += ls; t.w1n += ls;
        t.tw += (size_t)ls << p->logn; t.itw += (size_t)ls << p->logn;
        e = ntt_small_launch(fwd, (u64*)d, t, lc, batch, st);
    } else {
        e = ntt_flat_launch(fwd, (u64*)d, p->flat(), ls, lc, batch, p->chunk, st);
    }
    return e == cudaSuccess ? MFHE_OK : cuda_fail(e, fwd ? "ntt_forward" : "ntt_inverse");
}

extern "C" int mfhe_ntt_forward(const mfhe_ntt_plan* p, uint64_t* d, size_t batch, void* s) {
    return run(p, true, d, batch, 0, p ? p->L : 0, s);
}
extern "C" int mfhe_ntt_inverse(const mfhe_ntt_plan* p, uint64_t* d, size_t batch, void* s) {
    return run(p, false, d, batch, 0, p ? p->L : 0, s);
}
extern "C" int mfhe_ntt_forward_limbs(const mfhe_ntt_plan* p, uint64_t* d, size_t batch, int ls, int lc, void* s) {
    return run(p, true, d, batch, ls, lc, s);
}
extern "C" int mfhe_ntt_inverse_limbs(const mfhe_ntt_plan* p, uint64_t* d, size_t batch, int ls, int lc, void* s) {
    return run(p, false, d, batch, ls, lc, s);
}

// Host-buffer path: the batch is cut into chunks of whole polys and pipelined over three streams so
// that the H2D copy of chunk i+1, the transform of chunk i and the D2H copy of chunk i-1 overlap
// (PCIe is full duplex; with pinned host memory the step costs ~max(H2D, D2H) instead of their sum).
static int run_host(mfhe_ntt_plan* p, bool fwd, uint64_t* h, size_t batch) {
    if (!p || (!h && batch)) return fail(MFHE_EINVAL, "ntt_host: NULL plan or data");
    if (batch == 0) return MFHE_OK;
    const size_t poly_words = (size_t)p->L << p->logn;
    const size_t bytes = batch * poly_words * sizeof(uint64_t);
    cudaError_t e;
    if (bytes > p->stage_bytes) {
        cudaFree(p->d_stage); p->d_stage = nullptr; p->stage_bytes = 0;
        if ((e = cudaMalloc((void**)&p->d_stage, bytes)) != cudaSuccess) return cuda_fail(e, "ntt_host staging");
        p->stage_bytes = bytes;
    }
    cudaStream_t* st = p->host_streams;          // owned by the plan (a plan belongs to one device)
    if (!st[0])
        for (int i = 0; i < 3; ++i)
            if ((e = cudaStreamCreateWithFlags(&st[i], cudaStreamNonBlocking)) != cudaSuccess) return cuda_fail(e, "ntt_host stream");
    // ~32 MiB chunks, at least one poly, at most 64 chunks
    size_t per = std::max<size_t>(1, ((size_t)32 << 20) / (poly_words * 8));
    if ((batch + per - 1) / per > 64) per = (batch + 63) / 64;
    const size_t nchunks = (batch + per - 1) / per;
    std::vector<cudaEvent_t> up(nchunks), done(nchunks);
    int rc = MFHE_OK;
    for (size_t c = 0; c < nchunks; ++c) { cudaEventCreateWithFlags(&up[c], cudaEventDisableTiming); cudaEventCreateWithFlags(&done[c], cudaEventDisableTiming); }
    for (size_t c = 0; c < nchunks && rc == MFHE_OK; ++c) {
        const size_t b0 = c * per, nb = std::min(per, batch - b0), off = b0 * poly_words;
        if ((e = cudaMemcpyAsync(p->d_stage + off, h + off, nb * poly_words * 8, cudaMemcpyHostToDevice, st[0])) != cudaSuccess) { rc = cuda_fail(e, "ntt_host H2D"); break; }
        cudaEventRecord(up[c], st[0]);
        cudaStreamWaitEvent(st[1], up[c], 0);
        rc = run(p, fwd, p->d_stage + off, nb, 0, p->L, st[1]);
        if (rc) break;
        cudaEventRecord(done[c], st[1]);
        cudaStreamWaitEvent(st[2], done[c], 0);
        if ((e = cudaMemcpyAsync(h + off, p->d_stage + off, nb * poly_words * 8, cudaMemcpyDeviceToHost, st[2])) != cudaSuccess) rc = cuda_fail(e, "ntt_host D2H");
    }
    for (int i = 0; i < 3; ++i)
        if ((e = cudaStreamSynchronize(st[i])) != cudaSuccess && rc == MFHE_OK) rc = cuda_fail(e, "ntt_host sync");
    for (size_t c = 0; c < nchunks; ++c) { cudaEventDestroy(up[c]); cudaEventDestroy(done[c]); }
    return rc;
}
extern "C" int mfhe_ntt_forward_host(mfhe_ntt_plan* p, uint64_t* h, size_t batch) { return run_host(p, true, h, batch); }
extern "C" int mfhe_ntt_inverse_host(mfhe_ntt_plan* p, uint64_t* h, size_t batch) { return run_host(p, false, h, batch); }

// ---- pointwise ------------------------------------------------------------------------
namespace {
enum PwOp { kMul, kAdd, kSub };

template <int OP>
__global__ void __launch_bounds__(256)
pointwise_kernel(const u64* __restrict__ a, const u64* __restrict__ b, u64* __restrict__ out,
                 const u64* __restrict__ moduli, const ulonglong2* __restrict__ mu, int logn, int L,
                 size_t total) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const u64 q = moduli[(i >> logn) % L];
        const u64 x = a[i], y = b[i];
        u64 r;
        if (OP == kAdd) r = csub(x + y, q);
        else if (OP == kSub) r = csub(x + q - y, q);
        else {
            const ulonglong2 m = mu[(i >> logn) % L];
            r = mul_barrett(x, y, q, m.x, m.y);
        }
        out[i] = r;
    }
}
}  // namespace

static int pointwise(const mfhe_ntt_plan* p, int op, const uint64_t* a, const uint64_t* b, uint64_t* o,
                     size_t batch, void* stream) {
    if (!p || !a || !b || !o) return fail(MFHE_EINVAL, "pointwise: NULL argument");
    const size_t total = batch * p->L << p->logn;
    if (!total) return MFHE_OK;
    const unsigned grid = (unsigned)std::min<size_t>((total + 255) / 256, 148 * 16);
    cudaStream_t st = (cudaStream_t)stream;
    if (op == kMul) pointwise_kernel<kMul><<<grid, 256, 0, st>>>((const u64*)a, (const u64*)b, (u64*)o, (const u64*)p->d_moduli, p->d_mu, p->logn, p->L, total);
    else if (op == kAdd) pointwise_kernel<kAdd><<<grid, 256, 0, st>>>((const u64*)a, (const u64*)b, (u64*)o, (const u64*)p->d_moduli, p->d_mu, p->logn, p->L, total);
    else pointwise_kernel<kSub><<<grid, 256, 0, st>>>((const u64*)a, (const u64*)b, (u64*)o, (const u64*)p->d_moduli, p->d_mu, p->logn, p->L, total);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? MFHE_OK : cuda_fail(e, "pointwise");
}
extern "C" int mfhe_poly_mul(const mfhe_ntt_plan* p, const uint64_t* a, const uint64_t* b, uint64_t* o, size_t batch, void* s) { return pointwise(p, kMul, a, b, o, batch, s); }
extern "C" int mfhe_poly_add(const mfhe_ntt_plan* p, const uint64_t* a, const uint64_t* b, uint64_t* o, size_t batch, void* s) { return pointwise(p, kAdd, a, b, o, batch, s); }
extern "C" int mfhe_poly_sub(const mfhe_ntt_plan* p, const uint64_t* a, const uint64_t* b, uint64_t* o, size_t batch, void* s) { return pointwise(p, kSub, a, b, o, batch, s); }

// a <- INTT(NTT(a) (.) NTT(b)); b ends up in NTT form.  Flat plans multiply inside the first inverse pass.
extern "C" int mfhe_ntt_mul_intt(const mfhe_ntt_plan* p, uint64_t* a, uint64_t* b, int b_is_ntt, size_t batch, void* s) {
    if (!p || ((!a || !b) && batch)) return fail(MFHE_EINVAL, "ntt_mul_intt: NULL plan or data");
    if (a == b) return fail(MFHE_EINVAL, "ntt_mul_intt: a and b must be different buffers");
    if (batch == 0) return MFHE_OK;
    int rc = run(p, true, a, batch, 0, p->L, s);
    if (rc == MFHE_OK && !b_is_ntt) rc = run(p, true, b, batch, 0, p->L, s);
    if (rc != MFHE_OK) return rc;
    if (p->logn <= 8) {
        rc = pointwise(p, kMul, a, b, a, batch, s);
        return rc == MFHE_OK ? run(p, false, a, batch, 0, p->L, s) : rc;
    }
    cudaError_t e = ntt_flat_launch(false, (u64*)a, p->flat(), 0, p->L, batch, p->chunk, (cudaStream_t)s, (const u64*)b, p->d_mu);
    return e == cudaSuccess ? MFHE_OK : cuda_fail(e, "ntt_mul_intt");
}
