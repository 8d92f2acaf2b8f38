// bconv.cu -- fast RNS base conversion, Phantom's DBaseConverter::bConv_BEHZ semantics
// (ph/src/rns_bconv.cu:22-29 bconv_mult, :109-131 bconv_matmul, :212-229 driver):
//     y_i   = x_i * (Q/q_i)^-1 mod q_i                      i over the input base
//     out_j = ( sum_i y_i * (Q/q_i mod p_j) ) mod p_j       j over the output base
// One kernel instead of two and no [L_in][N] temporary in HBM: a CTA owns 128 coefficients, builds
// their y_i in shared memory, then every thread forms all L_out sums for its coefficient with a
// 128-bit lazy accumulator and a single Barrett reduction each (the accumulation trick of
// ph/include/rns_bconv.cuh:112-126).  The sharded form computes the partial sum over a window of
// input limbs (already reduced mod p_j, so partials from several GPUs add without overflow) and
// mfhe_bconv_reduce adds the partials: this is the only step of the whole path that needs a
// collective.
#include <algorithm>
#include <cstdlib>
#include <new>
#include <string>
#include <vector>

#include "../../include/mfhe.h"
#include "capi_common.h"
#include "modmath.cuh"
#include "numth_host.hpp"

using namespace mfhe;

struct mfhe_bconv_plan {
    int l_in = 0, l_out = 0;
    std::vector<uint64_t> q, p;
    u64 *d_q = nullptr, *d_p = nullptr;
    tw_t* d_qhat_inv = nullptr;        // [l_in] {(Q/q_i)^-1 mod q_i, shoup}
    u64* d_qhat_mod_p = nullptr;       // [l_out][l_in]
    ulonglong2* d_mu_p = nullptr;      // [l_out] floor(2^128 / p_j)
};

namespace {
constexpr int BC = 128;    // coefficients per CTA
constexpr int kScatterMax = 128;

// Destination table of the fused partial + exchange form (passed by value, 3 KiB of kernel parameters):
// output limb j of batch item b goes to base[j] + b * bstride[j]; base[j] may be a peer GPU's memory.
struct BconvScatter { u64* base[kScatterMax]; size_t bstride[kScatterMax]; };

// W1: the accumulated sum is below 2^(k+61) (k = bits of p_j; host-checked: in_count * max q <= 2^61), so ONE-word
// Barrett reduces it (same bound as mul_barrett1: qhat >= floor(v / p) - 2): one 64x64 high product instead of the
// five wide products of reduce128.  At in_count = 6 (config 5 on 8 ranks) the reduction is as expensive as the sums.
// IC > 0: in_count is the compile-time constant IC -- the y_i stay in registers (no shared memory), the sum over i is
// fully unrolled and two output limbs are formed per trip, so the loop and address arithmetic that made the generic
// form issue-bound (ncu: 4 200 instructions per coefficient for 96 multiply-accumulates, profiles/r1_21) disappears.
// IC == 0: any in_count, y_i in shared memory.
template <bool W1>
__device__ __forceinline__ u64 bconv_fold(u64 lo, u64 hi, u64 pj, ulonglong2 m) {
    if constexpr (W1) {
        const barrett1_t bc = make_barrett1(pj, m.x, m.y);
        const u64 x1 = (lo >> bc.sh) | (hi << (64 - bc.sh));
        const u64 r = lo - __umul64hi(x1, bc.mu1) * bc.q;
        return csub(csub(r, bc.q2), bc.q);
    } else {
        return reduce128(lo, hi, pj, m.x, m.y);
    }
}
// (hi:lo) += y * c on 32-bit halves with explicit carry chains: four 32x32->64 multiply-adds (ptxas pairs each
// mad.lo.cc / madc.hi into one IMAD.WIDE) + carry adds, instead of the separate low and high 64-bit products the
// compiler emits for  y * c  and  __umul64hi(y, c)  (about 16 IMAD + 10 IADD3 per term).
__device__ __forceinline__ void mac128(u64& lo, u64& hi, u64 y, u64 c) {
    u32 a0 = (u32)lo, a1 = (u32)(lo >> 32), a2 = (u32)hi, a3 = (u32)(hi >> 32);
    const u32 y0 = (u32)y, y1 = (u32)(y >> 32), c0 = (u32)c, c1 = (u32)(c >> 32);
    asm("mad.lo.cc.u32 %0, %4, %6, %0;\n\t"
        "madc.hi.cc.u32 %1, %4, %6, %1;\n\t"
        "addc.cc.u32 %2, %2, 0;\n\t"
        "addc.u32 %3, %3, 0;\n\t"
        "mad.lo.cc.u32 %1, %4, %7, %1;\n\t"
        "madc.hi.cc.u32 %2, %4, %7, %2;\n\t"
        "addc.u32 %3, %3, 0;\n\t"
        "mad.lo.cc.u32 %1, %5, %6, %1;\n\t"
        "madc.hi.cc.u32 %2, %5, %6, %2;\n\t"
        "addc.u32 %3, %3, 0;\n\t"
        "mad.lo.cc.u32 %2, %5, %7, %2;\n\t"
        "madc.hi.u32 %3, %5, %7, %3;"
        : "+r"(a0), "+r"(a1), "+r"(a2), "+r"(a3) : "r"(y0), "r"(y1), "r"(c0), "r"(c1));
    lo = ((u64)a1 << 32) | a0; hi = ((u64)a3 << 32) | a2;
}

template <bool SC, bool W1, int IC>
__global__ void __launch_bounds__(BC)
bconv_kernel(const u64* __restrict__ in, u64* __restrict__ out, size_t N, int l_in_total, int in_start, int in_count,
             int l_out, const u64* __restrict__ q, const tw_t* __restrict__ qhat_inv, const u64* __restrict__ qhat_mod_p,
             const u64* __restrict__ p, const ulonglong2* __restrict__ mu_p, const __grid_constant__ BconvScatter sc) {
    extern __shared__ u64 sy[];        // [in_count][BC], IC == 0 only
    const size_t d = (size_t)blockIdx.x * BC + threadIdx.x;
    const size_t b = blockIdx.y;
    const bool live = d < N;
    const u64* x = in + b * (size_t)in_count * N;
    u64 yr[IC > 0 ? IC : 1];
    if constexpr (IC > 0) {
        if (!live) return;
#pragma unroll
        for (int i = 0; i < IC; ++i) yr[i] = x[(size_t)i * N + d];
#pragma unroll
        for (int i = 0; i < IC; ++i) yr[i] = mul_shoup(yr[i], qhat_inv[in_start + i], q[in_start + i]);
    } else {
        for (int i = 0; i < in_count; ++i) {
            const int gi = in_start + i;
            sy[i * BC + threadIdx.x] = live ? mul_shoup(x[(size_t)i * N + d], qhat_inv[gi], q[gi]) : 0;
        }
        // each thread reads back only its own column: no barrier needed
        if (!live) return;
    }
    u64* o = SC ? nullptr : out + b * (size_t)l_out * N;
    auto put = [&](int j, u64 r) {
        if constexpr (SC) sc.base[j][b * sc.bstride[j] + d] = r;
        else              o[(size_t)j * N + d] = r;
    };
    if constexpr (IC > 0) {
        int j = 0;
        for (; j + 2 <= l_out; j += 2) {
            const u64* w0 = qhat_mod_p + (size_t)j * l_in_total + in_start;
            const u64* w1 = w0 + l_in_total;
            u64 lo0 = 0, hi0 = 0, lo1 = 0, hi1 = 0;
#pragma unroll
            for (int i = 0; i < IC; ++i) {
                mac128(lo0, hi0, yr[i], __ldg(w0 + i));
                mac128(lo1, hi1, yr[i], __ldg(w1 + i));
            }
            put(j, bconv_fold<W1>(lo0, hi0, p[j], mu_p[j]));
            put(j + 1, bconv_fold<W1>(lo1, hi1, p[j + 1], mu_p[j + 1]));
        }
        if (j < l_out) {
            const u64* w0 = qhat_mod_p + (size_t)j * l_in_total + in_start;
            u64 lo0 = 0, hi0 = 0;
#pragma unroll
            for (int i = 0; i < IC; ++i) mac128(lo0, hi0, yr[i], __ldg(w0 + i));
            put(j, bconv_fold<W1>(lo0, hi0, p[j], mu_p[j]));
        }
    } else {
        for (int j = 0; j < l_out; ++j) {
            u64 lo = 0, hi = 0;
            const u64* w = qhat_mod_p + (size_t)j * l_in_total + in_start;
            for (int i = 0; i < in_count; ++i) mac128(lo, hi, sy[i * BC + threadIdx.x], __ldg(w + i));
            put(j, bconv_fold<W1>(lo, hi, p[j], mu_p[j]));
        }
    }
}

}  // namespace
#include "bconv_umma.cuh"
namespace {
__global__ void bconv_reduce_kernel(const u64* __restrict__ parts, u64* __restrict__ out, size_t per_part, int nparts,
                                    size_t N, int out_start, int out_count, const u64* __restrict__ p,
                                    const ulonglong2* __restrict__ mu_p, size_t part_stride) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= per_part) return;
    const int j = out_start + (int)((idx / N) % out_count);
    u64 s = 0;
    for (int k = 0; k < nparts; ++k) s += parts[(size_t)k * part_stride + idx];      // nparts * p_j < 2^64
    out[idx] = reduce64(s, p[j], mu_p[j].y);
}
}  // namespace

extern "C" void mfhe_bconv_plan_destroy(mfhe_bconv_plan* b) {
    if (!b) return;
    cudaFree(b->d_q); cudaFree(b->d_p); cudaFree(b->d_qhat_inv); cudaFree(b->d_qhat_mod_p); cudaFree(b->d_mu_p);
    delete b;
}

extern "C" int mfhe_bconv_plan_create(int l_in, const uint64_t* q_in, int l_out, const uint64_t* p_out, mfhe_bconv_plan** out) {
    if (!out) return fail(MFHE_EINVAL, "bconv_plan_create: out is NULL");
    *out = nullptr;
    if (l_in < 1 || l_in > 128 || l_out < 1 || l_out > 128 || !q_in || !p_out) return fail(MFHE_EINVAL, "bconv_plan_create: bad sizes");
    mfhe_bconv_plan* b = new (std::nothrow) mfhe_bconv_plan;
    if (!b) return fail(MFHE_ENOMEM, "bconv_plan_create: host allocation failed");
    b->l_in = l_in; b->l_out = l_out; b->q.assign(q_in, q_in + l_in); b->p.assign(p_out, p_out + l_out);
    std::vector<host::TwPair> qi(l_in), mu(l_out);
    std::vector<uint64_t> qp((size_t)l_in * l_out);
    // lazy 128-bit accumulation needs l_in * q * p < 2^128
    long double bound = (long double)l_in;
    uint64_t qmax = 0, pmax = 0;
    for (int i = 0; i < l_in; ++i) qmax = std::max(qmax, q_in[i]);
    for (int j = 0; j < l_out; ++j) pmax = std::max(pmax, p_out[j]);
    bound *= (long double)qmax * (long double)pmax;
    if (qmax >= (1ull << 61) || pmax >= (1ull << 61) || bound >= 3.4028236692093846e38L) {
        delete b; return fail(MFHE_EUNSUPPORTED, "bconv_plan_create: l_in * q * p must stay below 2^128 and moduli below 2^61");
    }
    // the output moduli feed make_barrett1 (shifts by k - 2 and 66 - k) and the floor(2^128 / p) constants: odd, >= 2^3,
    // pairwise distinct
    for (int j = 0; j < l_out; ++j) {
        if (p_out[j] < 8 || !(p_out[j] & 1)) { delete b; return fail(MFHE_EINVAL, "bconv_plan_create: output moduli must be odd and >= 8"); }
        for (int k = 0; k < j; ++k)
            if (p_out[k] == p_out[j]) { delete b; return fail(MFHE_EINVAL, "bconv_plan_create: output moduli must be distinct"); }
    }
    for (int i = 0; i < l_in; ++i) {
        const uint64_t q = q_in[i];
        if (q < 3 || !(q & 1)) { delete b; return fail(MFHE_EINVAL, "bconv_plan_create: moduli must be odd"); }
        uint64_t pr = 1;
        for (int k = 0; k < l_in; ++k) if (k != i) pr = host::mulmod(pr, q_in[k] % q, q);
        if (!pr) { delete b; return fail(MFHE_EINVAL, "bconv_plan_create: input moduli are not coprime"); }
        const uint64_t inv = host::invmod(pr, q);
        qi[i] = host::TwPair{inv, host::shoup(inv, q)};
        for (int j = 0; j < l_out; ++j) {
            uint64_t pj = 1;
            for (int k = 0; k < l_in; ++k) if (k != i) pj = host::mulmod(pj, q_in[k] % p_out[j], p_out[j]);
            qp[(size_t)j * l_in + i] = pj;
        }
    }
    for (int j = 0; j < l_out; ++j) {
        const unsigned __int128 m = ~(unsigned __int128)0 / p_out[j];
        mu[j] = host::TwPair{(uint64_t)m, (uint64_t)(m >> 64)};
    }
    cudaError_t e;
#define UP(dst, src, bytes)                                                                 \
    if ((e = cudaMalloc((void**)&(dst), (bytes))) != cudaSuccess ||                          \
        (e = cudaMemcpy((dst), (src), (bytes), cudaMemcpyHostToDevice)) != cudaSuccess) {    \
        mfhe_bconv_plan_destroy(b); return cuda_fail(e, "bconv_plan_create upload"); }
    UP(b->d_q, b->q.data(), (size_t)l_in * 8); UP(b->d_p, b->p.data(), (size_t)l_out * 8);
    UP(b->d_qhat_inv, qi.data(), (size_t)l_in * 16); UP(b->d_qhat_mod_p, qp.data(), qp.size() * 8);
    UP(b->d_mu_p, mu.data(), (size_t)l_out * 16);
#undef UP
    *out = b;
    return MFHE_OK;
}

static int bconv_run(const mfhe_bconv_plan* b, const uint64_t* in, int in_start, int in_count, uint64_t* out, size_t N,
                     size_t batch, void* stream, const char* who, const BconvScatter* sc = nullptr) {
    if (!b || !in || (!out && !sc)) return fail(MFHE_EINVAL, std::string(who) + ": NULL argument");
    if (in_start < 0 || in_count < 1 || in_start + in_count > b->l_in) return fail(MFHE_EINVAL, std::string(who) + ": limb window out of range");
    if (!N || !batch) return MFHE_OK;
    if (batch > 65535) return fail(MFHE_EINVAL, std::string(who) + ": batch too large");
    {   // set on every call: cheap, and right for any device / thread (no process-wide one-shot flag)
        cudaError_t e = cudaSuccess;
        const void* fns[4] = {(const void*)bconv_kernel<false, false, 0>, (const void*)bconv_kernel<false, true, 0>,
                              (const void*)bconv_kernel<true, false, 0>, (const void*)bconv_kernel<true, true, 0>};
        for (const void* f : fns)
            if (e == cudaSuccess) e = cudaFuncSetAttribute(f, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
        if (e != cudaSuccess) return cuda_fail(e, who);
    }
    // one-word Barrett iff sum_i y_i c_i < in_count * qmax * p_j < 2^(k_j + 61) for every j: in_count * qmax <= 2^61
    uint64_t qmax = 0;
    for (int i = in_start; i < in_start + in_count; ++i) qmax = std::max(qmax, b->q[i]);
    static const bool force128 = getenv("MFHE_BCONV_128") != nullptr;
    static const bool generic = getenv("MFHE_BCONV_GENERIC") != nullptr;
    const bool w1 = !force128 && (unsigned __int128)qmax * (unsigned)in_count <= ((unsigned __int128)1 << 61);
    static const BconvScatter none{};
    const BconvScatter& tab = sc ? *sc : none;
    // tensor-core form (bconv_umma.cuh) whenever the coefficient count tiles by 128; MFHE_BCONV_PATH=c keeps the CUDA-core kernels
    const char* path = getenv("MFHE_BCONV_PATH");
    // It pays from 16 input limbs on (B200, N = 2^17, 16 output limbs, batch 16, ms per conversion; profiles/r2_14_bconv_umma.txt):
    //   input limbs      6      8      12     16     24     48
    //   CUDA cores     0.23   0.28   0.37   0.46   0.66   2.41 (no register-resident form beyond 24 limbs)
    //   tensor cores   0.40   0.40   0.40   0.42   0.50   0.67
    // below that the digit recombination (NDA + NDB - 1 partial sums per output) and the per-tile hand-offs cost more than the
    // few multiply-accumulates they replace.  MFHE_BCONV_PATH=t forces the tensor-core kernel for any window, =c the CUDA cores.
    const bool force_t = path && path[0] == 't';
    if (N % UB_M == 0 && !(path && path[0] == 'c') && !generic && (in_count >= 16 || force_t)) {
        uint64_t pmax = 0;
        for (uint64_t v : b->p) pmax = std::max(pmax, v);
        UbArgs a;
        a.in = (const u64*)in; a.out = (u64*)out; a.N = N; a.l_in_total = b->l_in; a.in_start = in_start; a.in_count = in_count;
        a.l_out = b->l_out; a.q = b->d_q; a.qhat_inv = b->d_qhat_inv; a.qhat_mod_p = b->d_qhat_mod_p; a.p = b->d_p; a.mu_p = b->d_mu_p;
        a.nda = ub_digits(qmax - 1); a.ndb = ub_digits(pmax - 1); a.kp = (in_count + 31) / 32 * 32;
        a.tiles_per_item = (int)(N / UB_M); a.ntiles = (int)(batch * (N / UB_M));
        const int smem = a.nda * UB_M * a.kp + a.ndb * UB_UN * a.kp + 512 + 24 * (b->l_out + in_count);
        if (smem <= 220 * 1024 && batch * (N / UB_M) < (1u << 30)) {
            int sms = 0, dev = 0;
            cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
            const int grid1 = std::min(a.ntiles, 2 * sms);
            cudaError_t e = cudaSuccess;
#define MFHE_UB_LAUNCH(SCV, W1V)                                                                                              \
            do {                                                                                                              \
                e = cudaFuncSetAttribute(bconv_umma_kernel<SCV, W1V>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);        \
                if (e == cudaSuccess) bconv_umma_kernel<SCV, W1V><<<grid1, UB_TH, smem, (cudaStream_t)stream>>>(a, tab);         \
            } while (0)
            if (sc) { if (w1) MFHE_UB_LAUNCH(true, true); else MFHE_UB_LAUNCH(true, false); }
            else    { if (w1) MFHE_UB_LAUNCH(false, true); else MFHE_UB_LAUNCH(false, false); }
#undef MFHE_UB_LAUNCH
            if (e == cudaSuccess) e = cudaGetLastError();
            return e == cudaSuccess ? MFHE_OK : cuda_fail(e, who);
        }
    }
    dim3 grid((unsigned)((N + BC - 1) / BC), (unsigned)batch);
#define MFHE_BCONV_LAUNCH(SCV, W1V, ICV)                                                                                   \
    bconv_kernel<SCV, W1V, ICV><<<grid, BC, (ICV) ? 0 : (size_t)in_count * BC * 8, (cudaStream_t)stream>>>(                  \
        (const u64*)in, (u64*)out, N, b->l_in, in_start, in_count, b->l_out, b->d_q, b->d_qhat_inv, b->d_qhat_mod_p, b->d_p,  \
        b->d_mu_p, tab)
#define MFHE_BCONV_IC(ICV)                                                                                 \
    do {                                                                                                   \
        if (sc) { if (w1) MFHE_BCONV_LAUNCH(true, true, ICV); else MFHE_BCONV_LAUNCH(true, false, ICV); }   \
        else    { if (w1) MFHE_BCONV_LAUNCH(false, true, ICV); else MFHE_BCONV_LAUNCH(false, false, ICV); } \
    } while (0)
    switch (generic ? 0 : in_count) {          // register-resident forms for the window sizes limb sharding produces
        case 2: MFHE_BCONV_IC(2); break;
        case 3: MFHE_BCONV_IC(3); break;
        case 4: MFHE_BCONV_IC(4); break;
        case 6: MFHE_BCONV_IC(6); break;
        case 8: MFHE_BCONV_IC(8); break;
        case 12: MFHE_BCONV_IC(12); break;
        case 16: MFHE_BCONV_IC(16); break;
        case 24: MFHE_BCONV_IC(24); break;
        default: MFHE_BCONV_IC(0); break;
    }
#undef MFHE_BCONV_IC
#undef MFHE_BCONV_LAUNCH
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? MFHE_OK : cuda_fail(e, who);
}
extern "C" int mfhe_bconv_plan_sizes(const mfhe_bconv_plan* b, int* l_in, int* l_out) {
    if (!b || !l_in || !l_out) return fail(MFHE_EINVAL, "bconv_plan_sizes: NULL argument");
    *l_in = b->l_in; *l_out = b->l_out;
    return MFHE_OK;
}
extern "C" int mfhe_bconv(const mfhe_bconv_plan* b, const uint64_t* in, uint64_t* out, size_t N, size_t batch, void* s) {
    return bconv_run(b, in, 0, b ? b->l_in : 0, out, N, batch, s, "bconv");
}
extern "C" int mfhe_bconv_partial(const mfhe_bconv_plan* b, const uint64_t* in, int in_start, int in_count, uint64_t* out,
                                  size_t N, size_t batch, void* s) {
    return bconv_run(b, in, in_start, in_count, out, N, batch, s, "bconv_partial");
}
extern "C" int mfhe_bconv_partial_scatter(const mfhe_bconv_plan* b, const uint64_t* in, int in_start, int in_count,
                                          uint64_t* const* dst, const size_t* dst_batch_stride, size_t N, size_t batch, void* s) {
    if (!b || !dst || !dst_batch_stride) return fail(MFHE_EINVAL, "bconv_partial_scatter: NULL argument");
    if (b->l_out > kScatterMax) return fail(MFHE_EUNSUPPORTED, "bconv_partial_scatter: more than 128 output limbs");
    BconvScatter sc{};
    for (int j = 0; j < b->l_out; ++j) {
        if (!dst[j]) return fail(MFHE_EINVAL, "bconv_partial_scatter: NULL destination");
        sc.base[j] = (u64*)dst[j]; sc.bstride[j] = dst_batch_stride[j];
    }
    return bconv_run(b, in, in_start, in_count, nullptr, N, batch, s, "bconv_partial_scatter", &sc);
}
// part_stride: words between consecutive parts (>= batch * out_count * N; a batch slice of a [nparts][batch][count][N] buffer)
extern "C" int mfhe_bconv_reduce_strided(const mfhe_bconv_plan* b, const uint64_t* parts, int nparts, size_t part_stride,
                                         int out_start, int out_count, uint64_t* out, size_t N, size_t batch, void* s) {
    if (!b || !parts || !out || nparts < 1 || out_start < 0 || out_count < 1 || out_start + out_count > b->l_out)
        return fail(MFHE_EINVAL, "bconv_reduce: bad argument");
    uint64_t pmax = 0;
    for (int j = out_start; j < out_start + out_count; ++j) pmax = std::max(pmax, b->p[j]);
    if ((unsigned __int128)pmax * (unsigned)nparts >> 64) return fail(MFHE_EINVAL, "bconv_reduce: nparts * p_j must stay below 2^64");
    const size_t per = batch * (size_t)out_count * N;
    if (!per) return MFHE_OK;
    if (part_stride < per) return fail(MFHE_EINVAL, "bconv_reduce: part stride smaller than a part");
    bconv_reduce_kernel<<<(unsigned)((per + 255) / 256), 256, 0, (cudaStream_t)s>>>((const u64*)parts, (u64*)out, per, nparts, N,
                                                                                  out_start, out_count, b->d_p, b->d_mu_p, part_stride);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? MFHE_OK : cuda_fail(e, "bconv_reduce");
}
extern "C" int mfhe_bconv_reduce(const mfhe_bconv_plan* b, const uint64_t* parts, int nparts, int out_start, int out_count,
                                 uint64_t* out, size_t N, size_t batch, void* s) {
    return mfhe_bconv_reduce_strided(b, parts, nparts, batch * (size_t)out_count * N, out_start, out_count, out, N, batch, s);
}
