// capi_modswitch.cu -- mod-up / mod-down with an auxiliary base P for the flat family.
//
// The reference reserves the base (P_MODULI / P_NUM_LIMBS = 3, include/core/config.h:22,46-52) and never uses it; Phantom
// has the operations fused into its key-switching (ph/src/ntt/ntt_modup.cu, ntt_moddown.cu; ph/src/rns_bconv.cu:530-830).
// Here they are compositions of this library's own kernels -- flat NTT on a limb window, base conversion with the
// scatter form (one destination per output limb, so the converted limbs land interleaved per polynomial without a
// pack pass) -- plus one fused subtract-and-scale pass:
//   mod-up    [batch][LQ][N] coefficients over Q  ->  [batch][LQ+LP][N] NTT form over Q u P      (x_P = bConv_{Q->P}(x))
//   mod-down  [batch][LQ+LP][N] NTT form over Q u P  ->  [batch][LQ][N] NTT form over Q:
//             out_i = (x_i - NTT(bConv_{P->Q}(INTT(x_P)))_i) * P^-1 mod q_i  =  (x - [x]_P) / P  up to the base
//             conversion's multiple-of-P slack (|error| < LP, the approximate mod-down of RNS-CKKS / Phantom).
#include <mutex>
#include <string>
#include <vector>

#include "capi_common.h"
#include "plan.h"

using namespace mfhe;

namespace {
struct ScaleConsts { u64 inv[64]; };

// out[b][i][k] = (x[b][i][k] - out[b][i][k]) * inv[i] mod q_i; x has xs limbs per poly, out has lq
__global__ void __launch_bounds__(256)
moddown_combine_kernel(const u64* __restrict__ x, u64* __restrict__ out, const u64* __restrict__ moduli,
                       const ulonglong2* __restrict__ mu_tab, int logn, int lq, int xs, size_t total,
                       const __grid_constant__ ScaleConsts k) {
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const size_t poly = idx >> logn, i = poly % lq, b = poly / lq, c = idx & (((size_t)1 << logn) - 1);
        const u64 q = moduli[i]; const ulonglong2 mu = mu_tab[i];
        const u64 v = x[((b * xs + i) << logn) + c];
        out[idx] = mul_barrett(sub_mod(v, out[idx], q), k.inv[i], q, mu.x, mu.y);
    }
}
}  // namespace

extern "C" int mfhe_modup(const mfhe_ntt_plan* qp, const mfhe_bconv_plan* q_to_p, const uint64_t* d_x, uint64_t* d_out,
                          size_t batch, void* stream) {
    if (!qp || !q_to_p || ((!d_x || !d_out) && batch)) return fail(MFHE_EINVAL, "modup: NULL argument");
    if (batch == 0) return MFHE_OK;
    int lq = 0, lp = 0;
    if (mfhe_bconv_plan_sizes(q_to_p, &lq, &lp) != MFHE_OK || lq + lp != qp->L || lq + lp > 128)
        return fail(MFHE_EINVAL, "modup: the plan's limbs must be the conversion's input base followed by its output base");
    const size_t N = (size_t)1 << qp->logn;
    cudaStream_t st = (cudaStream_t)stream;
    cudaError_t e = cudaMemcpy2DAsync(d_out, (size_t)(lq + lp) * N * 8, d_x, (size_t)lq * N * 8, (size_t)lq * N * 8, batch,
                                      cudaMemcpyDeviceToDevice, st);
    if (e != cudaSuccess) return cuda_fail(e, "modup copy");
    std::vector<uint64_t*> dst(lp); std::vector<size_t> stride(lp, (size_t)(lq + lp) * N);
    for (int j = 0; j < lp; ++j) dst[j] = d_out + (size_t)(lq + j) * N;
    int rc = mfhe_bconv_partial_scatter(q_to_p, d_x, 0, lq, dst.data(), stride.data(), N, batch, stream);
    if (rc != MFHE_OK) return rc;
    return mfhe_ntt_forward(qp, d_out, batch, stream);
}

extern "C" int mfhe_moddown(const mfhe_ntt_plan* qp, const mfhe_bconv_plan* p_to_q, const uint64_t* d_x, uint64_t* d_out,
                            uint64_t* d_scratch, size_t batch, void* stream) {
    if (!qp || !p_to_q || ((!d_x || !d_out || !d_scratch) && batch)) return fail(MFHE_EINVAL, "moddown: NULL argument");
    if (batch == 0) return MFHE_OK;
    int lp = 0, lq = 0;
    if (mfhe_bconv_plan_sizes(p_to_q, &lp, &lq) != MFHE_OK || lq + lp != qp->L || lq > 64)
        return fail(MFHE_EINVAL, "moddown: the plan's limbs must be the conversion's output base followed by its input base");
    const size_t N = (size_t)1 << qp->logn;
    cudaStream_t st = (cudaStream_t)stream;
    // P part -> packed scratch, back to coefficients
    cudaError_t e = cudaMemcpy2DAsync(d_scratch, (size_t)lp * N * 8, d_x + (size_t)lq * N, (size_t)(lq + lp) * N * 8,
                                      (size_t)lp * N * 8, batch, cudaMemcpyDeviceToDevice, st);
    if (e != cudaSuccess) return cuda_fail(e, "moddown copy");
    int rc = mfhe_ntt_inverse_limbs(qp, d_scratch, batch, lq, lp, stream);
    if (rc == MFHE_OK) rc = mfhe_bconv(p_to_q, d_scratch, d_out, N, batch, stream);
    if (rc == MFHE_OK) rc = mfhe_ntt_forward_limbs(qp, d_out, batch, 0, lq, stream);
    if (rc != MFHE_OK) return rc;
    ScaleConsts k{};
    for (int i = 0; i < lq; ++i) {
        const u64 q = qp->moduli[i];
        u64 pm = 1;
        for (int j = 0; j < lp; ++j) pm = host::mulmod(pm, qp->moduli[lq + j] % q, q);
        k.inv[i] = host::invmod(pm, q);
    }
    const size_t total = (batch * (size_t)lq) << qp->logn;
    const unsigned grid = (unsigned)std::min<size_t>((total + 255) / 256, 148 * 16);
    moddown_combine_kernel<<<grid, 256, 0, st>>>((const u64*)d_x, (u64*)d_out, qp->d_moduli, qp->d_mu, qp->logn, lq, lq + lp, total, k);
    e = cudaGetLastError();
    return e == cudaSuccess ? MFHE_OK : cuda_fail(e, "moddown");
}

// ---- mod-down without any allocation by anybody -------------------------------------------------------------------
// No counterpart in the reference: added for callers that cannot call cudaMalloc
// (CUDA graph capture, tight serving loops, harnesses that forbid it).  mfhe_moddown needs batch * LP * N words of scratch
// from its caller; here that scratch is a file-scope device array of a fixed, modest size that every GPU which loads
// the library holds until the process exits.
//   * limit: batch * LP * N <= MFHE_MODDOWN_STATIC_WORDS (2^21 words = 16 MiB; e.g. N = 2^16, LP = 4, batch = 8);
//   * in HOST code `g_moddown_scratch` is only a stand-in for the symbol, not a device address: the address is resolved
//     with cudaGetSymbolAddress once per device, with that device current, and that pointer is what the kernels get;
//   * every call shares the one array and it keeps the last call's contents.  Nothing here needs it zeroed: the region a
//     call uses (its first batch * LP * N words) is overwritten by the copy of the P limbs before anything reads it;
//   * calls that overlap on the GPU (different streams, or different host threads without ordering) RACE on the array:
//     serialise them on one stream or with events.
// The first call on a device resolves the symbol address (it may load the module); make that call before a graph capture.
namespace {
constexpr size_t kStaticWords = MFHE_MODDOWN_STATIC_WORDS;
static __device__ __align__(256) unsigned long long g_moddown_scratch[kStaticWords];
constexpr int kMaxDevices = 64;
uint64_t* g_scratch_ptr[kMaxDevices] = {};            // per device: resolved address of g_moddown_scratch
std::mutex g_scratch_mu;

int static_scratch(uint64_t** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return cuda_fail(e, "moddown_static: cudaGetDevice");
    if (dev < 0 || dev >= kMaxDevices) return fail(MFHE_EUNSUPPORTED, "moddown_static: device ordinal beyond 63");
    std::lock_guard<std::mutex> lock(g_scratch_mu);
    if (!g_scratch_ptr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_moddown_scratch);          // this device is current
        if (e != cudaSuccess || !p) return cuda_fail(e, "moddown_static: cudaGetSymbolAddress");
        g_scratch_ptr[dev] = (uint64_t*)p;
    }
    *out = g_scratch_ptr[dev];
    return MFHE_OK;
}
}  // namespace

extern "C" int mfhe_moddown_static(const mfhe_ntt_plan* qp, const mfhe_bconv_plan* p_to_q, const uint64_t* d_x,
                                   uint64_t* d_out, size_t batch, void* stream) {
    if (!qp || !p_to_q || ((!d_x || !d_out) && batch)) return fail(MFHE_EINVAL, "moddown_static: NULL argument");
    if (batch == 0) return MFHE_OK;
    int lp = 0, lq = 0;
    if (mfhe_bconv_plan_sizes(p_to_q, &lp, &lq) != MFHE_OK || lq + lp != qp->L)
        return fail(MFHE_EINVAL, "moddown_static: the plan's limbs must be the conversion's output base followed by its input base");
    const size_t N = (size_t)1 << qp->logn;
    if (batch > kStaticWords / ((size_t)lp * N))
        return fail(MFHE_EINVAL, "moddown_static: batch * LP * N = " + std::to_string(batch * (size_t)lp * N) +
                                 " words exceeds the static scratch of " + std::to_string(kStaticWords) +
                                 " words (16 MiB); use mfhe_moddown with caller scratch");
    uint64_t* scratch = nullptr;
    const int rc = static_scratch(&scratch);
    if (rc != MFHE_OK) return rc;
    return mfhe_moddown(qp, p_to_q, d_x, d_out, scratch, batch, stream);
}
