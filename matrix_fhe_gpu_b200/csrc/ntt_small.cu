// ntt_small.cu -- batched n-point transforms, n = 2^4..2^8, for the matrix pipeline's X axis.
//
// Replaces, with ONE launch for all batch_count*limbs limb-polys,
//   src/core/ntt_core.cu:443-460  xy_ntt_forward/backward_phantom (a host loop of
//       batch_count launches of ph/src/ntt/ntt_1d.cu:17-76 / :194-256), and
//   src/core/ntt_core.cu:394-431, 462-481  custom_ntt_* / xy_ntt_*_gl (copy + twist +
//       bit-reverse + log2(n) global-memory stages + scale, ~10 passes over the data).
// A limb-poly is owned by n/8 lanes (8 lanes for n = 64, four limb-polys per warp); the
// coefficients stay in registers across stages and cross lanes through a warp-private
// swizzled shared tile; global traffic is one read and one write per coefficient.
//
// All three conventions run the same merged Cooley-Tukey / Gentleman-Sande network: a node
// of the product tree X^m - c splits into X^(m/2) -+ sqrt(c) and the twiddle of the node is
// sqrt(c).  Only the host-built table (root constant -1, i or 1) and the output order differ:
//   kPhantomNegacyclic  X^n + 1, natural in -> bit-reversed out  (Phantom order)
//   kGlXnMinusI         X^n - i, natural in -> natural out: A[k] = sum a_j psi^j w^(jk)
//   kCyclicNatural      X^n - 1, natural in -> natural out (custom_ntt_forward alone)
#include "ntt_api.h"
#include "ntt_small.cuh"

namespace mfhe {

constexpr int kSmallThreads = 256;

template <int LOGN> __device__ __forceinline__ u32 bitrev_n(u32 x) { return __brev(x) >> (32 - LOGN); }

// AR = 1: butterflies on the FP64 pipe (every modulus below 2^50, T carries the {w, w/q} tables; modmath.cuh) -- same words
template <int LOGN, bool NATURAL, int AR>
__global__ void __launch_bounds__(kSmallThreads)
ntt_small_fwd_kernel(u64* __restrict__ data, SmallTables T, int limbs, size_t limb_polys) {
    using Net = SmallNet<LOGN>;
    constexpr int TPP = Net::TPP, PPB = kSmallThreads / TPP, STRIDE = Net::STRIDE;
    __shared__ __align__(16) u64 sm_all[PPB * STRIDE];
    const u32 u = threadIdx.x % TPP, pib = threadIdx.x / TPP;
    const size_t lp = (size_t)blockIdx.x * PPB + pib;
    if (lp >= limb_polys) return;                       // whole polys drop out together (TPP <= 32)
    const u32 limb = lp % limbs;
    const u64 q = T.moduli[limb], q2 = 2 * q;
    u64* g = data + (lp << LOGN);
    u64* sm = sm_all + pib * STRIDE;
    u64 v[8];
    if constexpr (AR) {
        const fconst_t fc = make_fconst(q);
        Net::forward_fp(v, g, sm, u, T.tw + ((size_t)limb << LOGN), fc);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = fp_store(fp_fold(d_of(v[j]), fc), fc);
    } else {
        Net::forward(v, g, sm, u, T.tw + ((size_t)limb << LOGN), q, q2);
    }
    // the thread now holds positions 8u .. 8u+7 of the bit-reversed-order result
    if constexpr (NATURAL) {
        __syncwarp();
#pragma unroll
        for (int j = 0; j < 8; ++j) sm[swz(bitrev_n<LOGN>(8 * u + j))] = v[j];
        __syncwarp();
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const ulonglong2 t = *reinterpret_cast<const ulonglong2*>(sm + swz(8 * u + 2 * k));
            v[2 * k] = t.x; v[2 * k + 1] = t.y;
        }
    }
    asm volatile("st.global.v4.u64 [%0], {%1,%2,%3,%4};" ::"l"(g + 8 * u), "l"(v[0]), "l"(v[1]), "l"(v[2]), "l"(v[3]) : "memory");
    asm volatile("st.global.v4.u64 [%0], {%1,%2,%3,%4};" ::"l"(g + 8 * u + 4), "l"(v[4]), "l"(v[5]), "l"(v[6]), "l"(v[7]) : "memory");
}

template <int LOGN, bool NATURAL, int AR>
__global__ void __launch_bounds__(kSmallThreads)
ntt_small_inv_kernel(u64* __restrict__ data, SmallTables T, int limbs, size_t limb_polys) {
    using Net = SmallNet<LOGN>;
    constexpr int TPP = Net::TPP, PPB = kSmallThreads / TPP, STRIDE = Net::STRIDE;
    __shared__ __align__(16) u64 sm_all[PPB * STRIDE];
    const u32 u = threadIdx.x % TPP, pib = threadIdx.x / TPP;
    const size_t lp = (size_t)blockIdx.x * PPB + pib;
    if (lp >= limb_polys) return;
    const u32 limb = lp % limbs;
    const u64 q = T.moduli[limb], q2 = 2 * q;
    u64* g = data + (lp << LOGN);
    u64* sm = sm_all + pib * STRIDE;
    u64 v[8];

    asm volatile("ld.global.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(v[0]), "=l"(v[1]), "=l"(v[2]), "=l"(v[3]) : "l"(g + 8 * u));
    asm volatile("ld.global.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(v[4]), "=l"(v[5]), "=l"(v[6]), "=l"(v[7]) : "l"(g + 8 * u + 4));
    if constexpr (NATURAL) {       // natural-order evaluations -> network (bit-reversed) positions
#pragma unroll
        for (int j = 0; j < 8; ++j) sm[swz(bitrev_n<LOGN>(8 * u + j))] = v[j];
        __syncwarp();
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const ulonglong2 t = *reinterpret_cast<const ulonglong2*>(sm + swz(8 * u + 2 * k));
            v[2 * k] = t.x; v[2 * k + 1] = t.y;
        }
    }
    u32 e;
    if constexpr (AR) {
        const fconst_t fc = make_fconst(q);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = u_of(fp_load(v[j]));
        e = Net::inverse_fp(v, sm, u, T.itw + ((size_t)limb << LOGN), fc, T.ninv[limb], T.w1n[limb]);
    } else {
        e = Net::inverse(v, sm, u, T.itw + ((size_t)limb << LOGN), q, q2, T.ninv[limb], T.w1n[limb]);
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) g[e + (j << Net::PL)] = v[j];
}

template <int LOGN>
static cudaError_t launch_small(bool fwd, bool natural, u64* data, const SmallTables& T, int limbs,
                                size_t limb_polys, cudaStream_t st) {
    constexpr int PPB = kSmallThreads / ((1 << LOGN) / 8);
    const unsigned grid = (unsigned)((limb_polys + PPB - 1) / PPB);
#define MFHE_SMALL_LAUNCH(K, NAT)                                                                       \
    do {                                                                                                \
        if (T.arith) K<LOGN, NAT, 1><<<grid, kSmallThreads, 0, st>>>(data, T, limbs, limb_polys);         \
        else         K<LOGN, NAT, 0><<<grid, kSmallThreads, 0, st>>>(data, T, limbs, limb_polys);         \
    } while (0)
    if (fwd) {
        if (natural) MFHE_SMALL_LAUNCH(ntt_small_fwd_kernel, true); else MFHE_SMALL_LAUNCH(ntt_small_fwd_kernel, false);
    } else {
        if (natural) MFHE_SMALL_LAUNCH(ntt_small_inv_kernel, true); else MFHE_SMALL_LAUNCH(ntt_small_inv_kernel, false);
    }
#undef MFHE_SMALL_LAUNCH
    return cudaGetLastError();
}

cudaError_t ntt_small_launch(bool forward, u64* data, const SmallTables& T, int limbs, size_t batch,
                             cudaStream_t st) {
    const size_t limb_polys = batch * (size_t)limbs;
    if (limb_polys == 0) return cudaSuccess;
    const bool natural = T.convention != kPhantomNegacyclic;
    switch (T.logn) {
        case 4: return launch_small<4>(forward, natural, data, T, limbs, limb_polys, st);
        case 5: return launch_small<5>(forward, natural, data, T, limbs, limb_polys, st);
        case 6: return launch_small<6>(forward, natural, data, T, limbs, limb_polys, st);
        case 7: return launch_small<7>(forward, natural, data, T, limbs, limb_polys, st);
        case 8: return launch_small<8>(forward, natural, data, T, limbs, limb_polys, st);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace mfhe
