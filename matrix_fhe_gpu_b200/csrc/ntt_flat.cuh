// ntt_flat.cuh -- flat RNS negacyclic NTT / INTT over [batch][L][N] u64, N = 2^12..2^17.
//
// Computes the same map, with the same table convention and output ordering, as
//   ph/src/ntt/fntt_2d.cu:620-653  nwt_2d_radix8_forward_inplace
//   ph/src/ntt/intt_2d.cu:724-757  nwt_2d_radix8_backward_inplace
// (forward: natural in -> bit-reversed out, canonical; inverse: the exact inverse
// including N^-1), but it is a different program:
//   * one launch covers every limb-poly of the batch (Phantom: one call per poly);
//   * a limb-poly is viewed as an R x 256 matrix, N = R*256.  The "column" kernel
//     runs the R-point sub-transforms of a 4096-element tile out of shared memory,
//     the "row" kernel gives every warp one contiguous 256-element row: the row
//     lives in that warp's registers (8 per lane) and crosses lanes through a
//     warp-private, XOR-swizzled shared-memory tile, so the row kernel needs no
//     block-level barrier at all;
//   * twiddles are {w, w'} pairs fetched with one LDG.128 (rows: LDG.256 for two);
//   * lazy arithmetic uses the 4 spare bits of q < 2^60: the Shoup quotient is only
//     approximated (3 wide multiplies instead of 6, result < 4q) and x is folded once
//     per two or three forward stages (modmath.cuh) instead of in every butterfly;
//   * results are stored with 256-bit stores.
#pragma once
#include "modmath.cuh"

namespace mfhe {

struct NttTables {            // device pointers, per plan
    const u64*  moduli;       // [L]
    const tw_t* tw;           // [L][N] forward, Phantom index order tw[m + i]
    const tw_t* itw;          // [L][N] inverse, raw (n^-1 NOT folded)
    const tw_t* ninv;         // [L]    {n^-1, shoup}
    const tw_t* w1n;          // [L]    {itw[1] * n^-1, shoup}: last inverse stage
    int logn;
    int L;
};

// ---------------------------------------------------------------------------
// One in-register radix-8 pass over three consecutive bits [P+2..P] of the
// sub-transform index.  v[j] holds the element whose triple equals j.
//   LOGN  size of the whole transform
//   GB    global bit position of triple bit 0
//   ACT   bit mask of the triple bits that are transformed in this pass
//   g0    global coefficient index of v[0]
// Stage for global bit b: s = LOGN-1-b, twiddle index 2^s + (g >> (b+1)).
// ---------------------------------------------------------------------------
template <int LOGN, int GB, int ACT>
__device__ __forceinline__ void fwd_pass(u64 (&v)[8], u32 g0, const tw_t* __restrict__ tw,
                                         u64 q, u64 q2) {
#pragma unroll
    for (int tb = 2; tb >= 0; --tb) {
        if (!((ACT >> tb) & 1)) continue;
        const int b = GB + tb;
        const int s = LOGN - 1 - b;
        const u32 base = (1u << s) + (g0 >> (b + 1));
#pragma unroll
        for (int jh = 0; jh < (4 >> tb); ++jh) {
            const tw_t w = __ldg(tw + base + jh);
#pragma unroll
            for (int jl = 0; jl < (1 << tb); ++jl) {
                const int jx = (jh << (tb + 1)) | jl;
                // stage number must be a constant expression for the bound schedule
                if (tb == 2) ct_bfly<LOGN, LOGN - 1 - (GB + 2)>(v[jx], v[jx | 4], w, q, q2);
                if (tb == 1) ct_bfly<LOGN, LOGN - 1 - (GB + 1)>(v[jx], v[jx | 2], w, q, q2);
                if (tb == 0) ct_bfly<LOGN, LOGN - 1 - (GB + 0)>(v[jx], v[jx | 1], w, q, q2);
            }
        }
        (void)s;
    }
}

template <int LOGN, int GB, int ACT>
__device__ __forceinline__ void inv_pass(u64 (&v)[8], u32 g0, const tw_t* __restrict__ itw,
                                         u64 q, u64 q2, tw_t ninv, tw_t w1n) {
#pragma unroll
    for (int tb = 0; tb <= 2; ++tb) {
        if (!((ACT >> tb) & 1)) continue;
        const int b = GB + tb;
        const int s = LOGN - 1 - b;
        const u32 base = (1u << s) + (g0 >> (b + 1));
#pragma unroll
        for (int jh = 0; jh < (4 >> tb); ++jh) {
            tw_t w = (b == LOGN - 1) ? w1n : __ldg(itw + base + jh);
#pragma unroll
            for (int jl = 0; jl < (1 << tb); ++jl) {
                const int jx = (jh << (tb + 1)) | jl;
                const int jy = jx | (1 << tb);
                if (b == LOGN - 1) {            // last stage: exact products, N^-1 folded in, canonical
                    const u64 t = v[jx] + 2 * q2 - v[jy], sum = v[jx] + v[jy];
                    v[jx] = mul_shoup(sum, ninv, q);
                    v[jy] = mul_shoup(t, w, q);
                } else {
                    gs_bfly(v[jx], v[jy], w, q, q2);
                }
            }
        }
    }
}

// position of element j of a thread whose remaining index bits are u, triple at P
__device__ __forceinline__ u32 deposit3(u32 u, int P) {
    return ((u >> P) << (P + 3)) | (u & ((1u << P) - 1u));
}

// XOR swizzle of a warp-private 256-element row (16-byte chunk granularity): keeps
// the three access patterns of the row kernel bank-conflict free.
__device__ __forceinline__ u32 swz(u32 e) {
    u32 chunk = e >> 1;
    return ((chunk ^ ((chunk >> 3) & 7u)) << 1) | (e & 1u);
}

}  // namespace mfhe
