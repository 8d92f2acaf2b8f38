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
//     approximated -- its cross terms come from two DFMAs on the FP64 pipe, leaving 3 IMAD.WIDE
//     + 4 IMAD on the binding fmaheavy pipe, result < 4q -- and x is folded once per two or
//     three forward stages (modmath.cuh) instead of in every butterfly;
//   * every thread works on the same positions of PB polys of the batch, so each twiddle
//     is fetched and prepared once per PB butterflies;
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
    int arith;                // 0: integer arithmetic, tw = {w, w'};  1: FP64-pipe arithmetic (q < 2^50), tw = {w, w/q} as doubles
};

// ---------------------------------------------------------------------------
// One in-register radix-8 pass over three consecutive bits [P+2..P] of the
// sub-transform index, applied to PB polys of the same limb at once: the
// twiddle of a butterfly depends only on its position, so it is fetched (one
// LDG.128) and prepared for the FP64-assisted quotient (fpq_prep) ONCE and used
// for the PB polys.  v[p][j] holds the element of poly p whose triple equals j.
//   LOGN  size of the whole transform
//   GB    global bit position of triple bit 0
//   ACT   bit mask of the triple bits that are transformed in this pass
//   g0    global coefficient index of v[.][0]
// Stage for global bit b: s = LOGN-1-b, twiddle index 2^s + (g >> (b+1)).
// ---------------------------------------------------------------------------
template <class QC> struct is_fp_arith { static constexpr bool value = false; };
template <> struct is_fp_arith<fconst_t> { static constexpr bool value = true; };

// QC = qconst_t: integer arithmetic (any q < 2^60); QC = fconst_t: FP64-pipe arithmetic (q < 2^50, tables of {w, w/q},
// v holds the bit patterns of doubles) -- same index structure, same twiddle fetches
template <int LOGN, int GB, int ACT, int PB, class QC>
__device__ __forceinline__ void fwd_pass(u64 (&v)[PB][8], u32 g0, const tw_t* __restrict__ tw,
                                         const QC& c) {
    constexpr bool FP = is_fp_arith<QC>::value;
#pragma unroll
    for (int tb = 2; tb >= 0; --tb) {
        if (!((ACT >> tb) & 1)) continue;
        const int b = GB + tb;
        const int s = LOGN - 1 - b;
        const u32 base = (1u << s) + (g0 >> (b + 1));
#pragma unroll
        for (int jh = 0; jh < (4 >> tb); ++jh) {
            const tw_t raw = __ldg(tw + base + jh);
            if constexpr (FP) {
#pragma unroll
                for (int p = 0; p < PB; ++p) {
#pragma unroll
                    for (int jl = 0; jl < (1 << tb); ++jl) {
                        const int jx = (jh << (tb + 1)) | jl;
                        if (tb == 2) ct_bfly_fp<LOGN, LOGN - 1 - (GB + 2)>(v[p][jx], v[p][jx | 4], raw, c);
                        if (tb == 1) ct_bfly_fp<LOGN, LOGN - 1 - (GB + 1)>(v[p][jx], v[p][jx | 2], raw, c);
                        if (tb == 0) ct_bfly_fp<LOGN, LOGN - 1 - (GB + 0)>(v[p][jx], v[p][jx | 1], raw, c);
                    }
                }
            } else {
            twf_t w;
            if (s != LOGN - 1) w = fpq_prep(raw);
#pragma unroll
            for (int p = 0; p < PB; ++p) {
#pragma unroll
                for (int jl = 0; jl < (1 << tb); ++jl) {
                    const int jx = (jh << (tb + 1)) | jl;
                    // stage number must be a constant expression for the bound schedule
                    if (tb == 2) ct_bfly<LOGN, LOGN - 1 - (GB + 2)>(v[p][jx], v[p][jx | 4], w, raw, c);
                    if (tb == 1) ct_bfly<LOGN, LOGN - 1 - (GB + 1)>(v[p][jx], v[p][jx | 2], w, raw, c);
                    if (tb == 0) ct_bfly<LOGN, LOGN - 1 - (GB + 0)>(v[p][jx], v[p][jx | 1], w, raw, c);
                }
            }
            }
        }
    }
}

template <int LOGN, int GB, int ACT, int PB, class QC>
__device__ __forceinline__ void inv_pass(u64 (&v)[PB][8], u32 g0, const tw_t* __restrict__ itw,
                                         const QC& c, tw_t ninv, tw_t w1n) {
    constexpr bool FP = is_fp_arith<QC>::value;
#pragma unroll
    for (int tb = 0; tb <= 2; ++tb) {
        if (!((ACT >> tb) & 1)) continue;
        const int b = GB + tb;
        const int s = LOGN - 1 - b;
        const u32 base = (1u << s) + (g0 >> (b + 1));
#pragma unroll
        for (int jh = 0; jh < (4 >> tb); ++jh) {
            if constexpr (FP) {
                tw_t raw = {0, 0};
                if (b != LOGN - 1) raw = __ldg(itw + base + jh);
#pragma unroll
                for (int p = 0; p < PB; ++p) {
#pragma unroll
                    for (int jl = 0; jl < (1 << tb); ++jl) {
                        const int jx = (jh << (tb + 1)) | jl;
                        const int jy = jx | (1 << tb);
                        if (b == LOGN - 1) gs_last_fp<LOGN>(v[p][jx], v[p][jy], ninv, w1n, c);
                        else if (tb == 0) gs_bfly_fp<LOGN, GB + 0>(v[p][jx], v[p][jy], raw, c);
                        else if (tb == 1) gs_bfly_fp<LOGN, GB + 1>(v[p][jx], v[p][jy], raw, c);
                        else              gs_bfly_fp<LOGN, GB + 2>(v[p][jx], v[p][jy], raw, c);
                    }
                }
            } else {
            twf_t w;
            if (b != LOGN - 1) w = fpq_prep(__ldg(itw + base + jh));
#pragma unroll
            for (int p = 0; p < PB; ++p) {
#pragma unroll
                for (int jl = 0; jl < (1 << tb); ++jl) {
                    const int jx = (jh << (tb + 1)) | jl;
                    const int jy = jx | (1 << tb);
                    if (b == LOGN - 1) {        // last stage: exact products, N^-1 folded in, canonical
                        const u64 t = v[p][jx] + c.q4 - v[p][jy], sum = v[p][jx] + v[p][jy];
                        // approximate quotient here too (3 wide + 2 high products instead of the ~7 wide ones of the
                        // exact 64 x 64 high word): result < 4q for any 64-bit operand, two conditional subtracts
                        // canonicalise (cuobjdump, ntt_cols_kernel<16,0,2>: IMAD.WIDE 269 -> 221)
                        v[p][jx] = csub(csub(mul_shoup_lazy4(sum, ninv, c.nq), c.q2), c.q);
                        v[p][jy] = csub(csub(mul_shoup_lazy4(t, w1n, c.nq), c.q2), c.q);
                    } else {
                        gs_bfly(v[p][jx], v[p][jy], w, c);
                    }
                }
            }
            }
        }
    }
}

// single-poly forms used by the n-point networks (ntt_small.cuh, scheme.cu)
template <int LOGN, int GB, int ACT>
__device__ __forceinline__ void fwd_pass(u64 (&v)[8], u32 g0, const tw_t* __restrict__ tw, u64 q, u64 /*q2*/) {
    const qconst_t c = make_qconst(q);
    fwd_pass<LOGN, GB, ACT, 1, qconst_t>(reinterpret_cast<u64 (&)[1][8]>(v), g0, tw, c);
}
template <int LOGN, int GB, int ACT>
__device__ __forceinline__ void inv_pass(u64 (&v)[8], u32 g0, const tw_t* __restrict__ itw, u64 q, u64 /*q2*/,
                                         tw_t ninv, tw_t w1n) {
    const qconst_t c = make_qconst(q);
    inv_pass<LOGN, GB, ACT, 1, qconst_t>(reinterpret_cast<u64 (&)[1][8]>(v), g0, itw, c, ninv, w1n);
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
