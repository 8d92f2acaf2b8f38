// ntt_small.cuh -- in-register n-point transform networks (n = 2^4..2^8) shared by the batched
// small-NTT kernels (ntt_small.cu) and the fused encrypt / decrypt row kernel (scheme.cu).
// A limb-poly is owned by n/8 lanes, 8 coefficients per lane; stages cross lanes through a
// warp-private XOR-swizzled shared tile of n + 8 words.  See ntt_small.cu for the conventions.
#pragma once
#include "ntt_flat.cuh"

namespace mfhe {

template <int LOGN>
struct SmallNet {
    static constexpr int N = 1 << LOGN, TPP = N / 8, STRIDE = N + 8;
    static constexpr int FP0 = LOGN - 3;
    static constexpr int FP1 = FP0 > 3 ? FP0 - 3 : 0;
    static constexpr int FP2 = FP1 > 3 ? FP1 - 3 : 0;
    static constexpr int PL = LOGN - 3;                 // position shift of the last inverse pass
    static constexpr int IP1 = PL < 3 ? PL : 3;
    static constexpr int IP2 = PL < 6 ? PL : 6;
    static_assert(FP2 == 0 && PL <= 6, "n <= 256");

    // natural-order row g -> canonical values at bit-reversed-order positions 8u .. 8u+7
    // gstride: distance (in words) between consecutive coefficients of the source row (1 = a contiguous row; the fused
    // matrix-product decryption transforms COLUMNS of a shared-memory tile)
    __device__ static __forceinline__ void forward(u64 (&v)[8], const u64* __restrict__ g, u64* sm, u32 u,
                                                   const tw_t* __restrict__ tw, u64 q, u64 q2, u32 gstride = 1) {
        {
            const u32 e0 = deposit3(u, FP0);
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = g[(e0 + (j << FP0)) * gstride];
            fwd_pass<LOGN, FP0, 7>(v, e0, tw, q, q2);
            if constexpr (FP0 > 0) {
#pragma unroll
                for (int j = 0; j < 8; ++j) sm[swz(e0 + (j << FP0))] = v[j];
            }
        }
        if constexpr (FP0 > 0) {
            __syncwarp();
            const u32 e1 = deposit3(u, FP1);
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = sm[swz(e1 + (j << FP1))];
            fwd_pass<LOGN, FP1, (1 << (FP0 - FP1)) - 1>(v, e1, tw, q, q2);
            if constexpr (FP1 > 0) {
#pragma unroll
                for (int j = 0; j < 8; ++j) sm[swz(e1 + (j << FP1))] = v[j];
                __syncwarp();
                const u32 e2 = deposit3(u, FP2);
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] = sm[swz(e2 + (j << FP2))];
                fwd_pass<LOGN, FP2, (1 << (FP1 - FP2)) - 1>(v, e2, tw, q, q2);
            }
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) { v[j] = csub(v[j], q2); v[j] = csub(v[j], q); }
    }

    // FP64-pipe forms (q < 2^50, tables of {w, w/q}; modmath.cuh).  forward_fp leaves integer-valued doubles (bit patterns)
    // with |v| <= fp_fwd_out_bound(LOGN) q/16 at the same positions; inverse_fp takes doubles with |v| <= 2q and returns
    // canonical residues exactly like inverse().
    __device__ static __forceinline__ void forward_fp(u64 (&v)[8], const u64* __restrict__ g, u64* sm, u32 u,
                                                      const tw_t* __restrict__ tw, const fconst_t& c, u32 gstride = 1) {
        u64 (&V)[1][8] = reinterpret_cast<u64 (&)[1][8]>(v);
        {
            const u32 e0 = deposit3(u, FP0);
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = u_of(fp_load(g[(e0 + (j << FP0)) * gstride]));
            fwd_pass<LOGN, FP0, 7, 1, fconst_t>(V, e0, tw, c);
            if constexpr (FP0 > 0) {
#pragma unroll
                for (int j = 0; j < 8; ++j) sm[swz(e0 + (j << FP0))] = v[j];
            }
        }
        if constexpr (FP0 > 0) {
            __syncwarp();
            const u32 e1 = deposit3(u, FP1);
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = sm[swz(e1 + (j << FP1))];
            fwd_pass<LOGN, FP1, (1 << (FP0 - FP1)) - 1, 1, fconst_t>(V, e1, tw, c);
            if constexpr (FP1 > 0) {
#pragma unroll
                for (int j = 0; j < 8; ++j) sm[swz(e1 + (j << FP1))] = v[j];
                __syncwarp();
                const u32 e2 = deposit3(u, FP2);
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] = sm[swz(e2 + (j << FP2))];
                fwd_pass<LOGN, FP2, (1 << (FP1 - FP2)) - 1, 1, fconst_t>(V, e2, tw, c);
            }
        }
    }
    __device__ static __forceinline__ u32 inverse_fp(u64 (&v)[8], u64* sm, u32 u, const tw_t* __restrict__ itw,
                                                     const fconst_t& c, tw_t ninv, tw_t w1n) {
        u64 (&V)[1][8] = reinterpret_cast<u64 (&)[1][8]>(v);
        inv_pass<LOGN, 0, 7, 1, fconst_t>(V, 8 * u, itw, c, ninv, w1n);
        if constexpr (PL > 0) {
            __syncwarp();
#pragma unroll
            for (int j = 0; j < 8; ++j) sm[swz(8 * u + j)] = v[j];
            __syncwarp();
            const u32 e1 = deposit3(u, IP1);
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = sm[swz(e1 + (j << IP1))];
            inv_pass<LOGN, IP1, 7 & ~((1 << (3 - IP1)) - 1), 1, fconst_t>(V, e1, itw, c, ninv, w1n);
            if constexpr (PL > 3) {
#pragma unroll
                for (int j = 0; j < 8; ++j) sm[swz(e1 + (j << IP1))] = v[j];
                __syncwarp();
                const u32 e2 = deposit3(u, IP2);
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] = sm[swz(e2 + (j << IP2))];
                inv_pass<LOGN, IP2, 7 & ~((1 << (IP1 + 3 - IP2)) - 1), 1, fconst_t>(V, e2, itw, c, ninv, w1n);
                return e2;
            }
            return e1;
        }
        return 8 * u;
    }

    // the value inverse() returns, known before the transform (lets the caller prefetch what it combines with)
    __device__ static __forceinline__ u32 out_base(u32 u) {
        if constexpr (PL > 3) return deposit3(u, IP2);
        else if constexpr (PL > 0) return deposit3(u, IP1);
        else return 8 * u;
    }
    // values at network positions 8u .. 8u+7 (< 2q) -> canonical natural-order coefficients;
    // on return v[j] is the coefficient at position  e + (j << PL)  where e is the return value.
    __device__ static __forceinline__ u32 inverse(u64 (&v)[8], u64* sm, u32 u, const tw_t* __restrict__ itw, u64 q,
                                                  u64 q2, tw_t ninv, tw_t w1n) {
        inv_pass<LOGN, 0, 7>(v, 8 * u, itw, q, q2, ninv, w1n);
        if constexpr (PL > 0) {
            __syncwarp();
#pragma unroll
            for (int j = 0; j < 8; ++j) sm[swz(8 * u + j)] = v[j];
            __syncwarp();
            const u32 e1 = deposit3(u, IP1);
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = sm[swz(e1 + (j << IP1))];
            inv_pass<LOGN, IP1, 7 & ~((1 << (3 - IP1)) - 1)>(v, e1, itw, q, q2, ninv, w1n);
            if constexpr (PL > 3) {
#pragma unroll
                for (int j = 0; j < 8; ++j) sm[swz(e1 + (j << IP1))] = v[j];
                __syncwarp();
                const u32 e2 = deposit3(u, IP2);
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] = sm[swz(e2 + (j << IP2))];
                inv_pass<LOGN, IP2, 7 & ~((1 << (IP1 + 3 - IP2)) - 1)>(v, e2, itw, q, q2, ninv, w1n);
                return e2;
            }
            return e1;
        }
        return 8 * u;
    }
};

}  // namespace mfhe
