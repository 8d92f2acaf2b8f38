// ntt_flat.cu -- kernels + launchers for the flat RNS negacyclic NTT (see ntt_flat.cuh).
#include "ntt_flat.cuh"
#include "ntt_api.h"

namespace mfhe {

constexpr int kColThreads = 512;   // column kernel: 4096-element tile, 8 per thread
constexpr int kRowWarps = 8;       // row kernel: 8 warps = 8 rows of 256 per CTA

// ===========================================================================
// Column kernel: R-point sub-transforms at stride 256, R = N/256.
// CTA tile = R rows x COLS columns (R*COLS = 4096 elements = 32 KiB), thread
// (c, u): column c, 8 rows selected by u.  Lanes run along c, so every global
// and shared access of a warp touches whole 128..256-byte row segments.
// ===========================================================================
template <int LOGN, bool FWD, int P, int ACT, bool FIRST, bool LAST>
__device__ __forceinline__ void col_pass(u64* __restrict__ gptr, u64* __restrict__ sm, u32 u,
                                         u32 c, const tw_t* __restrict__ tw, u64 q, u64 q2,
                                         tw_t ninv, tw_t w1n) {
    constexpr int LOGR = LOGN - 8;
    constexpr int COLS = 4096 >> LOGR;
    u64 v[8];
    const u32 row0 = deposit3(u, P);
    if constexpr (FIRST) {
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = gptr[(size_t)(row0 + (j << P)) << 8];
    } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = sm[(row0 + (j << P)) * COLS + c];
    }
    if constexpr (FWD) fwd_pass<LOGN, 8 + P, ACT>(v, row0 << 8, tw, q, q2);
    else               inv_pass<LOGN, 8 + P, ACT>(v, row0 << 8, tw, q, q2, ninv, w1n);
    if constexpr (LAST) {
#pragma unroll
        for (int j = 0; j < 8; ++j) gptr[(size_t)(row0 + (j << P)) << 8] = v[j];
    } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) sm[(row0 + (j << P)) * COLS + c] = v[j];
        __syncthreads();
    }
}

template <int LOGN, bool FWD>
__global__ void __launch_bounds__(kColThreads)
ntt_cols_kernel(u64* __restrict__ data, NttTables T, int limb_start, int limb_count) {
    constexpr int LOGR = LOGN - 8;
    constexpr int LOGC = 12 - LOGR;
    constexpr int COLS = 1 << LOGC;
    __shared__ u64 sm[4096];
    const u32 c = threadIdx.x & (COLS - 1);
    const u32 u = threadIdx.x >> LOGC;
    const u32 lp = blockIdx.y;
    const u32 limb = limb_start + lp % limb_count;
    const u64 q = T.moduli[limb], q2 = 2 * q;
    const tw_t* tw = (FWD ? T.tw : T.itw) + ((size_t)limb << LOGN);
    tw_t ninv = {0, 0}, w1n = {0, 0};
    if constexpr (!FWD) { ninv = T.ninv[limb]; w1n = T.w1n[limb]; }
    u64* g = data + ((size_t)lp << LOGN) + blockIdx.x * COLS + c;

    if constexpr (FWD) {
        // triples from the top bit down: P = LOGR-3, then steps of 3, last pass at P = 0
        constexpr int P0 = LOGR - 3;
        constexpr int P1 = P0 > 3 ? P0 - 3 : 0;
        constexpr int P2 = P1 > 3 ? P1 - 3 : 0;
        col_pass<LOGN, true, P0, 7, true, P0 == 0>(g, sm, u, c, tw, q, q2, ninv, w1n);
        if constexpr (P0 > 0)
            col_pass<LOGN, true, P1, (1 << (P0 - P1)) - 1, false, P1 == 0>(g, sm, u, c, tw, q, q2, ninv, w1n);
        if constexpr (P1 > 0)
            col_pass<LOGN, true, P2, (1 << (P1 - P2)) - 1, false, true>(g, sm, u, c, tw, q, q2, ninv, w1n);
        static_assert(P2 == 0, "LOGR <= 9 supported");
    } else {
        // triples from bit 0 up: P = 0, 3, ... last pass at P = LOGR-3
        constexpr int PL = LOGR - 3;
        constexpr int P1 = PL < 3 ? PL : 3;
        constexpr int P2 = PL < 6 ? PL : 6;
        col_pass<LOGN, false, 0, 7, true, PL == 0>(g, sm, u, c, tw, q, q2, ninv, w1n);
        if constexpr (PL > 0)
            col_pass<LOGN, false, P1, 7 & ~((1 << (3 - P1)) - 1), false, P1 == PL>(g, sm, u, c, tw, q, q2, ninv, w1n);
        if constexpr (PL > 3)
            col_pass<LOGN, false, P2, 7 & ~((1 << (P1 + 3 - P2)) - 1), false, true>(g, sm, u, c, tw, q, q2, ninv, w1n);
        static_assert(PL <= 6, "LOGR <= 9 supported");
    }
}

// ===========================================================================
// Row kernel: one warp per contiguous 256-element row, 8 elements per lane.
// ===========================================================================
__device__ __forceinline__ void st256(u64* p, u64 a, u64 b, u64 c, u64 d) {
    asm volatile("st.global.v4.u64 [%0], {%1,%2,%3,%4};" ::"l"(p), "l"(a), "l"(b), "l"(c), "l"(d) : "memory");
}
__device__ __forceinline__ void ld256(const u64* p, u64& a, u64& b, u64& c, u64& d) {
    asm volatile("ld.global.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(a), "=l"(b), "=l"(c), "=l"(d) : "l"(p));
}

template <int LOGN>
__global__ void __launch_bounds__(kRowWarps * 32)
ntt_rows_fwd_kernel(u64* __restrict__ data, NttTables T, int limb_start, int limb_count) {
    __shared__ __align__(16) u64 sm_all[kRowWarps * 256];
    const u32 lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    u64* sm = sm_all + warp * 256;
    const u32 lp = blockIdx.y;
    const u32 limb = limb_start + lp % limb_count;
    const u32 row = blockIdx.x * kRowWarps + warp;
    const u64 q = T.moduli[limb], q2 = 2 * q;
    const tw_t* tw = T.tw + ((size_t)limb << LOGN);
    u64* g = data + ((size_t)lp << LOGN) + ((size_t)row << 8);
    const u32 gbase = row << 8;
    u64 v[8];

    // pass A: column bits 7,6,5 -- element c = j*32 + lane (coalesced 256 B per load)
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = g[(j << 5) + lane];
    fwd_pass<LOGN, 5, 7>(v, gbase | lane, tw, q, q2);
#pragma unroll
    for (int j = 0; j < 8; ++j) sm[swz((j << 5) + lane)] = v[j];
    __syncwarp();

    // pass B: bits 4,3,2 -- c = (lane>>2)*32 + j*4 + (lane&3)
    const u32 cb = ((lane >> 2) << 5) | (lane & 3);
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = sm[swz(cb + (j << 2))];
    fwd_pass<LOGN, 2, 7>(v, gbase | cb, tw, q, q2);
#pragma unroll
    for (int j = 0; j < 8; ++j) sm[swz(cb + (j << 2))] = v[j];
    __syncwarp();

    // pass C: bits 1,0 (bit 2 passive) -- c = lane*8 + j, 16-byte shared loads
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const ulonglong2 t = *reinterpret_cast<const ulonglong2*>(sm + swz((lane << 3) + 2 * k));
        v[2 * k] = t.x; v[2 * k + 1] = t.y;
    }
    fwd_pass<LOGN, 0, 3>(v, gbase | (lane << 3), tw, q, q2);
#pragma unroll
    for (int j = 0; j < 8; ++j) { v[j] = csub(v[j], q2); v[j] = csub(v[j], q); }
    st256(g + (lane << 3), v[0], v[1], v[2], v[3]);
    st256(g + (lane << 3) + 4, v[4], v[5], v[6], v[7]);
}

template <int LOGN>
__global__ void __launch_bounds__(kRowWarps * 32)
ntt_rows_inv_kernel(u64* __restrict__ data, NttTables T, int limb_start, int limb_count) {
    __shared__ __align__(16) u64 sm_all[kRowWarps * 256];
    const u32 lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    u64* sm = sm_all + warp * 256;
    const u32 lp = blockIdx.y;
    const u32 limb = limb_start + lp % limb_count;
    const u32 row = blockIdx.x * kRowWarps + warp;
    const u64 q = T.moduli[limb], q2 = 2 * q;
    const tw_t* itw = T.itw + ((size_t)limb << LOGN);
    const tw_t ninv = T.ninv[limb], w1n = T.w1n[limb];
    u64* g = data + ((size_t)lp << LOGN) + ((size_t)row << 8);
    const u32 gbase = row << 8;
    u64 v[8];

    // pass A: bits 0,1,2 -- c = lane*8 + j, 256-bit global loads
    ld256(g + (lane << 3), v[0], v[1], v[2], v[3]);
    ld256(g + (lane << 3) + 4, v[4], v[5], v[6], v[7]);
    inv_pass<LOGN, 0, 7>(v, gbase | (lane << 3), itw, q, q2, ninv, w1n);
#pragma unroll
    for (int k = 0; k < 4; ++k)
        *reinterpret_cast<ulonglong2*>(sm + swz((lane << 3) + 2 * k)) = make_ulonglong2(v[2 * k], v[2 * k + 1]);
    __syncwarp();

    // pass B: bits 3,4,5 -- c = (lane>>3)*64 + j*8 + (lane&7)
    const u32 cb = ((lane >> 3) << 6) | (lane & 7);
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = sm[swz(cb + (j << 3))];
    inv_pass<LOGN, 3, 7>(v, gbase | cb, itw, q, q2, ninv, w1n);
#pragma unroll
    for (int j = 0; j < 8; ++j) sm[swz(cb + (j << 3))] = v[j];
    __syncwarp();

    // pass C: bits 6,7 (bit 5 passive) -- c = j*32 + lane
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = sm[swz((j << 5) + lane)];
    inv_pass<LOGN, 5, 6>(v, gbase | lane, itw, q, q2, ninv, w1n);
    if constexpr (LOGN == 8) { /* single-kernel size: values already canonical */ }
#pragma unroll
    for (int j = 0; j < 8; ++j) g[(j << 5) + lane] = v[j];
}

// ---------------------------------------------------------------------------
template <int LOGN>
static cudaError_t launch_fwd(u64* data, const NttTables& T, int limb_start, int limb_count,
                              size_t limb_polys, cudaStream_t st) {
    constexpr int LOGR = LOGN - 8;
    dim3 gc(256 / (4096 >> LOGR), (unsigned)limb_polys);
    ntt_cols_kernel<LOGN, true><<<gc, kColThreads, 0, st>>>(data, T, limb_start, limb_count);
    dim3 gr((1 << LOGR) / kRowWarps, (unsigned)limb_polys);
    ntt_rows_fwd_kernel<LOGN><<<gr, kRowWarps * 32, 0, st>>>(data, T, limb_start, limb_count);
    return cudaGetLastError();
}
template <int LOGN>
static cudaError_t launch_inv(u64* data, const NttTables& T, int limb_start, int limb_count,
                              size_t limb_polys, cudaStream_t st) {
    constexpr int LOGR = LOGN - 8;
    dim3 gr((1 << LOGR) / kRowWarps, (unsigned)limb_polys);
    ntt_rows_inv_kernel<LOGN><<<gr, kRowWarps * 32, 0, st>>>(data, T, limb_start, limb_count);
    dim3 gc(256 / (4096 >> LOGR), (unsigned)limb_polys);
    ntt_cols_kernel<LOGN, false><<<gc, kColThreads, 0, st>>>(data, T, limb_start, limb_count);
    return cudaGetLastError();
}

// limb_polys is chunked so that gridDim.y stays below 65536 and (optionally) so that a
// chunk's footprint stays L2-resident between the two kernels (chunk_limb_polys).
cudaError_t ntt_flat_launch(bool forward, u64* data, const NttTables& T, int limb_start,
                            int limb_count, size_t batch, size_t chunk_limb_polys, cudaStream_t st) {
    const size_t total = batch * (size_t)limb_count;
    size_t chunk = chunk_limb_polys ? chunk_limb_polys : total;
    // a chunk must be a whole number of polys (so that lp % limb_count is the limb) unless
    // it is smaller than one poly, in which case it must divide limb_count.
    if (chunk >= (size_t)limb_count) chunk -= chunk % limb_count;
    else while (limb_count % chunk) --chunk;
    if (chunk > 65535) { chunk = 65535 - 65535 % limb_count; if (!chunk) chunk = 1; }
    cudaError_t err = cudaSuccess;
    for (size_t off = 0; off < total && err == cudaSuccess; off += chunk) {
        const size_t cnt = (total - off < chunk) ? total - off : chunk;
        u64* p = data + (off << T.logn);
        const int ls = limb_start + (int)(off % limb_count);
        // when the chunk is a slice of one poly, limbs run ls .. ls+cnt-1
        const int lc = (chunk >= (size_t)limb_count) ? limb_count : (int)cnt;
        const int lstart = (chunk >= (size_t)limb_count) ? limb_start : ls;
#define MFHE_CASE(LN)                                                               \
        case LN: err = forward ? launch_fwd<LN>(p, T, lstart, lc, cnt, st)          \
                               : launch_inv<LN>(p, T, lstart, lc, cnt, st); break;
        switch (T.logn) {
            MFHE_CASE(12) MFHE_CASE(13) MFHE_CASE(14) MFHE_CASE(15) MFHE_CASE(16) MFHE_CASE(17)
            default: return cudaErrorInvalidValue;
        }
#undef MFHE_CASE
    }
    return err;
}

}  // namespace mfhe
