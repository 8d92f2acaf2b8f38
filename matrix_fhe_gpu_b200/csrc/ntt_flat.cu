// ntt_flat.cu -- kernels + launchers for the flat RNS negacyclic NTT (see ntt_flat.cuh).
#include "ntt_flat.cuh"
#include "ntt_api.h"

#ifndef MFHE_PB
#define MFHE_PB 2
#endif
// resident CTAs per SM asked of ptxas: the kernels are latency sensitive, A/B runs on B200 (profiles/r1_06_ntt_ab.txt)
// gave 2 / 4 as the best trade between occupancy and spills at PB = 2
#ifndef MFHE_COL_MINB
#define MFHE_COL_MINB 2
#endif
#ifndef MFHE_ROW_MINB
#define MFHE_ROW_MINB 4
#endif
// the FP64-pipe kernels (AR = 1) have their own tuning: cheaper butterflies leave them between the pipe and HBM
#ifndef MFHE_PB_FP
#define MFHE_PB_FP 2
#endif
#ifndef MFHE_COL_MINB_FP
#define MFHE_COL_MINB_FP 2
#endif
#ifndef MFHE_ROW_MINB_FP
#define MFHE_ROW_MINB_FP 4
#endif
// order of the (poly group, limb) blocks along gridDim.y, bit AR: 1 = limb-major (blocks that share a twiddle table run
// together).  Measured on the config-2 shape at 50 bits (profiles/r2_09_ntt_fp64_ab.txt): +1.7 % for the FP64-pipe kernels.
#ifndef MFHE_LIMB_MAJOR
#define MFHE_LIMB_MAJOR 2
#endif
namespace mfhe {

template <int AR>
__device__ __forceinline__ void block_coords(u32 by, u32 limb_count, u32& li, u32& pg) {
    if constexpr ((MFHE_LIMB_MAJOR >> AR) & 1) {
        const u32 groups = gridDim.y / limb_count;
        li = by / groups; pg = by % groups;
    } else {
        li = by % limb_count; pg = by / limb_count;
    }
}

constexpr int kColThreads = 512;   // column kernel: 4096-element tile per poly, 8 per thread
constexpr int kRowWarps = 8;       // row kernel: 8 warps = 8 rows of 256 per CTA (x PB polys)

// A CTA works on the same tile / rows of PB consecutive polys of the batch (same limb): poly p of
// the group lives `pstride` words after poly 0.
//
// ===========================================================================
// Column kernel: R-point sub-transforms at stride 256, R = N/256.
// CTA tile = R rows x COLS columns (R*COLS = 4096 elements = 32 KiB) per poly, thread
// (c, u): column c, 8 rows selected by u.  Lanes run along c, so every global
// and shared access of a warp touches whole 128..256-byte row segments.
// ===========================================================================
// FP64-pipe arithmetic (QC = fconst_t): residues become doubles where a transform reads its input (forward: this kernel's
// first pass) and canonical residues again where it writes its result (inverse: the last stage, inside inv_pass); what
// travels between the two kernels of a transform are the doubles' bit patterns.
template <int LOGN, bool FWD, int P, int ACT, bool FIRST, bool LAST, int PB, class QC>
__device__ __forceinline__ void col_pass(u64* __restrict__ gptr, size_t pstride, u64* __restrict__ sm, u32 u,
                                         u32 c, const tw_t* __restrict__ tw, const QC& qc,
                                         tw_t ninv, tw_t w1n) {
    constexpr int LOGR = LOGN - 8;
    constexpr int COLS = 4096 >> LOGR;
    u64 v[PB][8];
    const u32 row0 = deposit3(u, P);
#pragma unroll
    for (int p = 0; p < PB; ++p) {
        if constexpr (FIRST) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                v[p][j] = gptr[p * pstride + ((size_t)(row0 + (j << P)) << 8)];
                if constexpr (is_fp_arith<QC>::value && FWD) v[p][j] = u_of(fp_load(v[p][j]));
            }
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) v[p][j] = sm[p * 4096 + (row0 + (j << P)) * COLS + c];
        }
    }
    if constexpr (FWD) fwd_pass<LOGN, 8 + P, ACT, PB, QC>(v, row0 << 8, tw, qc);
    else               inv_pass<LOGN, 8 + P, ACT, PB, QC>(v, row0 << 8, tw, qc, ninv, w1n);
#pragma unroll
    for (int p = 0; p < PB; ++p) {
        if constexpr (LAST) {
#pragma unroll
            for (int j = 0; j < 8; ++j) gptr[p * pstride + ((size_t)(row0 + (j << P)) << 8)] = v[p][j];
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) sm[p * 4096 + (row0 + (j << P)) * COLS + c] = v[p][j];
        }
    }
    if constexpr (!LAST) __syncthreads();
}

template <int LOGN, bool FWD, int PB, int AR>
__global__ void __launch_bounds__(kColThreads, AR ? MFHE_COL_MINB_FP : MFHE_COL_MINB)
ntt_cols_kernel(u64* __restrict__ data, NttTables T, int limb_start, int limb_count) {
    typedef typename arith_const<AR>::type QC;
    constexpr int LOGR = LOGN - 8;
    constexpr int LOGC = 12 - LOGR;
    constexpr int COLS = 1 << LOGC;
    extern __shared__ __align__(16) u64 sm[];            // PB * 4096 words
    const u32 c = threadIdx.x & (COLS - 1);
    const u32 u = threadIdx.x >> LOGC;
    u32 li, pg;
    block_coords<AR>(blockIdx.y, (u32)limb_count, li, pg);
    const u32 limb = limb_start + li;
    const QC qc = arith_const<AR>::make(T.moduli[limb]);
    const tw_t* tw = (FWD ? T.tw : T.itw) + ((size_t)limb << LOGN);
    tw_t ninv = {0, 0}, w1n = {0, 0};
    if constexpr (!FWD) { ninv = T.ninv[limb]; w1n = T.w1n[limb]; }
    const size_t pstride = (size_t)limb_count << LOGN;
    u64* g = data + (((size_t)pg * PB * limb_count + li) << LOGN) + blockIdx.x * COLS + c;

    if constexpr (FWD) {
        // triples from the top bit down: P = LOGR-3, then steps of 3, last pass at P = 0
        constexpr int P0 = LOGR - 3;
        constexpr int P1 = P0 > 3 ? P0 - 3 : 0;
        constexpr int P2 = P1 > 3 ? P1 - 3 : 0;
        col_pass<LOGN, true, P0, 7, true, P0 == 0, PB, QC>(g, pstride, sm, u, c, tw, qc, ninv, w1n);
        if constexpr (P0 > 0)
            col_pass<LOGN, true, P1, (1 << (P0 - P1)) - 1, false, P1 == 0, PB, QC>(g, pstride, sm, u, c, tw, qc, ninv, w1n);
        if constexpr (P1 > 0)
            col_pass<LOGN, true, P2, (1 << (P1 - P2)) - 1, false, true, PB, QC>(g, pstride, sm, u, c, tw, qc, ninv, w1n);
        static_assert(P2 == 0, "LOGR <= 9 supported");
    } else {
        // triples from bit 0 up: P = 0, 3, ... last pass at P = LOGR-3
        constexpr int PL = LOGR - 3;
        constexpr int P1 = PL < 3 ? PL : 3;
        constexpr int P2 = PL < 6 ? PL : 6;
        col_pass<LOGN, false, 0, 7, true, PL == 0, PB, QC>(g, pstride, sm, u, c, tw, qc, ninv, w1n);
        if constexpr (PL > 0)
            col_pass<LOGN, false, P1, 7 & ~((1 << (3 - P1)) - 1), false, P1 == PL, PB, QC>(g, pstride, sm, u, c, tw, qc, ninv, w1n);
        if constexpr (PL > 3)
            col_pass<LOGN, false, P2, 7 & ~((1 << (P1 + 3 - P2)) - 1), false, true, PB, QC>(g, pstride, sm, u, c, tw, qc, ninv, w1n);
        static_assert(PL <= 6, "LOGR <= 9 supported");
    }
}

// ===========================================================================
// Row kernel: one warp per contiguous 256-element row (of PB polys), 8 elements per lane per poly.
// ===========================================================================
__device__ __forceinline__ void st256(u64* p, u64 a, u64 b, u64 c, u64 d) {
    asm volatile("st.global.v4.u64 [%0], {%1,%2,%3,%4};" ::"l"(p), "l"(a), "l"(b), "l"(c), "l"(d) : "memory");
}
__device__ __forceinline__ void ld256(const u64* p, u64& a, u64& b, u64& c, u64& d) {
    asm volatile("ld.global.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(a), "=l"(b), "=l"(c), "=l"(d) : "l"(p));
}

template <int LOGN, int PB, int AR>
__global__ void __launch_bounds__(kRowWarps * 32, AR ? MFHE_ROW_MINB_FP : MFHE_ROW_MINB)
ntt_rows_fwd_kernel(u64* __restrict__ data, NttTables T, int limb_start, int limb_count) {
    typedef typename arith_const<AR>::type QC;
    __shared__ __align__(16) u64 sm_all[kRowWarps * PB * 256];
    const u32 lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    u64* sm = sm_all + warp * (PB * 256);
    u32 li, pg;
    block_coords<AR>(blockIdx.y, (u32)limb_count, li, pg);
    const u32 limb = limb_start + li;
    const u32 row = blockIdx.x * kRowWarps + warp;
    const QC qc = arith_const<AR>::make(T.moduli[limb]);
    const tw_t* tw = T.tw + ((size_t)limb << LOGN);
    const size_t pstride = (size_t)limb_count << LOGN;
    u64* g = data + (((size_t)pg * PB * limb_count + li) << LOGN) + ((size_t)row << 8);
    const u32 gbase = row << 8;
    u64 v[PB][8];

    // pass A: column bits 7,6,5 -- element c = j*32 + lane (coalesced 256 B per load)
#pragma unroll
    for (int p = 0; p < PB; ++p)
#pragma unroll
        for (int j = 0; j < 8; ++j) v[p][j] = g[p * pstride + (j << 5) + lane];
    fwd_pass<LOGN, 5, 7, PB, QC>(v, gbase | lane, tw, qc);
    u32 off[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) off[j] = swz((j << 5) + lane);
#pragma unroll
    for (int p = 0; p < PB; ++p)
#pragma unroll
        for (int j = 0; j < 8; ++j) sm[p * 256 + off[j]] = v[p][j];
    __syncwarp();

    // pass B: bits 4,3,2 -- c = (lane>>2)*32 + j*4 + (lane&3)
    const u32 cb = ((lane >> 2) << 5) | (lane & 3);
#pragma unroll
    for (int j = 0; j < 8; ++j) off[j] = swz(cb + (j << 2));
#pragma unroll
    for (int p = 0; p < PB; ++p)
#pragma unroll
        for (int j = 0; j < 8; ++j) v[p][j] = sm[p * 256 + off[j]];
    fwd_pass<LOGN, 2, 7, PB, QC>(v, gbase | cb, tw, qc);
#pragma unroll
    for (int p = 0; p < PB; ++p)
#pragma unroll
        for (int j = 0; j < 8; ++j) sm[p * 256 + off[j]] = v[p][j];
    __syncwarp();

    // pass C: bits 1,0 (bit 2 passive) -- c = lane*8 + j, 16-byte shared loads
#pragma unroll
    for (int k = 0; k < 4; ++k) off[k] = swz((lane << 3) + 2 * k);
#pragma unroll
    for (int p = 0; p < PB; ++p)
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const ulonglong2 t = *reinterpret_cast<const ulonglong2*>(sm + p * 256 + off[k]);
            v[p][2 * k] = t.x; v[p][2 * k + 1] = t.y;
        }
    fwd_pass<LOGN, 0, 3, PB, QC>(v, gbase | (lane << 3), tw, qc);
#pragma unroll
    for (int p = 0; p < PB; ++p) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            if constexpr (AR) v[p][j] = fp_store(fp_fold(d_of(v[p][j]), qc), qc);
            else { v[p][j] = csub(v[p][j], qc.q2); v[p][j] = csub(v[p][j], qc.q); }
        }
        st256(g + p * pstride + (lane << 3), v[p][0], v[p][1], v[p][2], v[p][3]);
        st256(g + p * pstride + (lane << 3) + 4, v[p][4], v[p][5], v[p][6], v[p][7]);
    }
}

// MUL: the pointwise product with `mul` (same layout, NTT form, canonical) is applied to the loaded values, so that
// NTT . pointwise . INTT needs no pass of its own for the product (mfhe_ntt_mul_intt)
template <int LOGN, int PB, bool MUL, int AR>
__global__ void __launch_bounds__(kRowWarps * 32, AR ? MFHE_ROW_MINB_FP : MFHE_ROW_MINB)
ntt_rows_inv_kernel(u64* __restrict__ data, NttTables T, int limb_start, int limb_count,
                    const u64* __restrict__ mul, const ulonglong2* __restrict__ mu) {
    typedef typename arith_const<AR>::type QC;
    __shared__ __align__(16) u64 sm_all[kRowWarps * PB * 256];
    const u32 lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    u64* sm = sm_all + warp * (PB * 256);
    u32 li, pg;
    block_coords<AR>(blockIdx.y, (u32)limb_count, li, pg);
    const u32 limb = limb_start + li;
    const u32 row = blockIdx.x * kRowWarps + warp;
    const QC qc = arith_const<AR>::make(T.moduli[limb]);
    const tw_t* itw = T.itw + ((size_t)limb << LOGN);
    const tw_t ninv = T.ninv[limb], w1n = T.w1n[limb];
    const size_t pstride = (size_t)limb_count << LOGN;
    u64* g = data + (((size_t)pg * PB * limb_count + li) << LOGN) + ((size_t)row << 8);
    const u32 gbase = row << 8;
    u64 v[PB][8];
    u32 off[8];

    // pass A: bits 0,1,2 -- c = lane*8 + j, 256-bit global loads
#pragma unroll
    for (int p = 0; p < PB; ++p) {
        ld256(g + p * pstride + (lane << 3), v[p][0], v[p][1], v[p][2], v[p][3]);
        ld256(g + p * pstride + (lane << 3) + 4, v[p][4], v[p][5], v[p][6], v[p][7]);
    }
    if constexpr (MUL) {
        const ulonglong2 m = mu[limb];
        const barrett1_t bc = make_barrett1(T.moduli[limb], m.x, m.y);
        const u64* gm = mul + (g - data);
#pragma unroll
        for (int p = 0; p < PB; ++p) {
            u64 b[8];
            ld256(gm + p * pstride + (lane << 3), b[0], b[1], b[2], b[3]);
            ld256(gm + p * pstride + (lane << 3) + 4, b[4], b[5], b[6], b[7]);
#pragma unroll
            for (int j = 0; j < 8; ++j) v[p][j] = mul_barrett1(v[p][j], b[j], bc);
        }
    }
    if constexpr (AR) {
#pragma unroll
        for (int p = 0; p < PB; ++p)
#pragma unroll
            for (int j = 0; j < 8; ++j) v[p][j] = u_of(fp_load(v[p][j]));
    }
    inv_pass<LOGN, 0, 7, PB, QC>(v, gbase | (lane << 3), itw, qc, ninv, w1n);
#pragma unroll
    for (int k = 0; k < 4; ++k) off[k] = swz((lane << 3) + 2 * k);
#pragma unroll
    for (int p = 0; p < PB; ++p)
#pragma unroll
        for (int k = 0; k < 4; ++k)
            *reinterpret_cast<ulonglong2*>(sm + p * 256 + off[k]) = make_ulonglong2(v[p][2 * k], v[p][2 * k + 1]);
    __syncwarp();

    // pass B: bits 3,4,5 -- c = (lane>>3)*64 + j*8 + (lane&7)
    const u32 cb = ((lane >> 3) << 6) | (lane & 7);
#pragma unroll
    for (int j = 0; j < 8; ++j) off[j] = swz(cb + (j << 3));
#pragma unroll
    for (int p = 0; p < PB; ++p)
#pragma unroll
        for (int j = 0; j < 8; ++j) v[p][j] = sm[p * 256 + off[j]];
    inv_pass<LOGN, 3, 7, PB, QC>(v, gbase | cb, itw, qc, ninv, w1n);
#pragma unroll
    for (int p = 0; p < PB; ++p)
#pragma unroll
        for (int j = 0; j < 8; ++j) sm[p * 256 + off[j]] = v[p][j];
    __syncwarp();

    // pass C: bits 6,7 (bit 5 passive) -- c = j*32 + lane
#pragma unroll
    for (int j = 0; j < 8; ++j) off[j] = swz((j << 5) + lane);
#pragma unroll
    for (int p = 0; p < PB; ++p)
#pragma unroll
        for (int j = 0; j < 8; ++j) v[p][j] = sm[p * 256 + off[j]];
    inv_pass<LOGN, 5, 6, PB, QC>(v, gbase | lane, itw, qc, ninv, w1n);
#pragma unroll
    for (int p = 0; p < PB; ++p)
#pragma unroll
        for (int j = 0; j < 8; ++j) g[p * pstride + (j << 5) + lane] = v[p][j];
}

// ---------------------------------------------------------------------------
template <int LOGN, int PB, int AR>
static cudaError_t launch_pair(bool fwd, u64* data, const NttTables& T, int limb_start, int limb_count,
                               size_t polys, cudaStream_t st, const u64* mul, const ulonglong2* mu) {
    constexpr int LOGR = LOGN - 8;
    constexpr size_t col_smem = (size_t)PB * 4096 * sizeof(u64);
    {   // set on every call: cheap, and right for any device / thread (no process-wide one-shot flag)
        cudaFuncSetAttribute(ntt_cols_kernel<LOGN, true, PB, AR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)col_smem);
        cudaFuncSetAttribute(ntt_cols_kernel<LOGN, false, PB, AR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)col_smem);
    }
    const unsigned gy = (unsigned)(polys / PB) * (unsigned)limb_count;
    dim3 gc(256 / (4096 >> LOGR), gy);
    dim3 gr((1 << LOGR) / kRowWarps, gy);
    if (fwd) {
        ntt_cols_kernel<LOGN, true, PB, AR><<<gc, kColThreads, col_smem, st>>>(data, T, limb_start, limb_count);
        ntt_rows_fwd_kernel<LOGN, PB, AR><<<gr, kRowWarps * 32, 0, st>>>(data, T, limb_start, limb_count);
    } else {
        if (mul) ntt_rows_inv_kernel<LOGN, PB, true, AR><<<gr, kRowWarps * 32, 0, st>>>(data, T, limb_start, limb_count, mul, mu);
        else     ntt_rows_inv_kernel<LOGN, PB, false, AR><<<gr, kRowWarps * 32, 0, st>>>(data, T, limb_start, limb_count, nullptr, nullptr);
        ntt_cols_kernel<LOGN, false, PB, AR><<<gc, kColThreads, col_smem, st>>>(data, T, limb_start, limb_count);
    }
    return cudaGetLastError();
}

constexpr int kPB = MFHE_PB;       // polys per thread group (twiddle reuse factor)

template <int LOGN>
static cudaError_t launch_polys(bool fwd, u64* data, const NttTables& T, int limb_start, int limb_count,
                                size_t polys, cudaStream_t st, const u64* mul, const ulonglong2* mu) {
    // whole PB-groups first, then the remainder one poly at a time
    const size_t pb = T.arith ? (size_t)MFHE_PB_FP : (size_t)kPB;
    const size_t main = polys - polys % pb;
    cudaError_t err = cudaSuccess;
    if (main) err = T.arith ? launch_pair<LOGN, MFHE_PB_FP, 1>(fwd, data, T, limb_start, limb_count, main, st, mul, mu)
                            : launch_pair<LOGN, kPB, 0>(fwd, data, T, limb_start, limb_count, main, st, mul, mu);
    if (err == cudaSuccess && polys > main) {
        const size_t off = (main * (size_t)limb_count) << LOGN;
        err = T.arith ? launch_pair<LOGN, 1, 1>(fwd, data + off, T, limb_start, limb_count, polys - main, st, mul ? mul + off : nullptr, mu)
                      : launch_pair<LOGN, 1, 0>(fwd, data + off, T, limb_start, limb_count, polys - main, st, mul ? mul + off : nullptr, mu);
    }
    return err;
}

// {w, w'} -> {w centred as a double, w / q}: the twiddle format of the FP64-pipe arithmetic (per = entries per limb)
__global__ void fp_tables_kernel(const tw_t* __restrict__ src, tw_t* __restrict__ dst, const u64* __restrict__ moduli,
                                 size_t per, size_t total) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const u64 q = moduli[i / per], w = src[i].x;
    const long long wc = w > (q >> 1) ? (long long)w - (long long)q : (long long)w;
    const double wd = (double)wc;                                  // exact: |wc| < 2^49
    dst[i] = tw_t{u_of(wd), u_of(__ddiv_rn(wd, (double)(long long)q))};
}
cudaError_t ntt_fp_tables_launch(const tw_t* src, tw_t* dst, const u64* moduli, size_t per, int L, cudaStream_t st) {
    const size_t total = per * (size_t)L;
    fp_tables_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(src, dst, moduli, per, total);
    return cudaGetLastError();
}

// The batch is processed in chunks of whole polys so that gridDim.y stays below 65536 and
// (optionally, chunk_limb_polys != 0) so that a chunk's footprint stays L2-resident between the
// column and the row kernel.
cudaError_t ntt_flat_launch(bool forward, u64* data, const NttTables& T, int limb_start,
                            int limb_count, size_t batch, size_t chunk_limb_polys, cudaStream_t st,
                            const u64* mul, const ulonglong2* mu) {
    if (mul && (forward || !mu)) return cudaErrorInvalidValue;
    if (limb_count <= 0) return cudaErrorInvalidValue;
    size_t chunk = chunk_limb_polys ? chunk_limb_polys / (size_t)limb_count : batch;
    if (chunk == 0) chunk = 1;
    const size_t cap = 65535 / (size_t)limb_count;           // polys per launch (gridDim.y limit, PB = 1 worst case)
    if (cap == 0) return cudaErrorInvalidValue;
    if (chunk > cap) chunk = cap;
    if (chunk > kPB) chunk -= chunk % kPB;
    cudaError_t err = cudaSuccess;
    for (size_t off = 0; off < batch && err == cudaSuccess; off += chunk) {
        const size_t cnt = (batch - off < chunk) ? batch - off : chunk;
        const size_t woff = (off * (size_t)limb_count) << T.logn;
        u64* p = data + woff;
        const u64* pm = mul ? mul + woff : nullptr;
#define MFHE_CASE(LN) case LN: err = launch_polys<LN>(forward, p, T, limb_start, limb_count, cnt, st, pm, mu); break;
        switch (T.logn) {
            MFHE_CASE(12) MFHE_CASE(13) MFHE_CASE(14) MFHE_CASE(15) MFHE_CASE(16) MFHE_CASE(17)
            default: return cudaErrorInvalidValue;
        }
#undef MFHE_CASE
    }
    return err;
}

}  // namespace mfhe
