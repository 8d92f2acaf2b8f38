// trace_tc.cu -- the trace step's complex modular GEMM  C = n * A * B'^T  (src/core/trace.cu:63-130,
// batched_trace.cu:85-161 of the reference: one thread per output, four `unsigned __int128 %` per complex MAC)
// on the tensor cores, n = 64, exact.
//
// Per (lane, limb) the complex product is ONE real GEMM with K = 2n:
//     [Cr | Ci] (n x 2n)  =  [Ar | Ai] (n x 2n)  x  [ Br  q-Bi ; Bi  Br ]^T,
// the minus sign of Ai*Bi being carried by the residue q - Bi, so every operand is a non-negative residue.  As in
// wcrt_tc.cu both operands are cut into ND = ceil(bits(q)/8) unsigned 8-bit digits, every digit pair (u, v) is a
// u8 x u8 -> s32 GEMM (2n * 255^2 * ND < 2^27: no overflow) on mma.sync.m16n8k32 accumulating into position
// u + v, and the epilogue recombines the 2ND-1 partial sums into a 128-bit integer, multiplies by n (a shift)
// and applies ONE Barrett reduction: the same canonical residue as the reference's add_mod / mul_mod chain.
//
// CTA = one (lane, limb): A' digits [ND][64][128 B] stay in shared memory, the B' digits are staged in two halves
// (the Cr columns, then the Ci columns); 8 warps of 16 x 16 outputs, two column sub-passes per half.  Row stride
// 144 B keeps the fragment loads bank-conflict free.  Shared memory 2 * ND * 9 KiB -> two CTAs per SM.
#include "ctx.h"
#include "modmath.cuh"

namespace mfhe {

namespace {
constexpr int TN = 64, TK = 2 * TN, TRS = TK + 16, TTH = 256;
struct LimbList { int cnt; int limbs; unsigned char l[64]; };

__device__ __forceinline__ void mma_u8_t(int (&c)[4], const u32 (&a)[4], u32 b0, u32 b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// 64 x 64 residues (row-major, K contiguous) -> digit planes [ND][64][TRS] at K byte offset koff; NEG: q - x
// TR: the source is K-major (src[k][row], the encrypted matrix product contracts over the SLOW index of the
// ciphertext matrices): a thread still reads four consecutive words of the source, which are now four rows of one k,
// and scatters their digits with byte stores
template <int ND, bool NEG, bool TR>
__device__ __forceinline__ void stage_digits(unsigned char* dst, const u64* __restrict__ src, int koff, u64 q) {
    if constexpr (TR) {
        // lanes run along the rows (the source's contiguous axis: 256-byte segments per warp and k), a thread gathers the
        // four consecutive k of one row and stores them as one 32-bit word per digit plane
#pragma unroll
        for (int r = 0; r < TN * TN / 4 / TTH; ++r) {
            const int row = threadIdx.x & (TN - 1), kq = (threadIdx.x >> 6) + (TTH / TN) * r;
            u64 x[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) { x[j] = src[(4 * kq + j) * TN + row]; if (NEG) x[j] = q - x[j]; }
            unsigned char* d = dst + row * TRS + koff + kq * 4;
#pragma unroll
            for (int v = 0; v < ND; ++v) {
                const int sh = v < 4 ? 0 : 32;
                const u32 sel = (u32)(v & 3) | ((u32)(4 + (v & 3)) << 4);
                const u32 t01 = __byte_perm((u32)(x[0] >> sh), (u32)(x[1] >> sh), sel);
                const u32 t23 = __byte_perm((u32)(x[2] >> sh), (u32)(x[3] >> sh), sel);
                *reinterpret_cast<u32*>(d + v * TN * TRS) = __byte_perm(t01, t23, 0x5410);
            }
        }
        return;
    }
#pragma unroll
    for (int r = 0; r < TN * TN / 4 / TTH; ++r) {
        const int id = threadIdx.x + TTH * r, row = id >> 4, quad = id & 15;
        const ulonglong2 v01 = *reinterpret_cast<const ulonglong2*>(src + row * TN + quad * 4);
        const ulonglong2 v23 = *reinterpret_cast<const ulonglong2*>(src + row * TN + quad * 4 + 2);
        u64 x[4] = {v01.x, v01.y, v23.x, v23.y};
        if (NEG) {
#pragma unroll
            for (int j = 0; j < 4; ++j) x[j] = q - x[j];
        }
        unsigned char* d = dst + row * TRS + koff + quad * 4;
#pragma unroll
        for (int v = 0; v < ND; ++v) {
            const int sh = v < 4 ? 0 : 32;
            const u32 sel = (u32)(v & 3) | ((u32)(4 + (v & 3)) << 4);
            const u32 t01 = __byte_perm((u32)(x[0] >> sh), (u32)(x[1] >> sh), sel);
            const u32 t23 = __byte_perm((u32)(x[2] >> sh), (u32)(x[3] >> sh), sel);
            *reinterpret_cast<u32*>(d + v * TN * TRS) = __byte_perm(t01, t23, 0x5410);
        }
    }
}

template <int ND, bool TR>
__global__ void __launch_bounds__(TTH, 2)
trace_gemm_tc_kernel(const u64* __restrict__ Ar, const u64* __restrict__ Ai, const u64* __restrict__ Bpr,
                     const u64* __restrict__ Bpi, u64* __restrict__ Cr, u64* __restrict__ Ci,
                     const u64* __restrict__ moduli, const ulonglong2* __restrict__ mu_tab, LimbList ll, int log2n) {
    extern __shared__ __align__(16) unsigned char sm_tt[];
    unsigned char* sA = sm_tt;                          // [ND][64][TRS]: (Ar | Ai) digits
    unsigned char* sB = sm_tt + ND * TN * TRS;          // [ND][64][TRS]: one half of B'
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, tig = lane & 3;
    const int wm = warp & 3, wn = warp >> 2;
    const int b = blockIdx.x / ll.cnt, l = ll.l[blockIdx.x % ll.cnt];
    const size_t base = ((size_t)b * ll.limbs + l) * TN * TN;
    const u64 q = moduli[l];
    const ulonglong2 mu = mu_tab[l];

    stage_digits<ND, false, TR>(sA, Ar + base, 0, q);
    stage_digits<ND, false, TR>(sA, Ai + base, TN, q);
#pragma unroll 1
    for (int half = 0; half < 2; ++half) {              // 0: Cr = Ar Br + Ai (q - Bi);  1: Ci = Ar Bi + Ai Br
        if (half) __syncthreads();                      // every warp is done with the first half's B'
        if (half == 0) {
            stage_digits<ND, false, TR>(sB, Bpr + base, 0, q);
            stage_digits<ND, true, TR>(sB, Bpi + base, TN, q);
        } else {
            stage_digits<ND, false, TR>(sB, Bpi + base, 0, q);
            stage_digits<ND, false, TR>(sB, Bpr + base, TN, q);
        }
        __syncthreads();
        u64* __restrict__ C = (half ? Ci : Cr) + base;
#pragma unroll 1
        for (int sub = 0; sub < 2; ++sub) {
            const int col0 = sub * 32 + wn * 16;
            int acc[2 * ND - 1][2][4];
#pragma unroll
            for (int w = 0; w < 2 * ND - 1; ++w)
#pragma unroll
                for (int nt = 0; nt < 2; ++nt)
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc[w][nt][j] = 0;
            const unsigned char* pa = sA + (wm * 16 + g) * TRS + tig * 4;
            const unsigned char* pb = sB + (col0 + g) * TRS + tig * 4;
#pragma unroll
            for (int ks = 0; ks < TK / 32; ++ks) {
                u32 af[ND][4];
#pragma unroll
                for (int u = 0; u < ND; ++u) {
                    const unsigned char* p = pa + u * TN * TRS + ks * 32;
                    af[u][0] = *reinterpret_cast<const u32*>(p);
                    af[u][1] = *reinterpret_cast<const u32*>(p + 8 * TRS);
                    af[u][2] = *reinterpret_cast<const u32*>(p + 16);
                    af[u][3] = *reinterpret_cast<const u32*>(p + 8 * TRS + 16);
                }
#pragma unroll
                for (int v = 0; v < ND; ++v) {
                    const unsigned char* p = pb + v * TN * TRS + ks * 32;
                    const u32 b00 = *reinterpret_cast<const u32*>(p), b01 = *reinterpret_cast<const u32*>(p + 16);
                    const u32 b10 = *reinterpret_cast<const u32*>(p + 8 * TRS), b11 = *reinterpret_cast<const u32*>(p + 8 * TRS + 16);
#pragma unroll
                    for (int u = 0; u < ND; ++u) {
                        mma_u8_t(acc[u + v][0], af[u], b00, b01);
                        mma_u8_t(acc[u + v][1], af[u], b10, b11);
                    }
                }
            }
            // recombine, times n, one Barrett per output, 16-byte stores (two adjacent columns)
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) {
                const int col = col0 + nt * 8 + tig * 2;
#pragma unroll
                for (int hf = 0; hf < 2; ++hf) {
                    const int row = wm * 16 + g + hf * 8;
                    u64 r[2];
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        u64 lo = 0, hi = 0;
#pragma unroll
                        for (int w = 0; w < 2 * ND - 1; ++w) {
                            const u64 s = (u64)(u32)acc[w][nt][hf * 2 + e];
                            const int sh = 8 * w;
                            if (sh < 64) {
                                const u64 x = s << sh;
                                lo += x; hi += (lo < x);
                                if (sh > 32) hi += s >> (64 - sh);
                            } else {
                                hi += s << (sh - 64);
                            }
                        }
                        hi = (hi << log2n) | (lo >> (64 - log2n));
                        lo <<= log2n;
                        r[e] = reduce128(lo, hi, q, mu.x, mu.y);
                    }
                    *reinterpret_cast<ulonglong2*>(C + row * TN + col) = make_ulonglong2(r[0], r[1]);
                }
            }
        }
    }
}

template <int ND, bool TR>
cudaError_t launch_trace_nd(const u64* Ar, const u64* Ai, const u64* Bpr, const u64* Bpi, u64* Cr, u64* Ci, const u64* moduli,
                            const ulonglong2* mu, const LimbList& ll, int batch, cudaStream_t st) {
    constexpr int smem = 2 * ND * TN * TRS;
    cudaError_t e = cudaFuncSetAttribute(trace_gemm_tc_kernel<ND, TR>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    trace_gemm_tc_kernel<ND, TR><<<(unsigned)(batch * ll.cnt), TTH, smem, st>>>(Ar, Ai, Bpr, Bpi, Cr, Ci, moduli, mu, ll, 6);
    return cudaGetLastError();
}
}  // namespace

// n = 64, every modulus below 2^48 (at most six digits), at most 64 limbs
bool trace_gemm_tc_eligible(int n, int limbs, const int* h_ndig) {
    if (n != TN || limbs > 64 || !h_ndig) return false;
    for (int l = 0; l < limbs; ++l)
        if (h_ndig[l] > 6) return false;
    return true;
}

cudaError_t trace_gemm_tc_launch(const u64* Ar, const u64* Ai, const u64* Bpr, const u64* Bpi, u64* Cr, u64* Ci,
                                 const u64* moduli, const ulonglong2* mu, const int* h_ndig, int limbs, int batch,
                                 cudaStream_t st, int transposed) {
    cudaError_t e = cudaSuccess;
    for (int nd = 4; nd <= 6 && e == cudaSuccess; ++nd) {       // one launch per digit count (<= 4 digits run as 4)
        LimbList ll; ll.cnt = 0; ll.limbs = limbs;
        for (int l = 0; l < limbs; ++l)
            if ((h_ndig[l] <= 4 ? 4 : h_ndig[l]) == nd) ll.l[ll.cnt++] = (unsigned char)l;
        if (!ll.cnt) continue;
        if (transposed)
            e = nd == 4 ? launch_trace_nd<4, true>(Ar, Ai, Bpr, Bpi, Cr, Ci, moduli, mu, ll, batch, st)
              : nd == 5 ? launch_trace_nd<5, true>(Ar, Ai, Bpr, Bpi, Cr, Ci, moduli, mu, ll, batch, st)
                        : launch_trace_nd<6, true>(Ar, Ai, Bpr, Bpi, Cr, Ci, moduli, mu, ll, batch, st);
        else
            e = nd == 4 ? launch_trace_nd<4, false>(Ar, Ai, Bpr, Bpi, Cr, Ci, moduli, mu, ll, batch, st)
              : nd == 5 ? launch_trace_nd<5, false>(Ar, Ai, Bpr, Bpi, Cr, Ci, moduli, mu, ll, batch, st)
                        : launch_trace_nd<6, false>(Ar, Ai, Bpr, Bpi, Cr, Ci, moduli, mu, ll, batch, st);
    }
    return e;
}

}  // namespace mfhe
