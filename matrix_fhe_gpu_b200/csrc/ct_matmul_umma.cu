// ct_matmul_umma.cu -- the four Gaussian-integer GEMMs of the encrypted matrix product (mfhe_ct_matmul)
// on the 5th-generation tensor cores (tcgen05.mma kind::i8, accumulators in TMEM), n = 64, exact.
//
// Replaces, for this caller, four launches of the mma.sync kernel of trace_tc.cu (itself the tensor-core form of the
// reference's trace GEMM, src/core/trace.cu:63-130, batched_trace.cu:85-161: one thread per output, four u128 % per MAC).
//
// Per (lane w, limb l) the product has four components  D_{c,vc}[x][z] = n * sum_y U_c[y][x] * V'_vc[y][z]  (complex; c, vc in
// {b, a} = the halves of the two ciphertext pairs).  Written as real GEMMs with K = 2n = 128 (as in trace_tc.cu):
//     [Dr | Di] = [Ur | Ui] x [ Vr  q-Vi ; Vi  Vr ]^T
// and the two U components share their right-hand side, so they are STACKED along M:  A = [U_b ; U_a] is 128 x 128, exactly
// the M = 128 tile of a cta_group::1 instruction, and one problem is two GEMMs of 128 x 128 x 128 (vc = b, a) instead of four
// of 64 x 128 x 128.  Operands are cut into ND = ceil(bits(q)/8) unsigned 8-bit digits; with the digit-stack form of
// wcrt_umma.cu the B operand of an instruction is the stack of all ND digit planes of a 32-column block (N' = 32 ND), the
// instruction for A's digit u accumulates at TMEM column 32 u, so position w = u + v collects every digit pair with one
// instruction per (u, K step):  4 column blocks x 4 K steps x ND instructions per GEMM.
//
// CTA (persistent, one per SM, 512 threads) per problem:
//   stage A  : 2 components x (re, im) x 64 x 64 words -> ND planes [128 m][128 k] in the UMMA canonical K-major layout
//              (8 x 16-byte core matrices); the contraction index y is the SLOW index of the ciphertext matrices, so lanes
//              run along x (coalesced 256-byte segments) and a thread packs four consecutive y into one 32-bit store per plane
//   per vc   : stage B' = digit stacks of the four column blocks (each V word lands twice: Vr in the Dr and Di columns,
//              Vi and q - Vi once each), then per column block: one thread issues the 4 ND instructions, commits to an
//              mbarrier; all 16 warps read the 2ND-1 partial sums of their 32 rows x 8 columns from TMEM, recombine to 128
//              bits, multiply by n (shift), ONE Barrett reduction, 32-byte stores.
#include <cstdio>
#include "ctx.h"
#include "modmath.cuh"

#ifndef MFHE_MM_PREFETCH
#define MFHE_MM_PREFETCH 1
#endif
#ifndef MFHE_MM_DBG
#define MFHE_MM_DBG 0      // timing experiments only: 1 = no global loads while staging, 2 = no epilogue, 4 = no MMAs, 8 = no digit stores
#endif
namespace mfhe {

namespace {
constexpr int CN = 64, CM = 128, CK = 128, CUN = 32, CTH = 512;
constexpr int SR = 16 / (CTH / 64);                  // staging rounds: 16 four-k groups per row over CTH / 64 thread groups
constexpr int EC = CUN / (CTH / 128);                // epilogue: columns per thread (TMEM lane quarter = warp % 4)
constexpr int PLANE = CM * CK;                       // 16 KiB: one digit plane of A (128 rows x 128 k)

__device__ __forceinline__ u32 smem_u32(const void* p) { return (u32)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(u32 bar, u32 count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(u32 bar, u32 parity) {
    u32 ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    } while (!ok);
}
// K-major, no swizzle: core matrix = 8 rows x 16 B; LBO = step between the two 16-byte K halves of an instruction,
// SBO = step between 8-row groups (a row group holds 128 k = 8 core matrices = 1 KiB here)
__device__ __forceinline__ u64 umma_desc(u32 saddr) {
    return (u64)((saddr & 0x3FFFFu) >> 4) | ((u64)(128 >> 4) << 16) | ((u64)(1024 >> 4) << 32) | (1ull << 46);
}
__host__ __device__ constexpr u32 idesc(int n) { return (2u << 4) | ((u32)(n >> 3) << 17) | ((u32)(CM >> 4) << 24); }
__device__ __forceinline__ void umma_i8(u32 d_tmem, u64 adesc, u64 bdesc, u32 id, u32 accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(id), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(u32 bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// 16 lanes x 256 bits: thread T gets lane T/4 (r0, r1) and lane T/4 + 8 (r2, r3), columns 2 (T%4), 2 (T%4) + 1 of the
// 8-column group (cute Copy_Traits<SM100_TMEM_LOAD_16dp256b1x>)
__device__ __forceinline__ void tmem_ld_16x256(u32 taddr, u32 (&r)[4]) {
    asm volatile("tcgen05.ld.sync.aligned.16x256b.x1.b32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(taddr));
}

// a (re, im) pair of 64 x 64 matrices (2 x 32 KiB = 512 lines) towards L2, one line per thread: what the NEXT staging step
// reads, so that its loads find the lines on their way
__device__ __forceinline__ void prefetch_pair(const u64* re, const u64* im) {
#if MFHE_MM_PREFETCH
    const u64* p = (threadIdx.x < 256 ? re : im) + (threadIdx.x & 255) * 16;
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
#endif
}

// byte offset of (row, k) inside a K-major canonical operand whose rows hold 128 k
__device__ __forceinline__ u32 canon(u32 row, u32 k) { return (row >> 3) * 1024 + (k >> 4) * 128 + (row & 7) * 16 + (k & 15); }

// four residues (consecutive k of one row) -> one 32-bit word per digit plane
template <int ND>
__device__ __forceinline__ void digits4(const u64 (&x)[4], u32 (&w)[ND]) {
#pragma unroll
    for (int v = 0; v < ND; ++v) {
        const int sh = v < 4 ? 0 : 32;
        const u32 sel = (u32)(v & 3) | ((u32)(4 + (v & 3)) << 4);
        const u32 t01 = __byte_perm((u32)(x[0] >> sh), (u32)(x[1] >> sh), sel);
        const u32 t23 = __byte_perm((u32)(x[2] >> sh), (u32)(x[3] >> sh), sel);
        w[v] = __byte_perm(t01, t23, 0x5410);
    }
}

// s0 + 2^8 s1 + 2^16 s2 + 2^24 s3 + 2^32 s4 for partial sums below 2^28 (result below 2^61)
__device__ __forceinline__ u64 fold5(u32 s0, u32 s1, u32 s2, u32 s3, u32 s4) {
    u64 r = ((u64)s4 << 32) | s0;
    asm("mad.wide.u32 %0, %1, 256, %0;" : "+l"(r) : "r"(s1));
    asm("mad.wide.u32 %0, %1, 65536, %0;" : "+l"(r) : "r"(s2));
    asm("mad.wide.u32 %0, %1, 16777216, %0;" : "+l"(r) : "r"(s3));
    return r;
}

struct MatmulLimbs { int cnt; int limbs; unsigned char l[64]; };

template <int ND>
struct MmCfg {
    static constexpr int NP = ND * CUN;                          // B' rows of a column block = instruction width
    static constexpr int BBLK = NP * CK;                         // bytes of one column block's digit stack
    static constexpr int SMEM = ND * PLANE + 4 * BBLK + 256;     // A planes + 4 column blocks + barrier / TMEM slot
    static constexpr int ACC = (2 * ND - 1) * CUN;               // TMEM columns in use
    static_assert(ACC <= 512 && NP <= 256 && NP % 16 == 0 && EC == 8, "tile width");
};

template <int ND>
__global__ void __launch_bounds__(CTH, 1)
ct_matmul_umma_kernel(const u64* __restrict__ Ure, const u64* __restrict__ Uim, const u64* __restrict__ Vre,
                      const u64* __restrict__ Vim, u64* __restrict__ out, size_t T, const u64* __restrict__ moduli,
                      const ulonglong2* __restrict__ mu_tab, MatmulLimbs ll, int nprob) {
    using C = MmCfg<ND>;
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* smem = reinterpret_cast<unsigned char*>(((uintptr_t)smem_raw + 127) & ~(uintptr_t)127);
    unsigned char* sA = smem;                                    // [ND][128 m][128 k] canonical
    unsigned char* sB = smem + ND * PLANE;                       // [4 column blocks][ND * 32 rows][128 k] canonical
    unsigned char* ctrl = sB + 4 * C::BBLK;
    const u32 bar = smem_u32(ctrl);
    volatile u32* tmem_slot = reinterpret_cast<volatile u32*>(ctrl + 16);
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;

    if (t == 0) {
        mbar_init(bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        __syncwarp();
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32((const void*)tmem_slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const u32 tmem = *tmem_slot;
    u32 phase = 0;

    // staging geometry: lanes along the fast source index (x resp. z), a thread owns four consecutive y (= k) per round;
    // the 4-k group inside a 16-byte chunk is rotated by the 8-row group so that a warp's stores cover all 32 banks
    const int sx = t & 63;
    const int rot = (sx >> 3) & 3;

    for (int prob = blockIdx.x; prob < nprob; prob += gridDim.x) {
#ifdef MFHE_MM_TIMING
        long long tk[6] = {clock64(), 0, 0, 0, 0, 0}, acc_t[4] = {0, 0, 0, 0};
#define MM_TICK(k) tk[k] = clock64()
#define MM_ACC(i, a, b) acc_t[i] += tk[b] - tk[a]
#else
#define MM_TICK(k)
#define MM_ACC(i, a, b)
#endif
        const int w = prob / ll.cnt, l = ll.l[prob % ll.cnt];
        const size_t base = ((size_t)w * ll.limbs + l) * CN * CN;
        const u64 q = moduli[l], q2 = 2 * q;
        const ulonglong2 mu = mu_tab[l];
        const int kb = 64 - __clzll((long long)q);
        const u64 mu1 = (mu.y << (kb - 2)) | (mu.x >> (66 - kb));              // floor(2^128 / q) >> (66 - k) = floor(2^(k+62) / q)

        // ---- A = [U_b ; U_a]: row m = 64 c + x, k = y (re) / 64 + y (im) ----
#pragma unroll 1
        for (int c = 0; c < 2; ++c) {
            const u64* sr = Ure + c * T + base;
            const u64* si = Uim + c * T + base;
            if (c == 1) prefetch_pair(Vre + base, Vim + base);
#pragma unroll
            for (int r = 0; r < SR; ++r) {
                const int kq = (((t >> 6) & 3) ^ rot) + 4 * ((t >> 8) + (CTH / 256) * r);   // y = 4 kq .. 4 kq + 3
                u64 xr[4], xi[4];
#pragma unroll
#if MFHE_MM_DBG & 1
                for (int j = 0; j < 4; ++j) { xr[j] = (u64)(kq + j + t); xi[j] = (u64)(kq * j + t); }
#else
                for (int j = 0; j < 4; ++j) { xr[j] = sr[(4 * kq + j) * CN + sx]; xi[j] = si[(4 * kq + j) * CN + sx]; }
#endif
                u32 wr[ND], wi[ND];
                digits4<ND>(xr, wr); digits4<ND>(xi, wi);
                const u32 o_r = canon(64 * c + sx, 4 * kq), o_i = canon(64 * c + sx, 64 + 4 * kq);
#pragma unroll
                for (int v = 0; v < ND; ++v) {
                    *reinterpret_cast<u32*>(sA + v * PLANE + o_r) = wr[v];
                    *reinterpret_cast<u32*>(sA + v * PLANE + o_i) = wi[v];
                }
            }
        }

        MM_TICK(1);                                                             // tk[1] - tk[0]: A staging
#pragma unroll 1
        for (int vc = 0; vc < 2; ++vc) {
            MM_TICK(2);
            // ---- B': logical row n < 64: Dr column z = n  (k < 64: Vr, k >= 64: q - Vi);  n >= 64: Di column  (Vi, Vr).
            //      Column block cb = n / 32, stack row = 32 v + n % 32 ----
            const u64* sr = Vre + vc * T + base;
            const u64* si = Vim + vc * T + base;
            if (vc == 0) prefetch_pair(Vre + T + base, Vim + T + base);
#pragma unroll
            for (int r = 0; r < SR; ++r) {
                const int kq = (((t >> 6) & 3) ^ rot) + 4 * ((t >> 8) + (CTH / 256) * r);
                u64 xr[4], xi[4], xn[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
#if MFHE_MM_DBG & 1
                    xr[j] = (u64)(kq + j + t); xi[j] = (u64)(kq * j + t);
#else
                    xr[j] = sr[(4 * kq + j) * CN + sx]; xi[j] = si[(4 * kq + j) * CN + sx];
#endif
                    xn[j] = q - xi[j];
                }
                u32 wr[ND], wi[ND], wn[ND];
                digits4<ND>(xr, wr); digits4<ND>(xi, wi); digits4<ND>(xn, wn);
                const int cb0 = sx >> 5, nn = sx & 31;                         // Dr columns live in blocks 0, 1; Di in 2, 3
                unsigned char* b_re = sB + cb0 * C::BBLK;
                unsigned char* b_im = sB + (2 + cb0) * C::BBLK;
#pragma unroll
                for (int v = 0; v < ND; ++v) {
                    const u32 lo = canon(32 * v + nn, 4 * kq), hi = canon(32 * v + nn, 64 + 4 * kq);
                    *reinterpret_cast<u32*>(b_re + lo) = wr[v];
                    *reinterpret_cast<u32*>(b_re + hi) = wn[v];
                    *reinterpret_cast<u32*>(b_im + lo) = wi[v];
                    *reinterpret_cast<u32*>(b_im + hi) = wr[v];
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");       // generic-proxy stores -> visible to the MMA
            __syncthreads();
            MM_TICK(3); MM_ACC(0, 2, 3);                                        // B' staging + barrier

            // one thread: the 4 ND instructions of column block cb, then a commit on the block barrier
            auto issue = [&](int cb) {
#if MFHE_MM_DBG & 4
                return;
#endif
                if (t != 0) return;
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const u32 a0 = smem_u32(sA), b0 = smem_u32(sB + cb * C::BBLK);
#pragma unroll
                for (int ks = 0; ks < CK / 32; ++ks) {
                    const u64 bdesc = umma_desc(b0 + ks * 256);
                    if (ks == 0) {
                        // nothing to accumulate onto: position block u + ND - 1 is new for every u, the others were
                        // started by u - 1 -- accumulate piece + overwrite piece
                        umma_i8(tmem, umma_desc(a0), bdesc, idesc(C::NP), 0);
#pragma unroll
                        for (int u = 1; u < ND; ++u) {
                            const u64 adesc = umma_desc(a0 + u * PLANE);
                            umma_i8(tmem + (u32)(u * CUN), adesc, bdesc, idesc((ND - 1) * CUN), 1);
                            umma_i8(tmem + (u32)((u + ND - 1) * CUN), adesc, umma_desc(b0 + (ND - 1) * (CUN / 8) * 1024),
                                    idesc(CUN), 0);
                        }
                    } else {
#pragma unroll
                        for (int u = 0; u < ND; ++u)
                            umma_i8(tmem + (u32)(u * CUN), umma_desc(a0 + u * PLANE + ks * 256), bdesc, idesc(C::NP), 1);
                    }
                }
                umma_commit(bar);
            };

#pragma unroll 1
            for (int cb = 0; cb < 4; ++cb) {
                // block cb's MMAs were issued before the previous block's reductions (or just above, for cb = 0)
                MM_TICK(2);
                if (cb == 0) issue(0);
                if (vc == 1 && cb < 2 && prob + (int)gridDim.x < nprob) {       // the next problem's U components
                    const int p2 = prob + (int)gridDim.x;
                    const size_t b2 = ((size_t)(p2 / ll.cnt) * ll.limbs + ll.l[p2 % ll.cnt]) * CN * CN;
                    prefetch_pair(Ure + cb * T + b2, Uim + cb * T + b2);
                }
#if !(MFHE_MM_DBG & 4)
                mbar_wait(bar, phase);
                phase ^= 1;
#endif
                MM_TICK(3); MM_ACC(1, 2, 3);                                    // (issue of block 0 +) wait for the block's MMAs
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

                // ---- epilogue: warp = TMEM lane quarter (warp % 4) x 8 columns (warp / 4).  The 16x256b load shape gives a thread
                // rows lane/4 and lane/4 + 8 of a 16-row half and the column pair 2 (lane % 4): four threads share a row, so
                // one store instruction writes 64-byte row segments (8 lines) instead of 32 scattered 16-byte pieces ----
                const int c = (warp & 3) >> 1;                                  // rows 0..63: U's b component, 64..127: a
                const int ri = cb >> 1;                                         // 0: real part, 1: imaginary part
                u64* o = out + (size_t)(2 * (c + 2 * vc) + ri) * T + base + (size_t)(32 * (warp & 1) + (lane >> 2)) * CN
                         + (cb & 1) * 32 + (warp >> 2) * EC + 2 * (lane & 3);
                const u32 trow = tmem + ((u32)(32 * (warp & 3)) << 16) + (u32)((warp >> 2) * EC);
#if MFHE_MM_DBG & 2
                if (q == 12345) o[0] = trow;
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                __syncthreads();
                if (cb < 3) issue(cb + 1);
#else
                // phase 1 (holds TMEM): value = n * sum_w S_w 2^(8w) with S_w < 2^27 is folded into
                //   vlo = sum_{w<5} S_w 2^(8w)  (< 2^60),   vup = sum_{5<=w<10} S_w 2^(8(w-5)),   [vtop = S_10 for six digits]
                u64 vlo[EC], vup[EC];
                u32 vtop[(2 * ND - 1 > 10) ? EC : 1];
                u32 acc[2][2 * ND - 1][4];                                      // loads of half 1 fly while half 0 is folded
#pragma unroll
                for (int p = 0; p < 2 * ND - 1; ++p) tmem_ld_16x256(trow + (u32)(p * CUN), acc[0][p]);
#pragma unroll
                for (int grp = 0; grp < 2; ++grp) {                             // 16-row halves of the lane quarter
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    if (grp == 0) {
#pragma unroll
                        for (int p = 0; p < 2 * ND - 1; ++p) tmem_ld_16x256(trow + (16u << 16) + (u32)(p * CUN), acc[1][p]);
                    }
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        // S_0 + 2^8 S_1 + 2^16 S_2 + 2^24 S_3 as three 32 x 32 + 64 multiply-adds, 2^32 S_4 into the high word
                        const u32 (&a)[2 * ND - 1][4] = acc[grp & 1];
                        u64 lo = fold5(a[0][e], a[1][e], a[2][e], a[3][e], 2 * ND - 1 > 4 ? a[2 * ND - 1 > 4 ? 4 : 0][e] : 0u);
                        u64 up = 0;
                        if constexpr (2 * ND - 1 > 5)
                            up = fold5(a[5][e], 2 * ND - 1 > 6 ? a[2 * ND - 1 > 6 ? 6 : 0][e] : 0u, 2 * ND - 1 > 7 ? a[2 * ND - 1 > 7 ? 7 : 0][e] : 0u,
                                       2 * ND - 1 > 8 ? a[2 * ND - 1 > 8 ? 8 : 0][e] : 0u, 2 * ND - 1 > 9 ? a[2 * ND - 1 > 9 ? 9 : 0][e] : 0u);
                        vlo[4 * grp + e] = lo; vup[4 * grp + e] = up;
                        if constexpr (2 * ND - 1 > 10) vtop[4 * grp + e] = acc[grp & 1][10][e];
                    }
                }
                MM_TICK(4); MM_ACC(2, 3, 4);                                    // TMEM reads + fold
                // the accumulators are in registers: the next column block's MMAs run under the reductions below
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                __syncthreads();
                if (cb < 3) issue(cb + 1);
                // phase 2: value < 64 * 128 * q^2 < 2^(k+62) for k = bits(q) <= 48, so the ONE-word Barrett of the W-CRT
                // epilogue applies: qhat = mulhi(v >> (k-2), floor(2^(k+62) / q)) is within 2 of the quotient (HAC 14.42)
#pragma unroll
                for (int e = 0; e < EC; e += 2) {
                    u64 res[2];
#pragma unroll
                    for (int k2 = 0; k2 < 2; ++k2) {
                        const u64 xs = vup[e + k2] << 40;
                        u64 lo = vlo[e + k2] + xs;
                        u64 hi = (vup[e + k2] >> 24) + (lo < xs);
                        if constexpr (2 * ND - 1 > 10) hi += (u64)vtop[e + k2] << 16;
                        hi = (hi << 6) | (lo >> 58);                            // times n = 64
                        lo <<= 6;
                        const u64 x1 = (lo >> (kb - 2)) | (hi << (66 - kb));
                        const u64 rr = lo - __umul64hi(x1, mu1) * q;
                        res[k2] = csub(csub(rr, q2), q);
                    }
                    // e = 4 half + 2 hf: row 16 half + 8 hf (+ lane / 4), the thread's column pair
                    *reinterpret_cast<ulonglong2*>(o + (size_t)(16 * (e >> 2) + 8 * ((e >> 1) & 1)) * CN) = make_ulonglong2(res[0], res[1]);
                }
#endif
                MM_TICK(5); MM_ACC(3, 4, 5);                                    // barrier + issue of the next block + reductions + stores
            }
            __syncthreads();                                                    // every warp is done with this vc's B' stacks
        }
#ifdef MFHE_MM_TIMING
        if (blockIdx.x == 3 && t == 0 && prob == 3 + 10 * (int)gridDim.x)
            printf("problem chain (clk): A stage %lld  B stage (2x) %lld  mma wait (8x) %lld  tmem+fold (8x) %lld  issue+reduce+store (8x) %lld  total %lld\n",
                   tk[1] - tk[0], acc_t[0], acc_t[1], acc_t[2], acc_t[3], clock64() - tk[0]);
#endif
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
}

template <int ND>
cudaError_t launch_nd(const u64* Ure, const u64* Uim, const u64* Vre, const u64* Vim, u64* out, size_t T, const u64* moduli,
                      const ulonglong2* mu, const MatmulLimbs& ll, int phi, cudaStream_t st) {
    using C = MmCfg<ND>;
    cudaError_t e = cudaFuncSetAttribute(ct_matmul_umma_kernel<ND>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
    if (e != cudaSuccess) return e;
    int sms = 0, dev = 0;
    cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int nprob = phi * ll.cnt;
    ct_matmul_umma_kernel<ND><<<nprob < sms ? nprob : sms, CTH, C::SMEM, st>>>(Ure, Uim, Vre, Vim, out, T, moduli, mu, ll, nprob);
    return cudaGetLastError();
}
}  // namespace

// n = 64, every modulus below 2^48 (at most six digits), at most 64 limbs
bool ct_matmul_umma_eligible(int n, int limbs, const int* h_ndig) {
    if (n != CN || limbs > 64 || !h_ndig) return false;
    for (int l = 0; l < limbs; ++l)
        if (h_ndig[l] > 6) return false;
    return true;
}

// U*, V*: ciphertext pairs ([b | a] halves of T words each, MATRIX layout [w][l][y][x]); out: 8 poly sets
// (2 (c + 2 vc) + {re, im}) of T words, [w][l][x][z] -- the layout and order mfhe_ct_matmul documents
cudaError_t ct_matmul_umma_launch(const u64* Ure, const u64* Uim, const u64* Vre, const u64* Vim, u64* out, size_t T,
                                  const u64* moduli, const ulonglong2* mu, const int* h_ndig, int limbs, int phi, cudaStream_t st) {
    cudaError_t e = cudaSuccess;
    for (int nd = 4; nd <= 6 && e == cudaSuccess; ++nd) {       // one launch per digit count (<= 4 digits run as 4)
        MatmulLimbs ll; ll.cnt = 0; ll.limbs = limbs;
        for (int l = 0; l < limbs; ++l)
            if ((h_ndig[l] <= 4 ? 4 : h_ndig[l]) == nd) ll.l[ll.cnt++] = (unsigned char)l;
        if (!ll.cnt) continue;
        e = nd == 4 ? launch_nd<4>(Ure, Uim, Vre, Vim, out, T, moduli, mu, ll, phi, st)
          : nd == 5 ? launch_nd<5>(Ure, Uim, Vre, Vim, out, T, moduli, mu, ll, phi, st)
                    : launch_nd<6>(Ure, Uim, Vre, Vim, out, T, moduli, mu, ll, phi, st);
    }
    return e;
}

}  // namespace mfhe
