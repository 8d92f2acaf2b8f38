// wcrt_umma.cu -- the W-CRT modular GEMM on the 5th-generation tensor cores (tcgen05.mma kind::i8,
// accumulators in TMEM):   out[o][c] = sum_i tab[l][i][o] * in[i][c]  mod q_l,   bit-exact.
//
// Arithmetic (same digit scheme as wcrt_tc.cu): tab = sum_u a_u 2^(8u), in = sum_v b_v 2^(8v), ND digits each;
//     sum_i tab*in = sum_w 2^(8w) S_w,  S_w = sum_{u+v=w} sum_i a_u[i] b_v[i] < 2^28  (exact in s32).
// Mapping onto UMMA.  The B operand of one instruction is the STACK of all input digit planes
//     B' = [b_0; b_1; ...; b_{ND-1}]   (N' = ND*32 rows of 32 columns c each, K-major),
// and the accumulator of digit position w lives in TMEM columns [32w, 32w+32).  Issuing
//     D[columns 32u .. 32u + N') += A_u * B'^T                      for u = 0 .. ND-1
// adds a_u b_v into position u+v for every v at once: ND instructions of shape 128 x (32 ND) x 32 per K step
// instead of ND^2 small ones, and every A tile is read from shared memory once per K step.
//
//   CTA tile   128 outputs o  x  32 columns c, all K (phi) in chunks of 64, NS-stage ring in shared memory
//   warp 0     lane 0 streams the table digits: the table is stored pre-tiled in the UMMA canonical
//              (no-swizzle, K-major, 8x16-byte core matrices) layout, so one 8 KiB cp.async.bulk per digit plane
//              and chunk lands a ready-to-use A tile; completion through an mbarrier (complete_tx)
//   warps 4-7  produce B': fused source (memory in either layout / int64 / quantiser / samplers) -> canonical
//              u64 residues, four consecutive i per thread -> PRMT byte transpose -> K-major core matrices,
//              fence.proxy.async, mbarrier arrive
//   warp 1     lane 0 issues the tcgen05.mma's and commits them to the stage's "empty" barrier
//   all warps  epilogue: tcgen05.ld the 2ND-1 partial sums of 8 columns, recombine to 128 bits, ONE Barrett
//              reduction per output, 16-byte stores in either layout.
#include "ctx.h"
#include "wcrt_common.cuh"

namespace mfhe {

namespace {
constexpr int UM = 128, UN = 32, UKC = 64, UTHREADS = 256;
constexpr int A_PLANE = UM * UKC;            // 8 KiB: one digit plane of one A chunk
constexpr int B_PLANE = UN * UKC;            // 2 KiB

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
__device__ __forceinline__ void mbar_arrive(u32 bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(u32 bar, u32 bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(u32 dst, const void* src, u32 bytes, u32 bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
// K-major, no swizzle: 8 rows x 16 B core matrices; LBO = step between the two 16-byte K halves of one
// instruction, SBO = step between 8-row groups (cute/arch/mma_sm100_desc.hpp SmemDescriptor, version 1)
__device__ __forceinline__ u64 umma_desc(u32 saddr, u32 lbo, u32 sbo) {
    return (u64)((saddr & 0x3FFFFu) >> 4) | ((u64)(lbo >> 4) << 16) | ((u64)(sbo >> 4) << 32) | (1ull << 46);
}
__device__ __forceinline__ void umma_i8(u32 d_tmem, u64 adesc, u64 bdesc, u32 idesc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, 1, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc) : "memory");
}
__device__ __forceinline__ void umma_commit(u32 bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld8(u32 taddr, u32 (&r)[8]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr));
}
__device__ __forceinline__ void tmem_st8_zero(u32 taddr) {
    const u32 z = 0;
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%1,%1,%1,%1,%1,%1,%1};" ::"r"(taddr), "r"(z) : "memory");
}
}  // namespace

template <int ND> struct UmmaCfg {
    static constexpr int NS = (ND <= 5) ? 4 : 3;                       // ring stages
    static constexpr int STAGE = ND * (A_PLANE + B_PLANE);
    static constexpr int SMEM = NS * STAGE + 1024;                      // + barriers / TMEM pointer, 1 KiB align slack
    static constexpr int NPRIME = ND * UN;                              // N of one instruction
    static constexpr int ACC_COLS = (2 * ND - 1) * UN;
    // kind::i8 instruction descriptor: D = s32, A = B = unsigned 8 bit, both K-major, N' x 128
    static constexpr u32 IDESC = (2u << 4) | ((u32)(NPRIME >> 3) << 17) | ((u32)(UM >> 4) << 24);
};

template <int SRC, int ND>
__global__ void __launch_bounds__(UTHREADS, 1)
wcrt_umma_kernel(const WcrtArgs a) {
    using C = UmmaCfg<ND>;
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* smem = reinterpret_cast<unsigned char*>(((uintptr_t)smem_raw + 127) & ~(uintptr_t)127);
    unsigned char* ctrl = smem + C::NS * C::STAGE;                     // barriers: fullA[NS], fullB[NS], empty[NS], acc
    const u32 bar0 = smem_u32(ctrl);
    auto bar_fullA = [&](int s) { return bar0 + 8u * s; };
    auto bar_fullB = [&](int s) { return bar0 + 8u * (C::NS + s); };
    auto bar_empty = [&](int s) { return bar0 + 8u * (2 * C::NS + s); };
    const u32 bar_acc = bar0 + 8u * (3 * C::NS);
    volatile u32* tmem_slot = reinterpret_cast<volatile u32*>(ctrl + 8 * (3 * C::NS + 1));

    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int lrel = blockIdx.z + a.lz0, l = a.limb0 + lrel;
    const int ot = blockIdx.y, o0 = ot * UM, c0 = blockIdx.x * UN;
    const u64 q = a.moduli[l];
    const ulonglong2 mu = a.mu[l];
    const int kchunks = a.phi / UKC;
    const int lsrc = (SRC == kSrcMem) ? lrel : l;

    if (t == 0) {
        for (int s = 0; s < C::NS; ++s) { mbar_init(bar_fullA(s), 1); mbar_init(bar_fullB(s), 128); mbar_init(bar_empty(s), 1); }
        mbar_init(bar_acc, 1);
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
    const u32 tlane = tmem + ((u32)(32 * (warp & 3)) << 16);

    // zero the accumulator columns (every instruction accumulates)
    {
        constexpr int G = C::ACC_COLS / 8, H = (G + 1) / 2;
        const int g0 = (warp >> 2) * H, g1 = (g0 + H < G) ? g0 + H : G;
        for (int gi = g0; gi < g1; ++gi) tmem_st8_zero(tlane + 8u * gi);
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();

    if (warp == 0) {
        // ===== table-digit producer: one bulk copy per digit plane and chunk =====
        if (lane == 0) {
            const unsigned char* base = a.tabt + (size_t)l * 8 * a.phi * a.phi;
            for (int kt = 0; kt < kchunks; ++kt) {
                const int s = kt % C::NS;
                if (kt >= C::NS) mbar_wait(bar_empty(s), ((kt / C::NS) - 1) & 1);
                mbar_expect_tx(bar_fullA(s), ND * A_PLANE);
                const u32 dst = smem_u32(smem + s * C::STAGE);
#pragma unroll
                for (int u = 0; u < ND; ++u)
                    bulk_g2s(dst + u * A_PLANE,
                             base + ((size_t)u * a.phi * a.phi) + ((size_t)ot * kchunks + kt) * A_PLANE, A_PLANE, bar_fullA(s));
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (lane == 0) {
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            for (int kt = 0; kt < kchunks; ++kt) {
                const int s = kt % C::NS;
                const u32 par = (kt / C::NS) & 1;
                mbar_wait(bar_fullA(s), par);
                mbar_wait(bar_fullB(s), par);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const u32 sa = smem_u32(smem + s * C::STAGE), sb = sa + ND * A_PLANE;
#pragma unroll
                for (int ks = 0; ks < UKC / 32; ++ks) {
                    const u64 bdesc = umma_desc(sb + ks * 256, 128, 512);
#pragma unroll
                    for (int u = 0; u < ND; ++u)
                        umma_i8(tmem + (u32)(u * UN), umma_desc(sa + u * A_PLANE + ks * 256, 128, 512), bdesc, C::IDESC);
                }
                umma_commit(bar_empty(s));
            }
            umma_commit(bar_acc);
        }
        __syncwarp();
    } else if (warp >= 4) {
        // ===== input-digit producers (128 threads): item = (column c, four consecutive i) =====
        const int pt = t - 128;
        for (int kt = 0; kt < kchunks; ++kt) {
            const int s = kt % C::NS;
            if (kt >= C::NS) mbar_wait(bar_empty(s), ((kt / C::NS) - 1) & 1);
            unsigned char* sB = smem + s * C::STAGE + ND * A_PLANE;
            u64 rb[4][4];
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const int id = pt + 128 * r;                       // 512 items: c = id & 31 (low 3 bits + bits 5,6), kq
                const int c = ((id >> 5) & 3) * 8 + (id & 7), kq = ((id >> 7) & 3) * 4 + ((id >> 3) & 3);
#pragma unroll
                for (int j = 0; j < 4; ++j) rb[r][j] = wcrt_load_one<SRC>(a, kt * UKC + 4 * kq + j, lsrc, c0 + c, q, mu.y);
            }
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const int id = pt + 128 * r;
                const int c = ((id >> 5) & 3) * 8 + (id & 7), kq = ((id >> 7) & 3) * 4 + ((id >> 3) & 3);
                // row (v*32 + c) of B', bytes [4kq, 4kq+4): core matrix (row/8, kq/4), row%8, (kq%4)*4
                unsigned char* dst = sB + (c >> 3) * 512 + (kq >> 2) * 128 + (c & 7) * 16 + (kq & 3) * 4;
                const u32 l0 = (u32)rb[r][0], l1 = (u32)rb[r][1], l2 = (u32)rb[r][2], l3 = (u32)rb[r][3];
                const u32 h0 = (u32)(rb[r][0] >> 32), h1 = (u32)(rb[r][1] >> 32), h2 = (u32)(rb[r][2] >> 32), h3 = (u32)(rb[r][3] >> 32);
#pragma unroll
                for (int v = 0; v < ND; ++v) {
                    const u32 x0 = v < 4 ? l0 : h0, x1 = v < 4 ? l1 : h1, x2 = v < 4 ? l2 : h2, x3 = v < 4 ? l3 : h3;
                    const u32 sel = (u32)(v & 3) | ((u32)(4 + (v & 3)) << 4);
                    const u32 t01 = __byte_perm(x0, x1, sel), t23 = __byte_perm(x2, x3, sel);
                    *reinterpret_cast<u32*>(dst + v * B_PLANE) = __byte_perm(t01, t23, 0x5410);
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbar_arrive(bar_fullB(s));
        }
    }

    // ===== epilogue (all warps): lanes 32*(warp%4) .. +31 = rows, column half warp/4 =====
    mbar_wait(bar_acc, 0);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const int o = o0 + 32 * (warp & 3) + lane;
#pragma unroll 1
    for (int cg = 0; cg < 2; ++cg) {
        const int cc = (warp >> 2) * 16 + cg * 8;
        u32 acc[2 * ND - 1][8];
#pragma unroll
        for (int w = 0; w < 2 * ND - 1; ++w) tmem_ld8(tlane + (u32)(w * UN + cc), acc[w]);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        u64 r[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            u64 lo = 0, hi = 0;
#pragma unroll
            for (int w = 0; w < 2 * ND - 1; ++w) {
                const u64 sv = (u64)acc[w][e];
                const int sh = 8 * w;
                if (sh < 64) {
                    const u64 x = sv << sh;
                    lo += x; hi += (lo < x);
                    if (sh > 32) hi += sv >> (64 - sh);
                } else {
                    hi += sv << (sh - 64);
                }
            }
            r[e] = reduce128(lo, hi, q, mu.x, mu.y);
        }
#pragma unroll
        for (int e = 0; e < 8; e += 2) {
            const size_t off = poly_set_addr(a.out_layout, o, lrel, c0 + cc + e, a.n, a.cols, a.limbs);
            *reinterpret_cast<ulonglong2*>(a.out + off) = make_ulonglong2(r[e], r[e + 1]);
            if (a.out2) *reinterpret_cast<ulonglong2*>(a.out2 + off) = make_ulonglong2(r[e], r[e + 1]);
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
    }
}

template <int SRC, int ND>
static cudaError_t launch_umma_nd(const WcrtArgs& a, cudaStream_t st) {
    using C = UmmaCfg<ND>;
    static bool attr = false;
    if (!attr) {
        cudaError_t e = cudaFuncSetAttribute(wcrt_umma_kernel<SRC, ND>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
        if (e != cudaSuccess) return e;
        attr = true;
    }
    dim3 grid(a.cols / UN, a.phi / UM, a.lzn);
    wcrt_umma_kernel<SRC, ND><<<grid, UTHREADS, C::SMEM, st>>>(a);
    return cudaGetLastError();
}

template <int SRC>
static cudaError_t launch_umma_src(const WcrtArgs& a, int nd, cudaStream_t st) {
    switch (nd) {
        case 4: return launch_umma_nd<SRC, 4>(a, st);
        case 5: return launch_umma_nd<SRC, 5>(a, st);
        case 6: return launch_umma_nd<SRC, 6>(a, st);
        default: return cudaErrorInvalidValue;
    }
}

bool wcrt_umma_eligible(const WcrtArgs& a) {
    return a.tabt && a.h_ndig && a.phi % UM == 0 && a.phi % UKC == 0 && a.cols % UN == 0 && a.n % 2 == 0;
}

cudaError_t wcrt_umma_launch(const WcrtArgs& a, int nd, cudaStream_t st) {
    switch (a.source) {
        case kSrcMem:     return launch_umma_src<kSrcMem>(a, nd, st);
        case kSrcI64:     return launch_umma_src<kSrcI64>(a, nd, st);
        case kSrcQuantRe: return launch_umma_src<kSrcQuantRe>(a, nd, st);
        case kSrcQuantIm: return launch_umma_src<kSrcQuantIm>(a, nd, st);
        case kSrcUniform: return launch_umma_src<kSrcUniform>(a, nd, st);
        case kSrcGauss:   return launch_umma_src<kSrcGauss>(a, nd, st);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace mfhe
