// wcrt_tc.cu -- the W-CRT modular GEMM on the tensor cores:  out[o][c] = sum_i tab[l][i][o] * in[i][c]  mod q_l.
//
// Same map and same canonical residues as wcrt.cu (and as src/core/HE.cu:716-781, 1029-1114, 1245-1270 of the
// reference); exact integer arithmetic throughout.  Both operands are cut into 8-bit digits
//     tab = sum_u a_u 2^(8u),   in = sum_v b_v 2^(8v),      ND = ceil(bits(q) / 8) digits each,
// so that   sum_i tab*in = sum_w 2^(8w) S_w,   S_w = sum_{u+v=w} sum_i a_u[i] b_v[i]   (w = 0 .. 2ND-2).
// Every S_w is a u8 x u8 -> s32 GEMM (phi * 255^2 * ND < 2^28: no overflow), issued as
// mma.sync.m16n8k32.u8.u8 with one accumulator tile per w; the epilogue recombines the 2ND-1 partial sums
// into a 128-bit integer and applies ONE Barrett reduction per output.  A 35-bit limb costs 25 int8 MACs per
// modular MAC on the tensor pipe instead of 4 IMAD.WIDE (= 10 fmaheavy slots) on the integer pipe.
//
//   * table digits are precomputed at context creation as [limb][digit][o][i] bytes (i contiguous = the
//     K-major row an A fragment wants) and staged with cp.async;
//   * input digits are produced by the loader: the same fused sources as wcrt.cu (memory in either layout,
//     int64, quantiser, uniform / gaussian samplers) give canonical u64 residues in registers, four
//     consecutive i per thread, which are byte-transposed with PRMT into K-major words;
//   * CTA tile 64 (o) x 32 (c), K chunks of 64 double-buffered in shared memory (row stride 80 B keeps the
//     fragment loads and the loader stores bank-conflict free), 8 warps of 16 x 16 outputs each.
#include "ctx.h"
#include "wcrt_common.cuh"

namespace mfhe {

namespace {
constexpr int MC = 64, NC = 32, KC = 64, RS = 80, TCT = 256;

__device__ __forceinline__ void cp_async16(void* dst, const void* src) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit_wait_all() {
    asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
}
__device__ __forceinline__ void mma_u8(int (&c)[4], const u32 (&a)[4], u32 b0, u32 b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
}  // namespace

#ifndef MFHE_TC_MINB5
#define MFHE_TC_MINB5 2
#endif
template <int SRC, int ND>
__global__ void __launch_bounds__(TCT, (ND <= 4 ? 2 : ND == 5 ? MFHE_TC_MINB5 : 1))
wcrt_tc_kernel(const WcrtArgs a) {
    extern __shared__ __align__(16) unsigned char smem[];
    constexpr int A_STAGE = ND * MC * RS, B_STAGE = ND * NC * RS;
    unsigned char* sA = smem;                       // [2][ND][MC][RS]
    unsigned char* sB = smem + 2 * A_STAGE;         // [2][ND][NC][RS]
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int g = lane >> 2, tig = lane & 3;
    const int wm = warp & 3, wn = warp >> 2;
    const int lrel = blockIdx.z + a.lz0, l = a.limb0 + lrel;
    const int o0 = blockIdx.y * MC, c0 = blockIdx.x * NC;
    const u64 q = a.moduli[l];
    const ulonglong2 mu = a.mu[l];
    const unsigned char* tabd = a.tabd + (size_t)l * 8 * a.phi * a.phi;
    const int lsrc = (SRC == kSrcMem) ? lrel : l;

    // ---- loaders ------------------------------------------------------------------------------------
    auto issue_A = [&](int k0, int buf) {
        const int row = t >> 2, part = t & 3;
#pragma unroll
        for (int u = 0; u < ND; ++u)
            cp_async16(sA + buf * A_STAGE + (u * MC + row) * RS + part * 16,
                       tabd + ((size_t)u * a.phi + o0 + row) * a.phi + k0 + part * 16);
    };
    u64 rb[2][4];
    int bc[2], bkq[2];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        const int id = t + TCT * r;
        bc[r] = ((id >> 5) & 3) * 8 + (id & 7);
        bkq[r] = ((id >> 7) & 3) * 4 + ((id >> 3) & 3);
    }
    auto fetch_B = [&](int k0) {
#pragma unroll
        for (int r = 0; r < 2; ++r)
#pragma unroll
            for (int j = 0; j < 4; ++j) rb[r][j] = wcrt_load_one<SRC>(a, k0 + 4 * bkq[r] + j, lsrc, c0 + bc[r], q, mu.y);
    };
    auto stash_B = [&](int buf) {
#pragma unroll
        for (int r = 0; r < 2; ++r) {
            unsigned char* dst = sB + buf * B_STAGE + bc[r] * RS + bkq[r] * 4;
            const u32 l0 = (u32)rb[r][0], l1 = (u32)rb[r][1], l2 = (u32)rb[r][2], l3 = (u32)rb[r][3];
            const u32 h0 = (u32)(rb[r][0] >> 32), h1 = (u32)(rb[r][1] >> 32), h2 = (u32)(rb[r][2] >> 32), h3 = (u32)(rb[r][3] >> 32);
#pragma unroll
            for (int v = 0; v < ND; ++v) {
                const u32 x0 = v < 4 ? l0 : h0, x1 = v < 4 ? l1 : h1, x2 = v < 4 ? l2 : h2, x3 = v < 4 ? l3 : h3;
                const u32 sel = (u32)(v & 3) | ((u32)(4 + (v & 3)) << 4);           // byte v of x, byte v of y
                const u32 t01 = __byte_perm(x0, x1, sel), t23 = __byte_perm(x2, x3, sel);
                *reinterpret_cast<u32*>(dst + v * NC * RS) = __byte_perm(t01, t23, 0x5410);
            }
        }
    };

    int acc[2 * ND - 1][2][4];
#pragma unroll
    for (int w = 0; w < 2 * ND - 1; ++w)
#pragma unroll
        for (int nt = 0; nt < 2; ++nt)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[w][nt][j] = 0;

    issue_A(0, 0);
    fetch_B(0);
    stash_B(0);
    cp_async_commit_wait_all();
    __syncthreads();
    const int kchunks = a.phi / KC;
    for (int kt = 0; kt < kchunks; ++kt) {
        const int buf = kt & 1;
        if (kt + 1 < kchunks) { issue_A((kt + 1) * KC, buf ^ 1); fetch_B((kt + 1) * KC); }
        const unsigned char* pa = sA + buf * A_STAGE + (wm * 16 + g) * RS + tig * 4;
        const unsigned char* pb = sB + buf * B_STAGE + (wn * 16 + g) * RS + tig * 4;
#pragma unroll
        for (int ks = 0; ks < KC / 32; ++ks) {
            u32 af[ND][4];
#pragma unroll
            for (int u = 0; u < ND; ++u) {
                const unsigned char* p = pa + u * MC * RS + ks * 32;
                af[u][0] = *reinterpret_cast<const u32*>(p);
                af[u][1] = *reinterpret_cast<const u32*>(p + 8 * RS);
                af[u][2] = *reinterpret_cast<const u32*>(p + 16);
                af[u][3] = *reinterpret_cast<const u32*>(p + 8 * RS + 16);
            }
#pragma unroll
            for (int v = 0; v < ND; ++v) {
                const unsigned char* p = pb + v * NC * RS + ks * 32;
                const u32 b00 = *reinterpret_cast<const u32*>(p), b01 = *reinterpret_cast<const u32*>(p + 16);
                const u32 b10 = *reinterpret_cast<const u32*>(p + 8 * RS), b11 = *reinterpret_cast<const u32*>(p + 8 * RS + 16);
#pragma unroll
                for (int u = 0; u < ND; ++u) {
                    mma_u8(acc[u + v][0], af[u], b00, b01);
                    mma_u8(acc[u + v][1], af[u], b10, b11);
                }
            }
        }
        if (kt + 1 < kchunks) { stash_B(buf ^ 1); cp_async_commit_wait_all(); }
        __syncthreads();
    }

    // ---- epilogue: recombine the digit sums, one Barrett per output, 16-byte stores ---------------------
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
        const int c = c0 + wn * 16 + nt * 8 + tig * 2;
#pragma unroll
        for (int half = 0; half < 2; ++half) {
            const int o = o0 + wm * 16 + g + half * 8;
            u64 r[2];
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                u64 lo = 0, hi = 0;
#pragma unroll
                for (int w = 0; w < 2 * ND - 1; ++w) {
                    const u64 s = (u64)(u32)acc[w][nt][half * 2 + e];
                    const int sh = 8 * w;
                    if (sh < 64) {
                        const u64 x = s << sh;
                        lo += x; hi += (lo < x);
                        if (sh > 32) hi += s >> (64 - sh);
                    } else {
                        hi += s << (sh - 64);
                    }
                }
                r[e] = reduce128(lo, hi, q, mu.x, mu.y);
            }
            const size_t off = poly_set_addr(a.out_layout, o, lrel, c, a.n, a.cols, a.limbs);
            *reinterpret_cast<ulonglong2*>(a.out + off) = make_ulonglong2(r[0], r[1]);
            if (a.out2) *reinterpret_cast<ulonglong2*>(a.out2 + off) = make_ulonglong2(r[0], r[1]);
        }
    }
}

template <int SRC, int ND>
static cudaError_t launch_nd(const WcrtArgs& a, cudaStream_t st) {
    constexpr int smem = 2 * ND * (MC + NC) * RS;
    {   // set on every call: cheap, and right for any device / thread (no process-wide one-shot flag)
        cudaError_t e = cudaFuncSetAttribute(wcrt_tc_kernel<SRC, ND>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
    }
    dim3 grid(a.cols / NC, a.phi / MC, a.lzn);
    wcrt_tc_kernel<SRC, ND><<<grid, TCT, smem, st>>>(a);
    return cudaGetLastError();
}

template <int SRC>
static cudaError_t launch_src_tc(const WcrtArgs& a, int nd, cudaStream_t st) {
    switch (nd) {
        case 4: return launch_nd<SRC, 4>(a, st);
        case 5: return launch_nd<SRC, 5>(a, st);
        case 6: return launch_nd<SRC, 6>(a, st);
        default: return cudaErrorInvalidValue;
    }
}

bool wcrt_tc_eligible(const WcrtArgs& a) {
    return a.tabd && a.h_ndig && a.phi % MC == 0 && a.phi % KC == 0 && a.cols % NC == 0 && a.n % 2 == 0;
}

// one launch over the relative limbs [a.lz0, a.lz0 + a.lzn) that share the digit count nd (4, 5 or 6)
cudaError_t wcrt_tc_launch(const WcrtArgs& a, int nd, cudaStream_t st) {
    switch (a.source) {
        case kSrcMem:     return launch_src_tc<kSrcMem>(a, nd, st);
        case kSrcI64:     return launch_src_tc<kSrcI64>(a, nd, st);
        case kSrcQuantRe: return launch_src_tc<kSrcQuantRe>(a, nd, st);
        case kSrcQuantIm: return launch_src_tc<kSrcQuantIm>(a, nd, st);
        case kSrcUniform: return launch_src_tc<kSrcUniform>(a, nd, st);
        case kSrcGauss:   return launch_src_tc<kSrcGauss>(a, nd, st);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace mfhe
