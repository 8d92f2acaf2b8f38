// bconv_umma.cuh -- the base conversion's matrix product on the 5th-generation tensor cores (included by bconv.cu).
//
//     out_j[c] = ( sum_i y_i[c] * C[j][i] ) mod p_j,   y_i = x_i * (Q/q_i)^-1 mod q_i,   C[j][i] = Q/q_i mod p_j
// (ph/src/rns_bconv.cu:22-29 bconv_mult, :109-131 bconv_matmul) IS a GEMM: [coefficients x L_in] x [L_in x L_out], and on the
// CUDA cores it is multiplier-bound (one 64 x 64 -> 128 multiply-accumulate per (i, j, c): 0.7-1.2 T modular MAC/s,
// 8-25 % of HBM at config 5's 48 -> 16 limbs).  Here both factors are cut into unsigned 8-bit digits (y: NDA =
// ceil(bits(q)/8), C: NDB = ceil(bits(p)/8)), a tile is M = 128 coefficients x 16 output limbs with K = the input limbs of
// the window (padded to a multiple of 32), and -- the digit-stack form of wcrt_umma.cu -- the B operand of an instruction
// is the stack of C's NDB digit planes (N' = 16 NDB rows), the instruction for y's digit u accumulates at TMEM column 16 u:
// NDA instructions per K step put every digit pair (u, v) into position u + v.  The NDA + NDB - 1 partial sums (each below
// 2^26: exact in s32) are recombined to the 128-bit integer sum_i y_i C[j][i] and reduced once -- the same residue as the
// CUDA-core kernel's lazy 128-bit accumulation, word for word.
//
// CTA (256 threads, 2 per SM: 256 TMEM columns each), persistent over (batch item, 128-coefficient block) tiles:
//   stage   x -> y (one Shoup product per word) -> digits -> NDA planes [128 c][KP k] in the UMMA canonical K-major layout;
//           lanes run along the coefficients (coalesced 256-byte segments per limb), a thread packs four consecutive limbs
//           into one 32-bit store per plane, the 4-limb slot rotated by the row group (conflict-free)
//   MMAs    one thread: KP / 32 K steps x NDA instructions of 128 x 16 NDB x 32, commit to an mbarrier
//   fold    thread = coefficient (TMEM lane), 8 of the 16 output limbs: 32x32b loads, recombination, one-word or 128-bit
//           Barrett, and a store per output limb that is coalesced across the warp (32 consecutive coefficients)
// The matrix digits C' are rebuilt in shared memory per output tile of 16 limbs (a few hundred words).
#pragma once

namespace {
constexpr int UB_M = 128, UB_UN = 16, UB_TH = 256;
#ifndef MFHE_UB_G
#define MFHE_UB_G 2      // A/B on B200 (48 / 24 / 12 input limbs): 2 -> 0.669 / 0.498 / 0.401 ms, 4 -> 0.686 / 0.502 / 0.418, 8 -> 0.675 / 0.557 / 0.465
#endif
constexpr int UB_G = MFHE_UB_G;                          // 4-limb groups (x 4 loads) a staging thread keeps in flight

__device__ __forceinline__ u32 ub_smem_u32(const void* p) { return (u32)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void ub_mbar_init(u32 bar, u32 count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void ub_mbar_wait(u32 bar, u32 parity) {
    u32 ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    } while (!ok);
}
// K-major, no swizzle: core matrix = 8 rows x 16 B; LBO = 128 (next 16 k), SBO = bytes of an 8-row group = 8 * KP
__device__ __forceinline__ u64 ub_desc(u32 saddr, u32 sbo) {
    return (u64)((saddr & 0x3FFFFu) >> 4) | ((u64)(128 >> 4) << 16) | ((u64)(sbo >> 4) << 32) | (1ull << 46);
}
__device__ __forceinline__ u32 ub_idesc(int n) { return (2u << 4) | ((u32)(n >> 3) << 17) | ((u32)(UB_M >> 4) << 24); }
__device__ __forceinline__ void ub_mma(u32 d_tmem, u64 adesc, u64 bdesc, u32 id, u32 accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(id), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void ub_commit(u32 bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void ub_ld4(u32 taddr, u32 (&r)[4]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(taddr));
}
__device__ __forceinline__ void ub_ld8(u32 taddr, u32 (&r)[8]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr));
}
// byte offset of (row, k) in a canonical K-major operand whose rows hold KP bytes
__device__ __forceinline__ u32 ub_canon(u32 row, u32 k, u32 kp) { return (row >> 3) * (8 * kp) + (k >> 4) * 128 + (row & 7) * 16 + (k & 15); }
// s0 + 2^8 s1 + 2^16 s2 + 2^24 s3 + 2^32 s4 for partial sums below 2^28
__device__ __forceinline__ u64 ub_fold5(u32 s0, u32 s1, u32 s2, u32 s3, u32 s4) {
    u64 r = ((u64)s4 << 32) | s0;
    asm("mad.wide.u32 %0, %1, 256, %0;" : "+l"(r) : "r"(s1));
    asm("mad.wide.u32 %0, %1, 65536, %0;" : "+l"(r) : "r"(s2));
    asm("mad.wide.u32 %0, %1, 16777216, %0;" : "+l"(r) : "r"(s3));
    return r;
}

struct UbArgs {
    const u64* in; u64* out; size_t N; int l_in_total, in_start, in_count, l_out;
    const u64* q; const tw_t* qhat_inv; const u64* qhat_mod_p; const u64* p; const ulonglong2* mu_p;
    int nda, ndb, kp;                 // digit counts of y and of the matrix, K padded to a multiple of 32
    int tiles_per_item, ntiles;       // N / 128, batch * N / 128
};

template <bool SC, bool W1>
__global__ void __launch_bounds__(UB_TH, 2)
bconv_umma_kernel(const UbArgs a, const __grid_constant__ BconvScatter sc) {
    extern __shared__ __align__(1024) unsigned char ub_raw[];
    unsigned char* smem = reinterpret_cast<unsigned char*>(((uintptr_t)ub_raw + 127) & ~(uintptr_t)127);
    const int kp = a.kp, nda = a.nda, ndb = a.ndb, npos = nda + ndb - 1;
    const u32 aplane = (u32)UB_M * kp;                       // one digit plane of A
    unsigned char* sA = smem;                                 // [nda][128][kp]
    unsigned char* sB = smem + nda * aplane;                  // [ndb * 16][kp]: digit stack of one output tile
    unsigned char* ctrl = sB + (size_t)ndb * UB_UN * kp;
    const u32 bar = ub_smem_u32(ctrl);
    volatile u32* tmem_slot = reinterpret_cast<volatile u32*>(ctrl + 16);
    u64* spc = reinterpret_cast<u64*>(ctrl + 64);            // [l_out][3]: p_j and floor(2^128 / p_j), read by the fold
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    for (int j = t; j < a.l_out; j += UB_TH) { const ulonglong2 m = a.mu_p[j]; spc[3 * j] = a.p[j]; spc[3 * j + 1] = m.x; spc[3 * j + 2] = m.y; }
    u64* sqc = spc + 3 * a.l_out;                            // [in_count][3]: q_i and {(Q/q_i)^-1, shoup}, read by the staging
    for (int i = t; i < a.in_count; i += UB_TH) {
        const tw_t w = a.qhat_inv[a.in_start + i];
        sqc[3 * i] = a.q[a.in_start + i]; sqc[3 * i + 1] = w.x; sqc[3 * i + 2] = w.y;
    }

    if (t == 0) {
        ub_mbar_init(bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    // K padding (k >= in_count) stays zero for the whole kernel: only real limbs are ever written
    for (u32 i = t; i < (nda * aplane + (u32)ndb * UB_UN * kp) / 16; i += UB_TH) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
    if (warp == 0) {
        __syncwarp();
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" ::"r"(ub_smem_u32((const void*)tmem_slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const u32 tmem = *tmem_slot;
    u32 phase = 0;
    const int n_ot = (a.l_out + UB_UN - 1) / UB_UN;
    const int sc_ = t & 127, half = t >> 7, rot = (sc_ >> 3) & 3;
    int b_tile = -1;                                          // output tile whose digits sB holds

    for (int tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x) {
        const size_t b = (size_t)(tile / a.tiles_per_item), d0 = (size_t)(tile % a.tiles_per_item) * UB_M;
        const u64* x = a.in + b * (size_t)a.in_count * a.N + d0 + sc_;
#ifdef MFHE_UB_TIMING
        long long tk[8]; tk[0] = clock64();
#define UB_TICK(k) tk[k] = clock64()
#else
#define UB_TICK(k)
#endif
        // ---- stage A: y digits of 128 coefficients x in_count limbs; four 4-limb groups (16 loads) in flight per thread ----
        {   // the next tile of this CTA towards L2: 128 coefficients = 8 lines per limb
            const int nt = tile + (int)gridDim.x;
            if (nt < a.ntiles) {
                const u64* xn = a.in + (size_t)(nt / a.tiles_per_item) * (size_t)a.in_count * a.N + (size_t)(nt % a.tiles_per_item) * UB_M;
                for (int e = t; e < 8 * a.in_count; e += UB_TH)
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(xn + (size_t)(e >> 3) * a.N + (e & 7) * 16));
            }
        }
        for (int kb = half; kb < kp / 4; kb += 2 * UB_G) {    // kl = kb, kb + 2, ...: linear 4-limb groups, UB_G of them in flight
            u64 y[UB_G][4];
#pragma unroll
            for (int g = 0; g < UB_G; ++g) {
                const int kl = kb + 2 * g, i0 = 16 * (kl >> 2) + 4 * ((kl & 3) ^ rot);
#pragma unroll
                for (int j = 0; j < 4; ++j) y[g][j] = (kl < kp / 4 && i0 + j < a.in_count) ? x[(size_t)(i0 + j) * a.N] : 0;
            }
#pragma unroll
            for (int g = 0; g < UB_G; ++g) {
                const int kl = kb + 2 * g, i0 = 16 * (kl >> 2) + 4 * ((kl & 3) ^ rot);
                if (kl >= kp / 4 || i0 >= a.in_count) continue;
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (i0 + j < a.in_count) y[g][j] = mul_shoup(y[g][j], tw_t{sqc[3 * (i0 + j) + 1], sqc[3 * (i0 + j) + 2]}, sqc[3 * (i0 + j)]);
                const u32 off = ub_canon(sc_, i0, kp);
#pragma unroll
                for (int v = 0; v < 8; ++v) {
                    if (v >= nda) break;
                    const int sh = v < 4 ? 0 : 32;
                    const u32 sel = (u32)(v & 3) | ((u32)(4 + (v & 3)) << 4);
                    const u32 t01 = __byte_perm((u32)(y[g][0] >> sh), (u32)(y[g][1] >> sh), sel);
                    const u32 t23 = __byte_perm((u32)(y[g][2] >> sh), (u32)(y[g][3] >> sh), sel);
                    *reinterpret_cast<u32*>(sA + v * aplane + off) = __byte_perm(t01, t23, 0x5410);
                }
            }
        }
        UB_TICK(1);
        for (int ot = 0; ot < n_ot; ++ot) {
            // ---- the matrix digits of output tile ot: stack row 16 v + jj, byte k = digit v of C[16 ot + jj][in_start + k] ----
            if (b_tile != ot) {
                for (int e = t; e < UB_UN * a.in_count; e += UB_TH) {
                    const int jj = e / a.in_count, k = e % a.in_count, j = UB_UN * ot + jj;
                    const u64 c = j < a.l_out ? a.qhat_mod_p[(size_t)j * a.l_in_total + a.in_start + k] : 0;
                    for (int v = 0; v < ndb; ++v) sB[ub_canon(UB_UN * v + jj, k, kp)] = (unsigned char)(c >> (8 * v));
                }
                b_tile = ot;
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncthreads();
            UB_TICK(2);
            if (t == 0) {
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const u32 a0 = ub_smem_u32(sA), b0 = ub_smem_u32(sB), sbo = 8u * kp;
                const u32 id_full = ub_idesc(ndb * UB_UN);
                for (int ks = 0; ks < kp / 32; ++ks) {
                    const u64 bdesc = ub_desc(b0 + ks * 256, sbo);
                    if (ks == 0) {
                        // nothing to accumulate onto: position u + ndb - 1 is new for every u, the others were started by u - 1
                        ub_mma(tmem, ub_desc(a0, sbo), bdesc, id_full, 0);
                        for (int u = 1; u < nda; ++u) {
                            const u64 adesc = ub_desc(a0 + u * aplane, sbo);
                            if (ndb > 1) ub_mma(tmem + (u32)(u * UB_UN), adesc, bdesc, ub_idesc((ndb - 1) * UB_UN), 1);
                            ub_mma(tmem + (u32)((u + ndb - 1) * UB_UN), adesc, ub_desc(b0 + (ndb - 1) * (UB_UN / 8) * sbo, sbo), ub_idesc(UB_UN), 0);
                        }
                    } else {
                        for (int u = 0; u < nda; ++u)
                            ub_mma(tmem + (u32)(u * UB_UN), ub_desc(a0 + u * aplane + ks * 256, sbo), bdesc, id_full, 1);
                    }
                }
                ub_commit(bar);
            }
            UB_TICK(3);
            ub_mbar_wait(bar, phase);
            phase ^= 1;
            UB_TICK(4);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            // ---- fold: thread = coefficient d0 + 32 (warp % 4) + lane, output limbs 16 ot + 8 (warp / 4) + 0..7 ----
            const size_t d = d0 + 32 * (warp & 3) + lane;
            const u32 trow = tmem + ((u32)(32 * (warp & 3)) << 16) + (u32)((warp >> 2) * 8);
#pragma unroll
            // the 15 positions in three chunks of five (value = lo5 + 2^40 mid + 2^80 top), all 8 columns of the thread per load
            u64 part[3][8];
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                if (5 * c >= npos) {
#pragma unroll
                    for (int e = 0; e < 8; ++e) part[c][e] = 0;
                    continue;
                }
                u32 acc[5][8];
#pragma unroll
                for (int pp = 0; pp < 5; ++pp) {
                    if (5 * c + pp < npos) ub_ld8(trow + (u32)((5 * c + pp) * UB_UN), acc[pp]);
                    else {
#pragma unroll
                        for (int e = 0; e < 8; ++e) acc[pp][e] = 0;
                    }
                }
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int e = 0; e < 8; ++e) part[c][e] = ub_fold5(acc[0][e], acc[1][e], acc[2][e], acc[3][e], acc[4][e]);
            }
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const int j = UB_UN * ot + 8 * (warp >> 2) + e;
                const u64 xs = part[1][e] << 40;
                const u64 lo = part[0][e] + xs;                                   // value < 2^128 (plan creation), so top < 2^48
                const u64 hi = (part[1][e] >> 24) + (lo < xs) + (part[2][e] << 16);
                if (j < a.l_out) {
                    const u64 r = bconv_fold<W1>(lo, hi, spc[3 * j], make_ulonglong2(spc[3 * j + 1], spc[3 * j + 2]));
                    if constexpr (SC) sc.base[j][b * sc.bstride[j] + d] = r;
                    else              a.out[(b * (size_t)a.l_out + j) * a.N + d] = r;
                }
            }
            UB_TICK(5);
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncthreads();                                  // accumulators drained, sB free for the next output tile
            UB_TICK(6);
#ifdef MFHE_UB_TIMING
            if (blockIdx.x == 5 && t == 0 && tile == 5 + 20 * (int)gridDim.x)
                printf("tile chain (clk): stage %lld  sync %lld  issue %lld  mma-wait %lld  fold %lld  sync %lld  total %lld\n", tk[1] - tk[0],
                       tk[2] - tk[1], tk[3] - tk[2], tk[4] - tk[3], tk[5] - tk[4], tk[6] - tk[5], tk[6] - tk[0]);
#endif
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" ::"r"(tmem) : "memory");
}

inline int ub_digits(uint64_t v) { int b = 0; while (v) { ++b; v >>= 1; } return (b + 7) / 8; }
}  // namespace
