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
//   PERSISTENT CTAs (one per SM, 21 warps; 17 for ND = 6) walk over the 128 (o) x UN (c) tiles, UN = 48 columns (32 for
//   ND = 6; written 32 in the formulas above), the last column block of a limb ragged; K (phi) in chunks of 64 through an
//   NS-stage ring in shared memory that runs on across tile boundaries, so the next tile's operands are
//   already landing while the current tile is in its epilogue.
//   last warp  one thread issues the tcgen05.mma's (commit -> the stage's "empty" barrier, last chunk -> "tmem
//              full") and, while they run, refills the stage released by the previous chunk with table digits:
//              the table is stored pre-tiled in the UMMA canonical (no-swizzle, K-major, 8 x 16-byte core
//              matrices) layout, so one 8 KiB cp.async.bulk per digit plane and chunk lands a ready-to-use A
//              tile; completion through an mbarrier (complete_tx).
//   warps 0-7  produce B' (two groups of 4 warps on alternate chunks): fused source (memory in either layout /
//              int64 / quantiser / samplers) -> canonical u64 residues, LDG.128 of two columns x four
//              consecutive i per item -> PRMT byte transposes -> K-major core matrices, fence.proxy.async,
//              mbarrier arrive.
//   warps 8..  epilogue (4 lane quarters x UN/16 column parts): wait for the tile's last commit, tcgen05.ld the 2ND-1 partial sums of 8 columns,
//              recombine to 128 bits, ONE Barrett reduction per output, 16-byte stores in either layout; TMEM is
//              handed back to the issuer as soon as the last accumulator column is in registers.  No zero-fill:
//              the first K step of a tile is issued as overwrite / accumulate pieces.
#include <cstdio>
#include "ctx.h"
#include "wcrt_common.cuh"

#ifndef MFHE_UMMA_PREFETCH
#define MFHE_UMMA_PREFETCH 1      // chunks ahead of the loads that the producers prefetch into L2 (0.362 -> 0.355 ms)
#endif
#ifndef MFHE_UMMA_LOADER_WARP
#define MFHE_UMMA_LOADER_WARP 0      // 1: the table stream gets a warp of its own.  Measured (profiles/r2_11_wcrt_limiters.txt): barrier-only
                                     // skeleton 0.205 -> 0.193 ms, product kernel 0.354 -> 0.361 ms: the issuing thread is not the limiter
#endif
#ifndef MFHE_UMMA_DBG
#define MFHE_UMMA_DBG 0      // timing experiments only: 1 = no table copies, 2 = no input loads, 4 = no epilogue arithmetic,
                             // 16 = producers skip the digit stores, 32 = epilogue skips the TMEM reads, 64 = epilogue keeps TMEM until its stores are done
#endif

namespace mfhe {

namespace {
constexpr int UM = 128, UKC = 64;
constexpr int A_PLANE = UM * UKC;            // 8 KiB: one digit plane of one A chunk

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
__device__ __forceinline__ void umma_i8(u32 d_tmem, u64 adesc, u64 bdesc, u32 idesc, u32 accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(u32 bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld8(u32 taddr, u32 (&r)[8]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld4(u32 taddr, u32 (&r)[4]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(taddr));
}
// 16 lanes x 256 bits: thread T gets lane T/4 (r0, r1) and lane T/4 + 8 (r2, r3), columns 2 (T%4), 2 (T%4) + 1 of the
// 8-column group (cute Copy_Traits<SM100_TMEM_LOAD_16dp256b1x>): four threads share a row, so the epilogue's
// 16-byte stores of one instruction form 64-byte row segments instead of 32 scattered 16-byte pieces
__device__ __forceinline__ void tmem_ld_16x256(u32 taddr, u32 (&r)[4]) {
    asm volatile("tcgen05.ld.sync.aligned.16x256b.x1.b32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(taddr));
}
__device__ __forceinline__ void tmem_st8_zero(u32 taddr) {
    const u32 z = 0;
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%1,%1,%1,%1,%1,%1,%1};" ::"r"(taddr), "r"(z) : "memory");
}
}  // namespace

template <int ND, int SRC> struct UmmaCfg {
    // Tile width.  A 128-row kind::i8 instruction costs ~130-160 clk whatever its N (profiles/r1_10_wcrt_limiters.txt,
    // third series), so the instruction is made as wide as TMEM allows: N' = ND * UN <= 256, (2ND-1) * UN <= 512.
    static constexpr int UN = (ND >= 6) ? 40 : 48;
    static constexpr int B_PLANE = UN * UKC;                            // one input digit plane of one chunk
    static constexpr int NCG = UN / 8;                                  // 8-column groups
    static constexpr int PROD_ROUNDS = (2 * NCG + 3) / 4;               // producer items per thread and chunk
    static constexpr int EPI_PARTS = (UN + 15) / 16;                    // epilogue: 16 columns per warp (the last part may hold 8)
    static constexpr int EPI_WARPS = 4 * EPI_PARTS;
    // producer groups of 4 warps, each taking every PG-th chunk: plain memory input keeps up with one group (17 warps,
    // 96 registers, no spills); the computing sources (int64 reduction, samplers) need two (21 warps, 80 registers)
    static constexpr int PG = (SRC == kSrcMem) ? 1 : 2, PROD_WARPS = 4 * PG;
    static constexpr int ISSUE_WARP = PROD_WARPS + EPI_WARPS;
    // MFHE_UMMA_LOADER_WARP = 1 gives the table stream a warp of its own instead of the issuing thread refilling the ring
    // (an experiment on whether the ONE issuing thread's serial work per chunk limits the MMA stream: it does not)
    static constexpr int LOADER_WARP = ISSUE_WARP + 1;
    static constexpr int THREADS = (ISSUE_WARP + 1 + MFHE_UMMA_LOADER_WARP) * 32;
    static constexpr int NS = (ND <= 5) ? 4 : 3;                       // ring stages
    static constexpr int STAGE = ND * (A_PLANE + B_PLANE);
    // first K step of a tile: if the accumulate / overwrite pieces would not be multiples of 16 columns (ND = 6,
    // UN = 40) the upper positions are instead cleared by one extra instruction with an all-zero A tile
    static constexpr bool ZERO_FIRST = ((ND - 1) * UN) % 16 != 0;
    static constexpr int ZPAD = ZERO_FIRST ? A_PLANE : 0;
    static constexpr int SMEM = NS * STAGE + ZPAD + 1024;               // + barriers / TMEM pointer, alignment slack
    static constexpr int ACC_COLS = (2 * ND - 1) * UN;
    static_assert(ACC_COLS <= 512 && ND * UN <= 256 && (ND * UN) % 16 == 0 && UN % 8 == 0, "tile width");
    // kind::i8 instruction descriptor: D = s32, A = B = unsigned 8 bit, both K-major, 128 x N
    __host__ __device__ static constexpr u32 idesc(int n) { return (2u << 4) | ((u32)(n >> 3) << 17) | ((u32)(UM >> 4) << 24); }
};

// tile id -> (relative limb, column block, output block); the output block runs fastest so that the CTAs
// working at the same time share the table panels in L2 and read every input column block once from HBM
struct TileIdx { int lz, cblk, ot; };
__device__ __forceinline__ TileIdx tile_decode(int tile, int nOT, int nCB) {
    TileIdx r; r.ot = tile % nOT; const int x = tile / nOT; r.cblk = x % nCB; r.lz = x / nCB; return r;
}

template <int SRC, int ND>
__global__ void __launch_bounds__(UmmaCfg<ND, SRC>::THREADS, 1)
wcrt_umma_kernel(const WcrtArgs a, int ntiles) {
    using C = UmmaCfg<ND, SRC>;
    constexpr int UN = C::UN, B_PLANE = C::B_PLANE;
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* smem = reinterpret_cast<unsigned char*>(((uintptr_t)smem_raw + 127) & ~(uintptr_t)127);
    unsigned char* zpad = smem + C::NS * C::STAGE;        // all-zero A tile (ZERO_FIRST only)
    unsigned char* ctrl = zpad + C::ZPAD;                 // barriers: fullA[NS], fullB[NS], empty[NS], tmem_full, tmem_empty
    const u32 bar0 = smem_u32(ctrl);
    auto bar_fullA = [&](int s) { return bar0 + 8u * s; };
    auto bar_fullB = [&](int s) { return bar0 + 8u * (C::NS + s); };
    auto bar_empty = [&](int s) { return bar0 + 8u * (2 * C::NS + s); };
    const u32 bar_tfull = bar0 + 8u * (3 * C::NS), bar_tempty = bar0 + 8u * (3 * C::NS + 1);
    volatile u32* tmem_slot = reinterpret_cast<volatile u32*>(ctrl + 8 * (3 * C::NS + 2));

    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int kchunks = a.phi / UKC, nOT = a.phi / UM, nCB = (a.cols + UN - 1) / UN;     // the last column block may be ragged
    const int my_tiles = (ntiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
    const int my_chunks = my_tiles * kchunks;

    if (t == 0) {
        for (int s = 0; s < C::NS; ++s) { mbar_init(bar_fullA(s), 1); mbar_init(bar_fullB(s), 128); mbar_init(bar_empty(s), 1); }
        mbar_init(bar_tfull, 1); mbar_init(bar_tempty, C::EPI_WARPS);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if constexpr (C::ZERO_FIRST) {
        for (int i = t; i < C::ZPAD / 16; i += C::THREADS) reinterpret_cast<uint4*>(zpad)[i] = make_uint4(0, 0, 0, 0);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
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

    // table digits of chunk g -> its ring stage: one bulk copy per digit plane, completion on the stage's fullA barrier
    auto issue_A = [&](int g) {
        const int s = g % C::NS, kt = g % kchunks;
        const TileIdx ti = tile_decode((int)blockIdx.x + (g / kchunks) * (int)gridDim.x, nOT, nCB);
        const unsigned char* base = a.tabt + (size_t)(a.limb0 + a.lz0 + ti.lz) * 8 * a.phi * a.phi;
#if MFHE_UMMA_DBG & 1
        mbar_arrive(bar_fullA(s)); return;
#endif
        mbar_expect_tx(bar_fullA(s), ND * A_PLANE);
        const u32 dst = smem_u32(smem + s * C::STAGE);
#pragma unroll
        for (int u = 0; u < ND; ++u)
            bulk_g2s(dst + u * A_PLANE,
                     base + ((size_t)u * a.phi * a.phi) + ((size_t)ti.ot * kchunks + kt) * A_PLANE, A_PLANE, bar_fullA(s));
    };

    if (MFHE_UMMA_LOADER_WARP && warp == C::LOADER_WARP) {
        // ===== one thread: streams the table digits, one stage ahead of nothing but the ring itself =====
        if (lane == 0) {
            for (int g = 0; g < my_chunks; ++g) {
                if (g >= C::NS) mbar_wait(bar_empty(g % C::NS), ((g / C::NS) - 1) & 1);
                issue_A(g);
            }
        }
        __syncwarp();
    } else if (warp == C::ISSUE_WARP) {
        // ===== one thread: issues the MMAs (and, without a loader warp, refills the ring with table digits) =====
        if (lane == 0) {
            if (!MFHE_UMMA_LOADER_WARP)
                for (int g = 0; g < C::NS && g < my_chunks; ++g) issue_A(g);
            const u32 smem0 = smem_u32(smem);
            const u32 idesc_full = C::idesc(ND * UN);
            int s = 0, kt = 0, it = 0;
            u32 par = 0;
#ifdef MFHE_UMMA_TIMING
            long long w_te = 0, w_a = 0, w_b = 0, w_refill = 0, t_all = clock64(), tq;
#define WT(acc, stmt) do { tq = clock64(); stmt; acc += clock64() - tq; } while (0)
#else
#define WT(acc, stmt) stmt
#endif
            for (int g = 0; g < my_chunks; ++g) {
                if (kt == 0 && it > 0) WT(w_te, mbar_wait(bar_tempty, (it - 1) & 1));      // epilogue has drained the accumulators
                WT(w_a, mbar_wait(bar_fullA(s), par));
                WT(w_b, mbar_wait(bar_fullB(s), par));
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const u32 sa = smem0 + (u32)s * (u32)C::STAGE, sb = sa + ND * A_PLANE;
#pragma unroll
                for (int ks = 0; ks < UKC / 32; ++ks) {
                    const u64 bdesc = umma_desc(sb + ks * 256, 128, 512);
                    if (kt == 0 && ks == 0) {
                        // first K step of a tile: nothing to accumulate onto.  Position block u+ND-1 is new for
                        // every u, the others were started by u-1: split the instruction accordingly.
                        if constexpr (C::ZERO_FIRST) {
                            umma_i8(tmem + (u32)((ND - 1) * UN), umma_desc(smem_u32(zpad), 128, 512), bdesc, idesc_full, 0);
                            umma_i8(tmem, umma_desc(sa, 128, 512), bdesc, idesc_full, 0);
#pragma unroll
                            for (int u = 1; u < ND; ++u)
                                umma_i8(tmem + (u32)(u * UN), umma_desc(sa + u * A_PLANE, 128, 512), bdesc, idesc_full, 1);
                        } else {
                        umma_i8(tmem, umma_desc(sa, 128, 512), bdesc, idesc_full, 0);
#pragma unroll
                        for (int u = 1; u < ND; ++u) {
                            const u64 adesc = umma_desc(sa + u * A_PLANE, 128, 512);
                            umma_i8(tmem + (u32)(u * UN), adesc, bdesc, C::idesc((ND - 1) * UN), 1);
                            umma_i8(tmem + (u32)((u + ND - 1) * UN), adesc, umma_desc(sb + (ND - 1) * B_PLANE, 128, 512),
                                    C::idesc(UN), 0);
                        }
                        }
                    } else {
#ifdef MFHE_UMMA_DBGN      // operand-fetch experiment: every instruction with N = MFHE_UMMA_DBGN (results are garbage)
#pragma unroll
                        for (int u = 0; u < ND; ++u)
                            umma_i8(tmem + (u32)(u * UN), umma_desc(sa + u * A_PLANE + ks * 256, 128, 512), bdesc,
                                    C::idesc(MFHE_UMMA_DBGN), 1);
#else
#pragma unroll
                        for (int u = 0; u < ND; ++u)
                            umma_i8(tmem + (u32)(u * UN), umma_desc(sa + u * A_PLANE + ks * 256, 128, 512), bdesc,
                                    idesc_full, 1);
#endif
                    }
                }
                umma_commit(bar_empty(s));
                if (kt == kchunks - 1) umma_commit(bar_tfull);
                if (!MFHE_UMMA_LOADER_WARP) {
                    // while chunk g runs, refill the stage chunk g-1 has just released
                    if (g >= 1 && g - 1 + C::NS < my_chunks) {
                        WT(w_refill, mbar_wait(bar_empty((g - 1) % C::NS), ((g - 1) / C::NS) & 1));
                        issue_A(g - 1 + C::NS);
                    }
                }
                if (++s == C::NS) { s = 0; par ^= 1; }
                if (++kt == kchunks) { kt = 0; ++it; }
            }
#ifdef MFHE_UMMA_TIMING
            if (blockIdx.x == 7)
                printf("issuer (clk, %d chunks): total %lld  wait tmem-empty %lld  wait table (fullA) %lld  wait digits (fullB) %lld  wait previous chunk (refill) %lld\n",
                       my_chunks, clock64() - t_all, w_te, w_a, w_b, w_refill);
#endif
        }
        __syncwarp();
    } else if (warp < C::PROD_WARPS) {
        // ===== input-digit producers: two groups of 128 threads take alternate chunks.  A chunk of one digit stack is
        // NCG column groups x 2 halves of the 16 four-k groups = 2 NCG warp-items; a lane of a warp-item owns (column
        // pair, four consecutive i): 4 LDG.128, PRMT byte transposes, K-major core-matrix stores.  One register set
        // per item: right after an item is converted its registers take the loads of the group's NEXT chunk (two MMA
        // chunk times ahead), so the loads are in flight while the rest is converted and while the stage is awaited. =====
        constexpr int R = C::PROD_ROUNDS, WI = 2 * C::NCG;
        const int grp = warp >> 2, pt = t & 127;
        int ic[R], ikq[R], isw[R];
        bool live[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int id = pt + 128 * r, wi = id >> 5;             // warp-item: column group wi % NCG, k half wi / NCG
            live[r] = wi < WI;
            isw[r] = (id >> 4) & 1;
            ic[r] = (wi % C::NCG) * 8 + (id & 3) * 2;
            ikq[r] = ((id >> 2) & 3) | (isw[r] << 2) | ((wi / C::NCG) << 3);
        }
        const size_t istride = (size_t)a.limbs * a.cols;
        // fetch-side cursor over this group's chunks
        int f_kt = grp, f_it = 0;
        size_t f_off[R];
        bool f_ok[R];
        int f_lrel = 0, f_l = 0, f_c0 = 0;
        u64 f_q = 0, f_mu = 0;
        auto f_tile = [&]() {
            const TileIdx ti = tile_decode((int)blockIdx.x + f_it * (int)gridDim.x, nOT, nCB);
            f_lrel = a.lz0 + ti.lz; f_l = a.limb0 + f_lrel; f_c0 = ti.cblk * UN;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                f_ok[r] = live[r] && f_c0 + ic[r] < a.cols;        // ragged last column block: zeros, no loads
                f_off[r] = 0;
                if constexpr (SRC == kSrcMem)
                    if (f_ok[r]) f_off[r] = poly_set_addr(a.in_layout, 0, f_lrel, f_c0 + ic[r], a.n, a.cols, a.limbs);
            }
            if constexpr (SRC != kSrcMem) { f_q = a.moduli[f_l]; f_mu = a.mu[f_l].y; }
        };
        auto fetch_item = [&](int r, u64 (&rb)[8]) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int i = f_kt * UKC + 4 * ikq[r] + j;
#if MFHE_UMMA_DBG & 2
                rb[2 * j] = (u64)(i + f_lrel); rb[2 * j + 1] = (u64)i;
#else
                if (!f_ok[r]) { rb[2 * j] = 0; rb[2 * j + 1] = 0; }
                else if constexpr (SRC == kSrcMem) {
                    const u64* src = (const u64*)a.in + (size_t)i * istride + f_off[r];
                    const ulonglong2 v = __ldg(reinterpret_cast<const ulonglong2*>(src));
                    rb[2 * j] = v.x; rb[2 * j + 1] = v.y;
#if MFHE_UMMA_PREFETCH
                    // the rows this item needs MFHE_UMMA_PREFETCH chunks later (same tile): start them towards L2 now
                    if (f_kt + MFHE_UMMA_PREFETCH * C::PG < kchunks)
                        asm volatile("prefetch.global.L2 [%0];" ::"l"(src + (size_t)(MFHE_UMMA_PREFETCH * C::PG * UKC) * istride));
#endif
                } else {
                    rb[2 * j] = wcrt_load_one<SRC>(a, i, f_l, f_c0 + ic[r], f_q, f_mu);
                    rb[2 * j + 1] = wcrt_load_one<SRC>(a, i, f_l, f_c0 + ic[r] + 1, f_q, f_mu);
                }
#endif
            }
        };
        auto f_advance = [&]() {
            f_kt += C::PG;
            if (f_kt >= kchunks) { f_kt = grp; ++f_it; }
        };
        auto stash_item = [&](unsigned char* sB, int r, const u64 (&rb)[8]) {
            // rows (v*UN + c), (v*UN + c + 1) of B', bytes [4kq, 4kq+4): core matrix (row/8, kq/4), row%8, (kq%4)*4.
            // Lanes with isw set store the odd row first, so that one store instruction covers all 32 banks.
            unsigned char* dst = sB + (ic[r] >> 3) * 512 + (ikq[r] >> 2) * 128 + (ic[r] & 7) * 16 + (ikq[r] & 3) * 4;
            unsigned char* d1 = dst + (isw[r] ? 16 : 0);
            unsigned char* d2 = dst + (isw[r] ? 0 : 16);
#pragma unroll
            for (int v = 0; v < ND; ++v) {
                const int sh = (v < 4) ? 0 : 32;
                const u32 sel = (u32)(v & 3) | ((u32)(4 + (v & 3)) << 4);
                u32 wd[2];
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const u32 x0 = (u32)(rb[0 + e] >> sh), x1 = (u32)(rb[2 + e] >> sh);
                    const u32 x2 = (u32)(rb[4 + e] >> sh), x3 = (u32)(rb[6 + e] >> sh);
                    wd[e] = __byte_perm(__byte_perm(x0, x1, sel), __byte_perm(x2, x3, sel), 0x5410);
                }
                *reinterpret_cast<u32*>(d1 + v * B_PLANE) = isw[r] ? wd[1] : wd[0];
                *reinterpret_cast<u32*>(d2 + v * B_PLANE) = isw[r] ? wd[0] : wd[1];
            }
        };
        u64 rb[R][8];
        if (grp < my_chunks) {
            f_tile();
#pragma unroll
            for (int r = 0; r < R; ++r) fetch_item(r, rb[r]);
            f_advance();
        }
        for (int g = grp; g < my_chunks; g += C::PG) {           // kchunks is even, so is my_chunks
            const int s = g % C::NS;
            const bool more = g + C::PG < my_chunks;
            if (g >= C::NS) mbar_wait(bar_empty(s), ((g / C::NS) - 1) & 1);
            unsigned char* sB = smem + s * C::STAGE + ND * A_PLANE;
            if (more && f_kt == grp) f_tile();                   // (only reads of the OLD f_ok / ic remain below: see order)
#pragma unroll
            for (int r = 0; r < R; ++r) {
#if !(MFHE_UMMA_DBG & 16)
                if (live[r]) stash_item(sB, r, rb[r]);
#endif
                if (more) fetch_item(r, rb[r]);
            }
            if (more) f_advance();
#if !(MFHE_UMMA_DBG & 8)
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
#endif
            mbar_arrive(bar_fullB(s));
        }
    } else {
        // ===== epilogue (4 x UN/16 warps): TMEM lanes 32*(warp%4).. = output rows, 16-column part (warp-8)/4.  Sub-group sg =
        // (16-lane half, 8-column group): thread holds rows lane/4 and lane/4 + 8, columns 2 (lane%4) + {0, 1} =====
        const u32 tlane = tmem + ((u32)(32 * (warp & 3)) << 16);
        for (int it = 0; it < my_tiles; ++it) {
            const TileIdx ti = tile_decode((int)blockIdx.x + it * (int)gridDim.x, nOT, nCB);
            const int lrel = a.lz0 + ti.lz, l = a.limb0 + lrel;
            const u64 q = a.moduli[l];
            const ulonglong2 mu = a.mu[l];
            const int o = ti.ot * UM + 32 * (warp & 3) + (lane >> 2);
            mbar_wait(bar_tfull, it & 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            // phase 1 (holds TMEM): fold the 2ND-1 partial sums of each of this thread's 16 outputs into
            //   lo = sum_{w<5} S_w 2^(8w)  (< 2^61)   and   up = sum_{5<=w<10} S_w 2^(8(w-5))  (< 2^61)   [+ S_10 for ND = 6],
            // 4 columns at a time to keep the register footprint small
            const int ccol = ((warp - C::PROD_WARPS) >> 2) * 16;
            u64 vlo[16], vup[16];
            u32 vtop[(2 * ND - 1 > 10) ? 16 : 1];         // position 10 (scale 2^80) exists for ND = 6 only
#pragma unroll
            for (int sg = 0; sg < 4; ++sg) {
                if (ccol + 8 * (sg & 1) >= UN) continue;          // last part of a tile whose width is not a multiple of 16
                u32 acc[2 * ND - 1][4];
#if MFHE_UMMA_DBG & 32
#pragma unroll
                for (int w = 0; w < 2 * ND - 1; ++w) { acc[w][0] = lane + w; acc[w][1] = lane ^ w; acc[w][2] = sg; acc[w][3] = it; }
#else
#pragma unroll
                for (int w = 0; w < 2 * ND - 1; ++w)
                    tmem_ld_16x256(tlane + ((u32)(16 * (sg >> 1)) << 16) + (u32)(w * UN + ccol + 8 * (sg & 1)), acc[w]);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#endif
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    u64 lo = acc[0][e], up = acc[5 < 2 * ND - 1 ? 5 : 0][e];
#pragma unroll
                    for (int w = 1; w < 5 && w < 2 * ND - 1; ++w) lo += (u64)acc[w][e] << (8 * w);
#pragma unroll
                    for (int w = 6; w < 2 * ND - 1 && w < 10; ++w) up += (u64)acc[w][e] << (8 * (w - 5));
                    if (2 * ND - 1 <= 5) up = 0;
                    vlo[4 * sg + e] = lo; vup[4 * sg + e] = up;
                    if constexpr (2 * ND - 1 > 10) vtop[4 * sg + e] = acc[10][e];
                }
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_tempty);                 // the issuer may start the next tile
#if MFHE_UMMA_DBG & 4
            if (vlo[0] == 0x12345678u) a.out[0] = vup[1];
            continue;
#endif
            // phase 2 (overlaps the next tile's MMAs): value = lo + up * 2^40, one Barrett reduction, 16-byte stores.
            // The value is the exact integer sum_i tab*in < phi q^2 < 2^(k+62) (k = bits of q <= 48 here), so a
            // ONE-word Barrett suffices: qhat = mulhi(v >> (k-2), floor(2^(k+62) / q)) is within 2 of the quotient
            // (HAC 14.42 with the bounds v / 2^(k+62) < 1/2 and 2^(k-2) / q <= 1/2), r = v - qhat q < 3q.
            const int kb = 64 - __clzll((long long)q);
            const u64 mu1 = (mu.y << (kb - 2)) | (mu.x >> (66 - kb));        // floor(2^128 / q) >> (66 - k)
            const u64 q2 = 2 * q;
#pragma unroll
            for (int e = 0; e < 16; e += 2) {
                u64 r[2];
#pragma unroll
                for (int k = 0; k < 2; ++k) {
                    const u64 x = vup[e + k] << 40;
                    const u64 lo = vlo[e + k] + x;
                    u64 hi = (vup[e + k] >> 24) + (lo < x);
                    if constexpr (2 * ND - 1 > 10) hi += (u64)vtop[e + k] << 16;
                    const u64 x1 = (lo >> (kb - 2)) | (hi << (66 - kb));
                    const u64 rr = lo - __umul64hi(x1, mu1) * q;
                    r[k] = csub(csub(rr, q2), q);
                }
                // e = 4 sg + 2 p: rows o + 16 (sg / 2) + 8 p, column pair 8 (sg % 2) + 2 (lane % 4)
                const size_t off = poly_set_addr(a.out_layout, o + 16 * (e >> 3) + 8 * ((e >> 1) & 1), lrel,
                                                 ti.cblk * UN + ccol + 8 * ((e >> 2) & 1) + 2 * (lane & 3), a.n, a.cols, a.limbs);
                if (ccol + 8 * ((e >> 2) & 1) < UN &&
                    ti.cblk * UN + ccol + 8 * ((e >> 2) & 1) + 2 * (lane & 3) < a.cols) {      // ragged last column block
                    *reinterpret_cast<ulonglong2*>(a.out + off) = make_ulonglong2(r[0], r[1]);
                    if (a.out2) *reinterpret_cast<ulonglong2*>(a.out2 + off) = make_ulonglong2(r[0], r[1]);
                }
            }
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
    using C = UmmaCfg<ND, SRC>;
    {   // set on every call: cheap, and right for any device / thread (no process-wide one-shot flag)
        cudaError_t e = cudaFuncSetAttribute(wcrt_umma_kernel<SRC, ND>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
        if (e != cudaSuccess) return e;
    }
    int sms = 0, dev = 0;
    cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int ntiles = ((a.cols + C::UN - 1) / C::UN) * (a.phi / UM) * a.lzn;
    wcrt_umma_kernel<SRC, ND><<<ntiles < sms ? ntiles : sms, C::THREADS, C::SMEM, st>>>(a, ntiles);
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
    return a.tabt && a.h_ndig && a.phi % UM == 0 && a.phi % (2 * UKC) == 0 && a.cols % 16 == 0 && a.n % 2 == 0;
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
