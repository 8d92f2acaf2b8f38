// umma_pair.cu -- does cta_group::2 shorten the W-CRT's tcgen05.mma kind::i8 stream?  
// MMA-only skeleton of wcrt_umma.cu's issuer: per K chunk of 64, 2 K-steps x ND instructions  D[u*UN ...] += A_u * B'^T
// with both operands in shared memory (no-swizzle K-major core matrices, same descriptors as the kernel), a 4-stage ring of
// operand buffers, a commit per chunk, the issuer never more than 4 chunks ahead.  Operand contents are whatever shared
// memory holds: only the instruction stream is timed.
//   PAIR = 0: every CTA issues cta_group::1 instructions of 128 x N x 32 (what the kernel does today, N = 240)
//   PAIR = 1: clusters of two CTAs, the leader issues cta_group::2 instructions of 256 x N x 32: each SM keeps its own 128 rows
//             of A and HALF of B' (N/2 rows); kind::i8 wants N % 32 == 0 at M = 256, so the 240-row digit stack is padded to 256
// Both variants run 148 CTAs in clusters of 2, one per SM.  Reported: ns and SM clocks per instruction, int8 MAC/clk/SM.
// Build + run:  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o umma_pair umma_pair.cu && timeout 60 ./umma_pair
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef unsigned int u32; typedef unsigned long long u64;

constexpr int ND = 5, UN = 48, NS = 4;
constexpr int A_PLANE = 128 * 64, B_BYTES = 256 * 64;            // one digit plane of A per chunk; B' (up to 256 rows) per chunk
constexpr int STAGE = ND * A_PLANE + B_BYTES;                      // 56 KiB
constexpr int SMEM = NS * STAGE + 1024;

__device__ __forceinline__ u32 smem_u32(const void* p) { return (u32)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(u32 bar, u32 count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory"); }
__device__ __forceinline__ void mbar_wait(u32 bar, u32 parity) {
    u32 ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    } while (!ok);
}
__device__ __forceinline__ void mbar_poll(u32 bar, u32 parity) {          // non-suspending form
    u32 ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    } while (!ok);
}
__device__ __forceinline__ void mbar_wait_sleep(u32 bar, u32 parity, unsigned ns) {   // poll, sleep between polls
    u32 ok;
    for (;;) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
        if (ok) break;
        __nanosleep(ns);
    }
}
__device__ __forceinline__ u64 umma_desc(u32 saddr, u32 lbo, u32 sbo) {
    return (u64)((saddr & 0x3FFFFu) >> 4) | ((u64)(lbo >> 4) << 16) | ((u64)(sbo >> 4) << 32) | (1ull << 46);
}
__host__ __device__ constexpr u32 idesc(int m, int n) { return (2u << 4) | ((u32)(n >> 3) << 17) | ((u32)(m >> 4) << 24); }
__device__ __forceinline__ void cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

template <int PAIR, int N>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(32 * 13, 1) skeleton(int chunks, u32* sink, int depth, int tile_chunks) {
    const int flags = (int)(size_t)sink;
    extern __shared__ __align__(1024) unsigned char smem[];
    unsigned char* ctrl = smem + NS * STAGE;
    const u32 bar0 = smem_u32(ctrl);
    volatile u32* tmem_slot = reinterpret_cast<volatile u32*>(ctrl + 128);
    u32 rank;
    asm("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
    const int t = threadIdx.x, warp = t >> 5;
    if (t == 0) {
        for (int s = 0; s < NS; ++s) mbar_init(bar0 + 8u * s, 1);
        mbar_init(bar0 + 8u * NS, 1); mbar_init(bar0 + 8u * (NS + 1), 12);        // tmem full / tmem empty (12 epilogue warps)
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        __syncwarp();
        if (PAIR) {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32((const void*)tmem_slot)) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
        } else {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32((const void*)tmem_slot)) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    cluster_sync();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const u32 tmem = *tmem_slot;

    if (t == 0 && (!PAIR || rank == 0)) {
        const u32 id = idesc(PAIR ? 256 : 128, N);
        for (int g = 0; g < chunks; ++g) {
            const int s = g % NS;
            if (tile_chunks && !(flags & 96) && g % tile_chunks == 0 && g > 0) { if (flags & 4) mbar_poll(bar0 + 8u * (NS + 1), ((g / tile_chunks) - 1) & 1); else mbar_wait(bar0 + 8u * (NS + 1), ((g / tile_chunks) - 1) & 1); }   // accumulators drained
            // at most `depth` chunks in flight: chunk g - depth has retired (depth = NS: only the ring itself limits the issuer;
            // depth = 2 is wcrt_umma.cu's issuer up to round 2, which waited for chunk g-1 before going on to chunk g+1)
            if (g >= depth) mbar_wait(bar0 + 8u * ((g - depth) % NS), ((g - depth) / NS) & 1);
            const u32 sa = smem_u32(smem + s * STAGE), sb = sa + ND * A_PLANE;
#pragma unroll
            for (int ks = 0; ks < 2; ++ks) {
                const u64 bdesc = umma_desc(sb + ks * 256, 128, 512);
#pragma unroll
                for (int u = 0; u < ND; ++u) {
                    const u64 adesc = umma_desc(sa + u * A_PLANE + ks * 256, 128, 512);
                    const u32 d = tmem + (u32)(u * UN), acc = 1;
                    if (PAIR) {
                        const u32 z = 0;
                        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                                     "tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5, %5, %5, %5, %5}, p;\n\t}"
                                     ::"r"(d), "l"(adesc), "l"(bdesc), "r"(id), "r"(acc), "r"(z) : "memory");
                    } else {
                        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                                     "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
                                     ::"r"(d), "l"(adesc), "l"(bdesc), "r"(id), "r"(acc) : "memory");
                    }
                }
            }
            if (PAIR) {
                const unsigned short mask = 1;                                   // only the leader waits on it
                asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                             ::"r"(bar0 + 8u * s), "h"(mask) : "memory");
            } else {
                asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar0 + 8u * s) : "memory");
                if (tile_chunks && g % tile_chunks == tile_chunks - 1)
                    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar0 + 8u * NS) : "memory");
            }
        }
        for (int g = (chunks > NS ? chunks - NS : 0); g < chunks; ++g) mbar_wait(bar0 + 8u * (g % NS), (g / NS) & 1);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        
    }
    if (!PAIR && tile_chunks && warp >= 1) {
        // stand-in for the epilogue warps: wait for the tile's last commit, (optionally read TMEM), hand the accumulators back
        const int tiles = chunks / tile_chunks;
        if (flags & 32) {                                            // idle but alive: no barrier, no shared memory
            for (int i = 0; i < chunks / 2; ++i) __nanosleep(1000);
        } else if (flags & 64) {                                     // busy on the ALU only
            u32 x = t;
            for (int i = 0; i < chunks * 300; ++i) x = x * 1664525u + 1013904223u;
            if (x == 0x12345u) tmem_slot[1] = x;
        } else
        for (int it = 0; it < tiles; ++it) {
            if (flags & 16) {                                        // one warp watches the mbarrier, the others sit at a hardware barrier
                if (warp == 1) mbar_wait(bar0 + 8u * NS, it & 1);
                asm volatile("bar.sync 1, 384;" ::: "memory");
            } else if (flags & 8) mbar_wait_sleep(bar0 + 8u * NS, it & 1, 64 << ((flags >> 8) & 7));
            else if (flags & 2) mbar_poll(bar0 + 8u * NS, it & 1); else mbar_wait(bar0 + 8u * NS, it & 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            if (flags & 1) {                                   // read the 9 x 16 columns this warp owns, like phase 1
                u32 r[4], acc = 0;
                const u32 tl = tmem + ((u32)(32 * (warp & 3)) << 16);
                for (int sg = 0; sg < 4; ++sg) {
#pragma unroll
                    for (int w = 0; w < 9; ++w) {
                        asm volatile("tcgen05.ld.sync.aligned.16x256b.x1.b32 {%0,%1,%2,%3}, [%4];"
                                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(tl + ((u32)(16 * (sg >> 1)) << 16) + (u32)(w * UN + ((warp - 1) >> 2) * 16 + 8 * (sg & 1))));
                        acc += r[0] + r[1] + r[2] + r[3];
                    }
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                }
                if (acc == 0x12345u) tmem_slot[1] = acc;
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if ((t & 31) == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar0 + 8u * (NS + 1)) : "memory");
        }
    }
    __syncthreads();
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    cluster_sync();                                                   // the peer's shared memory and TMEM stay alive until here
    if (warp == 0) {
        if (PAIR) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
        else      asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
    }
}

template <int PAIR, int N>
void run(const char* name, int sms, double useful_cols, int depth = NS, int tile_chunks = 0, int reads = 0) {
    const int chunks = 4096;
    cudaFuncSetAttribute(skeleton<PAIR, N>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    const int grid = sms & ~1;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    skeleton<PAIR, N><<<grid, 32 * 13, SMEM>>>(64, (u32*)(size_t)reads, depth, tile_chunks);
    cudaError_t err = cudaDeviceSynchronize();
    if (err != cudaSuccess) { printf("%-44s FAILED: %s\n", name, cudaGetErrorString(err)); return; }
    cudaEventRecord(e0);
    skeleton<PAIR, N><<<grid, 32 * 13, SMEM>>>(chunks, (u32*)(size_t)reads, depth, tile_chunks);
    cudaEventRecord(e1);
    err = cudaEventSynchronize(e1);
    if (err != cudaSuccess) { printf("%-44s FAILED: %s\n", name, cudaGetErrorString(err)); return; }
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double instr = (double)chunks * 2 * ND;                       // per issuing CTA
    const double ns = ms * 1e6 / instr;
    // per-SM MACs of one instruction: 128 rows x N columns x 32 either way (the pair's 256 x N is shared by two SMs)
    const double macs_sm = 128.0 * N * 32, useful_sm = 128.0 * useful_cols * 32;
    printf("%-44s %7.3f ms  %6.1f ns/instr  %6.0f MAC/ns/SM issued  %6.0f useful  (chip: %.2f P int8 MAC/s useful)\n", name, ms, ns,
           macs_sm / ns, useful_sm / ns, useful_sm / ns * sms / 1e6);
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    printf("%s, %d SMs; ND = %d digit planes, 2 K-steps of 32 per chunk, %d-stage ring\n", p.name, p.multiProcessorCount, ND, NS);
    const int sms = p.multiProcessorCount;
    run<0, 240>("cta_group::1  128 x 240 x 32 (kernel today)", sms, 240);
    run<0, 160>("cta_group::1  128 x 160 x 32", sms, 160);
    run<0, 256>("cta_group::1  128 x 256 x 32", sms, 240);
    run<1, 256>("cta_group::2  256 x 256 x 32 (240 useful)", sms, 240);
    run<1, 160>("cta_group::2  256 x 160 x 32", sms, 160);
    run<1, 224>("cta_group::2  256 x 224 x 32", sms, 224);
    run<0, 240>("cta_group::1  128 x 240, <= 3 chunks in flight", sms, 240, 3);
    run<0, 240>("cta_group::1  128 x 240, <= 2 chunks in flight", sms, 240, 2);
    run<0, 240>("cta_group::1  128 x 240, <= 1 chunk in flight", sms, 240, 1);
    run<0, 240>("1-CTA 240, depth 2, tile handshake / 8 chunks", sms, 240, 2, 8);
    run<0, 240>("1-CTA 240, depth 2, handshake + TMEM reads", sms, 240, 2, 8, 1);
    run<0, 240>("handshake, epilogue polls (test_wait)", sms, 240, 2, 8, 2);
    run<0, 240>("handshake, issuer polls tempty", sms, 240, 2, 8, 4);
    run<0, 240>("handshake, both poll", sms, 240, 2, 8, 6);
    run<0, 240>("handshake, epilogue sleeps 64 ns between polls", sms, 240, 2, 8, 8);
    run<0, 240>("handshake, epilogue sleeps 256 ns", sms, 240, 2, 8, 8 | (2 << 8));
    run<0, 240>("handshake, epilogue sleeps 1024 ns", sms, 240, 2, 8, 8 | (4 << 8));
    run<0, 240>("handshake, 1 warp on the mbarrier + bar.sync", sms, 240, 2, 8, 16);
    run<0, 240>("same + TMEM reads", sms, 240, 2, 8, 17);
    run<0, 240>("no handshake, 12 warps asleep (nanosleep)", sms, 240, 2, 8, 32);
    run<0, 240>("no handshake, 12 warps busy on the ALU", sms, 240, 2, 8, 64);
    run<0, 240>("handshake / 16 chunks", sms, 240, 2, 16, 0);
    run<0, 240>("handshake / 4 chunks", sms, 240, 2, 4, 0);
    printf("err=%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
