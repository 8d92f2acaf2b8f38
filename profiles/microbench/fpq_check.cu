// fpq_check.cu -- exhaustive-ish check of mul_fpq (FP64-assisted approximate Shoup quotient) against the
// exact Shoup product, over random and adversarial x, twiddles and moduli.  Prints mismatches (must be 0)
// and the largest lazy result seen in units of q (must be < 4).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../matrix_fhe_gpu_b200/csrc/modmath.cuh"
using namespace mfhe;
__device__ u64 sm64(u64 z) { z += 0x9E3779B97F4A7C15ull; z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; return z ^ (z >> 31); }
__global__ void check(u64 q, u64 seed, unsigned long long* bad, unsigned long long* maxr) {
    const u64 gid = blockIdx.x * (u64)blockDim.x + threadIdx.x;
    u64 w = sm64(seed + gid * 3) % q;
    if ((gid & 1023) == 0) w = q - 1;
    if ((gid & 1023) == 1) w = 1;
    if ((gid & 1023) == 2) w = 0;
    // shoup word floor(w * 2^64 / q) by long division on the device
    u64 ws = 0, rem = w;
    for (int i = 0; i < 64; ++i) { u64 top = rem >> 63; rem <<= 1; ws <<= 1; if (top || rem >= q) { rem -= q; ws |= 1; } }
    const tw_t tw = {w, ws};
    const twf_t twf = fpq_prep(tw);
    const u64 cneg = fpq_cneg(q), nq = 0 - q;
    u64 lbad = 0, lmax = 0;
    for (int it = 0; it < 256; ++it) {
        u64 x = sm64(seed * 77 + gid * 256 + it);
        switch (it & 15) {
            case 0: x = 0; break; case 1: x = ~0ull; break; case 2: x = q - 1; break; case 3: x = 0xFFFFFFFFull; break;
            case 4: x = 0xFFFFFFFF00000000ull; break; case 5: x &= 0xFFFFFFFFull; break; case 6: x |= 0xFFFFFFFF00000000ull; break;
            case 7: x = 1; break; case 8: x = 1ull << 32; break; default: break;
        }
        const u64 r = mul_fpq(x, twf, nq, cneg);
        const u64 e = mul_shoup(x, tw, q);
        u64 rr = r; 
        if (rr >= 4 * q) ++lbad;
        rr = csub(rr, 2 * q); rr = csub(rr, q);
        if (rr != e) ++lbad;
        if (r > lmax) lmax = r;
    }
    atomicAdd(bad, lbad); atomicMax(maxr, lmax);
}
int main() {
    const u64 qs[] = {1152921504606584833ull, 1152921504538951681ull, 1125899903827969ull, 17592186435073ull, 17182765057ull,
                      36028797005856769ull, 576460752303439873ull, 1152921504606830593ull, 12289ull};
    unsigned long long *bad, *maxr; cudaMalloc(&bad, 8); cudaMalloc(&maxr, 8);
    for (u64 q : qs) {
        cudaMemset(bad, 0, 8); cudaMemset(maxr, 0, 8);
        check<<<148 * 64, 256>>>(q, q ^ 12345, bad, maxr);
        unsigned long long hb, hm; cudaMemcpy(&hb, bad, 8, cudaMemcpyDeviceToHost); cudaMemcpy(&hm, maxr, 8, cudaMemcpyDeviceToHost);
        printf("q=%llu  products=%llu  mismatches=%llu  max r/q=%.4f  (%s)\n", (unsigned long long)q, 148ull * 64 * 256 * 256, hb,
               (double)hm / (double)q, cudaGetErrorString(cudaGetLastError()));
    }
    return 0;
}
