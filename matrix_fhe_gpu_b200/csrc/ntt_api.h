// ntt_api.h -- internal C++ interface between the C-ABI layer and the kernels.
#pragma once
#include <cstddef>
#include <cuda_runtime.h>
#include "ntt_flat.cuh"

namespace mfhe {

// flat negacyclic NTT, N = 2^12..2^17 (ntt_flat.cu)
// inverse only: `mul` (same shape as data, NTT form) is multiplied in while the first pass loads, mu = floor(2^128/q)
cudaError_t ntt_flat_launch(bool forward, u64* data, const NttTables& T, int limb_start,
                            int limb_count, size_t batch, size_t chunk_limb_polys,
                            cudaStream_t st, const u64* mul = nullptr, const ulonglong2* mu = nullptr);

// twiddle tables of the FP64-pipe arithmetic from the integer ones ({w, w'} -> {centred w, w / q} as doubles)
cudaError_t ntt_fp_tables_launch(const tw_t* src, tw_t* dst, const u64* moduli, size_t per, int L, cudaStream_t st);

// small transforms, n = 2^3..2^8, one warp-slice per limb-poly (ntt_small.cu)
enum SmallConvention { kPhantomNegacyclic = 0, kGlXnMinusI = 1, kCyclicNatural = 2 };
struct SmallTables {
    const u64*  moduli;   // [L]
    const tw_t* tw;       // [L][n]  forward twiddles (convention specific)
    const tw_t* itw;      // [L][n]
    const tw_t* ninv;     // [L]
    const tw_t* w1n;      // [L]
    const tw_t* twist;    // [L][n]  gl: psi^j         (nullptr otherwise)
    const tw_t* itwist;   // [L][n]  gl: psi^-j * n^-1 (nullptr otherwise)
    int logn;
    int L;
    int convention;
    int arith;            // 1: tw / itw / ninv / w1n are the FP64-pipe tables {w, w/q} (fused row kernel of scheme.cu only)
};
cudaError_t ntt_small_launch(bool forward, u64* data, const SmallTables& T, int limbs,
                             size_t batch, cudaStream_t st);

}  // namespace mfhe
