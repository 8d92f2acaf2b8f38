// plan.h -- internal definition of mfhe_ntt_plan, shared by the C-ABI translation units.
#pragma once
#include <vector>
#include "../../include/mfhe.h"
#include "ntt_api.h"
#include "numth_host.hpp"

using mfhe::u64; using mfhe::tw_t; using mfhe::NttTables; using mfhe::SmallTables;

struct mfhe_ntt_plan {
    int logn = 0, L = 0, convention = 0;
    std::vector<uint64_t> moduli, roots;
    u64* d_moduli = nullptr;
    tw_t *d_tw = nullptr, *d_itw = nullptr, *d_ninv = nullptr, *d_w1n = nullptr;
    // FP64-pipe arithmetic (flat plans whose moduli are all below 2^50): {centred w, w / q} as doubles, same shapes
    tw_t *d_twd = nullptr, *d_itwd = nullptr, *d_ninvd = nullptr, *d_w1nd = nullptr;
    int arith = 0;                             // 0 integer, 1 FP64 pipe (mfhe_ntt_plan_set_arith)
    ulonglong2* d_mu = nullptr;                // [L] floor(2^128 / q) as {lo, hi} (Barrett)
    std::vector<mfhe::host::TwPair> h_tw, h_itw;     // kept for mfhe_ntt_plan_get_table
    size_t chunk = 0;
    uint64_t* d_stage = nullptr;               // staging for the *_host entry points
    size_t stage_bytes = 0;
    cudaStream_t host_streams[3] = {nullptr, nullptr, nullptr};   // H2D / compute / D2H of the *_host entry points (lazy)
    NttTables flat() const {
        return arith ? NttTables{d_moduli, d_twd, d_itwd, d_ninvd, d_w1nd, logn, L, 1}
                     : NttTables{d_moduli, d_tw, d_itw, d_ninv, d_w1n, logn, L, 0};
    }
    // n <= 256: the same merged network for all three conventions, on the FP64 pipe when `arith` (every modulus < 2^50)
    SmallTables small() const {
        if (arith) return SmallTables{d_moduli, d_twd, d_itwd, d_ninvd, d_w1nd, nullptr, nullptr, logn, L, convention, 1};
        return SmallTables{d_moduli, d_tw, d_itw, d_ninv, d_w1n, nullptr, nullptr, logn, L, convention, 0};
    }
    // the fused encrypt / decrypt row kernels (scheme.cu) use the negacyclic tables
    SmallTables small_rows() const { return small(); }
};

