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
    ulonglong2* d_mu = nullptr;                // [L] floor(2^128 / q) as {lo, hi} (Barrett)
    std::vector<mfhe::host::TwPair> h_tw, h_itw;     // kept for mfhe_ntt_plan_get_table
    size_t chunk = 0;
    uint64_t* d_stage = nullptr;               // staging for the *_host entry points
    size_t stage_bytes = 0;
    cudaStream_t host_streams[3] = {nullptr, nullptr, nullptr};   // H2D / compute / D2H of the *_host entry points (lazy)
    NttTables flat() const { return NttTables{d_moduli, d_tw, d_itw, d_ninv, d_w1n, logn, L}; }
    SmallTables small() const {
        return SmallTables{d_moduli, d_tw, d_itw, d_ninv, d_w1n, nullptr, nullptr, logn, L, convention};
    }
};

