// comm.cu -- the multi-GPU exchange of the limb-sharded base conversion behind the C ABI (SURVEY 8(b) "multi-GPU" group,
// 8(e) case 3).  One process per GPU; no NCCL and no Python in the data path: every rank owns two receive buffers and a
// flag array allocated with cudaMalloc, exports them as CUDA IPC handles, and opens its peers' (the host program moves the
// 192-byte handles between the processes by whatever transport it has -- the tests use a torch.distributed store, a C++
// host would use its own).  A step is then
//     INTT of the rank's Q-limbs  ->  base-conversion kernel whose result stores go straight into slot [rank] of the
//     owner's receive buffer over NVLink (mfhe_bconv_partial_scatter)  ->  flag barrier (release / acquire at system
//     scope)  ->  sum of the world's partials mod p_j  ->  NTT of the rank's P-limbs
// with the receive buffers double-buffered: a rank passes the barrier of step k only after every peer has enqueued
// (stream order) its reduce of step k-1, so step k+1's stores never meet step k-1's reads of the same buffer.
#include <algorithm>
#include <new>
#include <vector>

#include "capi_common.h"
#include "plan.h"

using namespace mfhe;

namespace { constexpr int kMaxRanks = 64; }

struct mfhe_comm {
    int rank = 0, world = 1;
    size_t recv_words = 0;
    u64* recv[2] = {nullptr, nullptr};
    unsigned long long* flags = nullptr;                 // [world] : flags[r] = last epoch rank r signalled to us
    u64* peer_recv[2][kMaxRanks] = {};
    unsigned long long* peer_flags[kMaxRanks] = {};
    bool opened[kMaxRanks] = {};
    bool imported = false;
    unsigned long long epoch = 0;
    int phase = 0;
    int groups = 1;
    cudaStream_t side = nullptr;                         // reduce + NTT of slice s next to INTT + scatter of slice s+1
    cudaEvent_t ev_bar[16] = {}, ev_done = nullptr;
};

namespace {
struct BarrierArgs { unsigned long long* peer[kMaxRanks]; };

__global__ void comm_barrier_kernel(const __grid_constant__ BarrierArgs a, unsigned long long* local, int rank, int world,
                                    unsigned long long epoch) {
    const int t = threadIdx.x;
    if (t >= world) return;
    __threadfence_system();                              // everything this stream wrote before (peer stores included)
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(a.peer[t] + rank), "l"(epoch) : "memory");
    unsigned long long v;
    do {
        asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(local + t) : "memory");
    } while (v < epoch);
}

void window(int total, int world, int r, int* start, int* count) {     // sharded.py limb_windows
    const int base = total / world, extra = total % world;
    *count = base + (r < extra ? 1 : 0);
    *start = r * base + std::min(r, extra);
}
}  // namespace

extern "C" void mfhe_comm_destroy(mfhe_comm* c) {
    if (!c) return;
    for (int r = 0; r < c->world; ++r) {
        if (!c->opened[r]) continue;
        cudaIpcCloseMemHandle(c->peer_recv[0][r]); cudaIpcCloseMemHandle(c->peer_recv[1][r]); cudaIpcCloseMemHandle(c->peer_flags[r]);
    }
    cudaFree(c->recv[0]); cudaFree(c->recv[1]); cudaFree(c->flags);
    if (c->side) cudaStreamDestroy(c->side);
    for (cudaEvent_t e : c->ev_bar) if (e) cudaEventDestroy(e);
    if (c->ev_done) cudaEventDestroy(c->ev_done);
    delete c;
}

extern "C" int mfhe_comm_create(int rank, int world, size_t recv_words, mfhe_comm** out) {
    if (!out) return fail(MFHE_EINVAL, "comm_create: out is NULL");
    *out = nullptr;
    if (world < 1 || world > kMaxRanks || rank < 0 || rank >= world || !recv_words) return fail(MFHE_EINVAL, "comm_create: bad argument");
    mfhe_comm* c = new (std::nothrow) mfhe_comm;
    if (!c) return fail(MFHE_ENOMEM, "comm_create: host allocation failed");
    c->rank = rank; c->world = world; c->recv_words = recv_words;
    cudaError_t e = cudaMalloc((void**)&c->recv[0], recv_words * 8);
    if (e == cudaSuccess) e = cudaMalloc((void**)&c->recv[1], recv_words * 8);
    if (e == cudaSuccess) e = cudaMalloc((void**)&c->flags, kMaxRanks * sizeof(unsigned long long));
    if (e == cudaSuccess) e = cudaMemset(c->flags, 0, kMaxRanks * sizeof(unsigned long long));
    if (e != cudaSuccess) { mfhe_comm_destroy(c); return cuda_fail(e, "comm_create"); }
    c->peer_recv[0][rank] = c->recv[0]; c->peer_recv[1][rank] = c->recv[1]; c->peer_flags[rank] = c->flags;
    c->imported = (world == 1);
    *out = c;
    return MFHE_OK;
}

// h_handle: MFHE_COMM_HANDLE_BYTES = 3 cudaIpcMemHandle_t (receive buffer 0, receive buffer 1, flags)
extern "C" int mfhe_comm_export(const mfhe_comm* c, void* h_handle) {
    if (!c || !h_handle) return fail(MFHE_EINVAL, "comm_export: NULL argument");
    cudaIpcMemHandle_t* h = (cudaIpcMemHandle_t*)h_handle;
    cudaError_t e = cudaIpcGetMemHandle(&h[0], c->recv[0]);
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h[1], c->recv[1]);
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h[2], c->flags);
    return e == cudaSuccess ? MFHE_OK : cuda_fail(e, "comm_export");
}

// h_all: world x MFHE_COMM_HANDLE_BYTES in rank order (this rank's own entry is ignored)
extern "C" int mfhe_comm_import(mfhe_comm* c, const void* h_all) {
    if (!c || !h_all) return fail(MFHE_EINVAL, "comm_import: NULL argument");
    if (c->imported) return MFHE_OK;
    const cudaIpcMemHandle_t* h = (const cudaIpcMemHandle_t*)h_all;
    for (int r = 0; r < c->world; ++r) {
        if (r == c->rank) continue;
        cudaError_t e = cudaIpcOpenMemHandle((void**)&c->peer_recv[0][r], h[3 * r + 0], cudaIpcMemLazyEnablePeerAccess);
        if (e == cudaSuccess) e = cudaIpcOpenMemHandle((void**)&c->peer_recv[1][r], h[3 * r + 1], cudaIpcMemLazyEnablePeerAccess);
        if (e == cudaSuccess) e = cudaIpcOpenMemHandle((void**)&c->peer_flags[r], h[3 * r + 2], cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) return cuda_fail(e, "comm_import (cudaIpcOpenMemHandle: peers must be NVLink / P2P reachable)");
        c->opened[r] = true;
    }
    c->imported = true;
    return MFHE_OK;
}

extern "C" int mfhe_comm_barrier(mfhe_comm* c, void* stream) {
    if (!c || !c->imported) return fail(MFHE_EINVAL, "comm_barrier: communicator not imported");
    BarrierArgs a{};
    for (int r = 0; r < c->world; ++r) a.peer[r] = c->peer_flags[r];
    comm_barrier_kernel<<<1, kMaxRanks, 0, (cudaStream_t)stream>>>(a, c->flags, c->rank, c->world, ++c->epoch);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? MFHE_OK : cuda_fail(e, "comm_barrier");
}

extern "C" int mfhe_comm_set_groups(mfhe_comm* c, int groups) {
    if (!c || groups < 1 || groups > 16) return fail(MFHE_EINVAL, "comm_set_groups: 1..16 slices");
    if (groups > 1 && !c->side) {
        cudaError_t e = cudaStreamCreateWithFlags(&c->side, cudaStreamNonBlocking);
        for (int i = 0; i < 16 && e == cudaSuccess; ++i) e = cudaEventCreateWithFlags(&c->ev_bar[i], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_done, cudaEventDisableTiming);
        if (e != cudaSuccess) return cuda_fail(e, "comm_set_groups");
    }
    c->groups = groups;
    return MFHE_OK;
}

extern "C" size_t mfhe_bconv_sharded_recv_words(int world, int l_out, size_t N, size_t batch) {
    int s = 0, cmax = 0;
    for (int r = 0; r < world; ++r) { int c; window(l_out, world, r, &s, &c); cmax = std::max(cmax, c); }
    return (size_t)world * batch * (size_t)cmax * N;
}

// BASELINE config 5 on one rank.  plan_q / plan_p: flat plans over the FULL input / output bases; d_x_local
// [batch][q_count][N] in NTT form over this rank's window of the input base (destroyed); d_out [batch][p_count][N] in
// NTT form over this rank's window of the output base.  Windows are the balanced contiguous split of sharded.py.
extern "C" int mfhe_bconv_sharded_step(mfhe_comm* c, const mfhe_ntt_plan* plan_q, const mfhe_ntt_plan* plan_p,
                                       const mfhe_bconv_plan* bc, uint64_t* d_x_local, uint64_t* d_out, size_t batch, void* stream) {
    if (!c || !plan_q || !plan_p || !bc || !d_x_local || !d_out || !batch) return fail(MFHE_EINVAL, "bconv_sharded_step: bad argument");
    if (!c->imported) return fail(MFHE_EINVAL, "bconv_sharded_step: communicator not imported");
    int l_in = 0, l_out = 0;
    mfhe_bconv_plan_sizes(bc, &l_in, &l_out);
    if (l_in != plan_q->L || l_out != plan_p->L || plan_q->logn != plan_p->logn || l_out > 128)
        return fail(MFHE_EINVAL, "bconv_sharded_step: plans and conversion do not match");
    const size_t N = (size_t)1 << plan_q->logn;
    if (c->world > l_in || c->world > l_out) return fail(MFHE_EINVAL, "bconv_sharded_step: more ranks than limbs in a base");
    if (mfhe_bconv_sharded_recv_words(c->world, l_out, N, batch) > c->recv_words) return fail(MFHE_EINVAL, "bconv_sharded_step: receive buffers too small");
    int qs, qc, ps, pc;
    window(l_in, c->world, c->rank, &qs, &qc);
    window(l_out, c->world, c->rank, &ps, &pc);
    const int ph = c->phase; c->phase ^= 1;
    std::vector<uint64_t*> dst(l_out), dsl(l_out); std::vector<size_t> stride(l_out);
    for (int r = 0; r < c->world; ++r) {
        int s, cnt; window(l_out, c->world, r, &s, &cnt);
        for (int j = s; j < s + cnt; ++j) {
            dst[j] = (uint64_t*)c->peer_recv[ph][r] + ((size_t)c->rank * batch * cnt + (size_t)(j - s)) * N;
            stride[j] = (size_t)cnt * N;
        }
    }
    const int G = (int)std::min<size_t>((size_t)c->groups, batch);
    int rc = MFHE_OK;
    if (G <= 1) {
        rc = mfhe_ntt_inverse_limbs(plan_q, d_x_local, batch, qs, qc, stream);
        if (rc == MFHE_OK) rc = mfhe_bconv_partial_scatter(bc, d_x_local, qs, qc, dst.data(), stride.data(), N, batch, stream);
        if (rc == MFHE_OK) rc = mfhe_comm_barrier(c, stream);
        if (rc == MFHE_OK) rc = mfhe_bconv_reduce(bc, (const uint64_t*)c->recv[ph], c->world, ps, pc, d_out, N, batch, stream);
        if (rc == MFHE_OK) rc = mfhe_ntt_forward_limbs(plan_p, d_out, batch, ps, pc, stream);
        return rc;
    }
    // software pipeline over batch slices: the caller's stream does INTT -> scatter -> barrier per slice, the side stream
    // reduce -> NTT of a slice as soon as its barrier has passed; the caller's stream rejoins at the end
    cudaStream_t st = (cudaStream_t)stream;
    for (int g = 0; g < G && rc == MFHE_OK; ++g) {
        const size_t b0 = batch * g / G, nb = batch * (g + 1) / G - b0;
        if (!nb) continue;
        uint64_t* xs = d_x_local + b0 * (size_t)qc * N;
        for (int j = 0; j < l_out; ++j) dsl[j] = dst[j] + b0 * stride[j];
        rc = mfhe_ntt_inverse_limbs(plan_q, xs, nb, qs, qc, stream);
        if (rc == MFHE_OK) rc = mfhe_bconv_partial_scatter(bc, xs, qs, qc, dsl.data(), stride.data(), N, nb, stream);
        if (rc == MFHE_OK) rc = mfhe_comm_barrier(c, stream);
        if (rc != MFHE_OK) break;
        cudaEventRecord(c->ev_bar[g], st);
        cudaStreamWaitEvent(c->side, c->ev_bar[g], 0);
        uint64_t* os = d_out + b0 * (size_t)pc * N;
        rc = mfhe_bconv_reduce_strided(bc, (const uint64_t*)c->recv[ph] + b0 * (size_t)pc * N, c->world, batch * (size_t)pc * N, ps, pc,
                                       os, N, nb, c->side);
        if (rc == MFHE_OK) rc = mfhe_ntt_forward_limbs(plan_p, os, nb, ps, pc, c->side);
    }
    cudaEventRecord(c->ev_done, c->side);
    cudaStreamWaitEvent(st, c->ev_done, 0);
    return rc;
}
