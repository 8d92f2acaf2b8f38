"""Multi-GPU side of the hot path: one process per GPU, torch.distributed for the plumbing.

Two cases exist on this path:

* batch sharding (flat NTT, the matrix pipeline): every limb-poly / (re, im) pair is independent, so
  rank r simply processes items r, r+k, ... -- `batch_slice` -- and there is NO data-path collective;
* limb-sharded base conversion (BASELINE config 5): rank r owns a window of the input base Q and a
  window of the output base P.  Base conversion needs every input limb of a coefficient, so this is
  the one real exchange step.  Each rank forms the partial sums of ITS input limbs for ALL output
  limbs (mfhe_bconv_partial: already reduced mod p_j, so partials add without overflow), the partials
  are exchanged so that every rank receives the world's partials for ITS output limbs only
  (`exchange_partials`: an all-to-all; (k-1)/k * batch * L_out * N * 8 bytes leave each rank instead
  of the (k-1) * batch * L_in * N * 8 an all-gather of the inputs would move), and mfhe_bconv_reduce
  adds them mod p_j.  The local INTT before and NTT after are Phantom's
  nwt_2d_radix8_backward/forward_inplace on a limb window (ph/src/ntt/intt_2d.cu:724-757,
  fntt_2d.cu:620-653); the conversion is DBaseConverter::bConv_BEHZ (ph/src/rns_bconv.cu:212-229).

Three forms of the exchange exist (`tools/run_sharded.py --exchange`):
  nccl       partial -> pack -> one all-to-all (NCCL) -> reduce                      (`step`)
  pipelined  the same per batch group, the all-to-all of group g overlapping the INTT + partial sums of
             group g+1 and the reduce + NTT of group g-1                             (`step_pipelined`)
  p2p        FUSED compute + exchange: mfhe_bconv_partial_scatter stores every partial sum straight into
             the receive buffer of the rank that owns the output limb (CUDA symmetric memory over NVLink /
             NVSwitch peer mappings, `scatter_table` gives the destination of each output limb), so there is
             no pack pass, no send buffer and no collective -- one device-side barrier orders the step
             (`step_p2p`; double-buffered so that step k+1's stores never meet step k's reads).

`limb_windows`, `batch_slice`, `scatter_table` and `exchange_partials` are pure torch / torch.distributed (they run on
CPU tensors over gloo in the test-suite and on CUDA tensors over NCCL on the box); the compute calls
go through the C ABI and need the GPU.
"""
from __future__ import annotations

import os
from typing import List, Sequence, Tuple


def limb_windows(total: int, world: int) -> List[Tuple[int, int]]:
    """Contiguous balanced windows (start, count) of `total` limbs over `world` ranks; the first
    total % world ranks own one limb more.  Config 5: 48 Q-limbs over 8 ranks -> 6 each, 16 P-limbs -> 2."""
    if world < 1 or total < 0:
        raise ValueError("bad partition")
    base, extra = divmod(total, world)
    out, start = [], 0
    for r in range(world):
        cnt = base + (1 if r < extra else 0)
        out.append((start, cnt))
        start += cnt
    return out


def batch_slice(batch: int, rank: int, world: int) -> range:
    """Items of a batch owned by `rank`: r, r + world, ... (pairs r, r+k, ...)."""
    return range(rank, batch, world)


def pack_for_exchange(partial, out_windows: Sequence[Tuple[int, int]]):
    """partial [batch][L_out][N] -> flat send buffer ordered by destination rank, each chunk
    [batch][count_r][N]; returns (send, split_sizes)."""
    import torch
    chunks = [partial[:, s:s + c, :].reshape(-1) for s, c in out_windows]
    return torch.cat(chunks), [int(ch.numel()) for ch in chunks]


def exchange_partials(partial, out_windows: Sequence[Tuple[int, int]], rank: int, world: int, group=None):
    """All-to-all of the base-conversion partial sums.

    partial: [batch][L_out][N] int64 (u64 bit patterns) on this rank.
    Returns parts [world][batch][count_rank][N]: every rank's partial for THIS rank's output window,
    in rank order -- the layout mfhe_bconv_reduce consumes."""
    import torch
    import torch.distributed as dist
    batch, _, n = partial.shape
    send, in_splits = pack_for_exchange(partial, out_windows)
    mine = out_windows[rank][1]
    recv = torch.empty(world * batch * mine * n, dtype=partial.dtype, device=partial.device)
    if world == 1:
        recv.copy_(send)
    else:
        dist.all_to_all_single(recv, send, output_split_sizes=[batch * mine * n] * world,
                               input_split_sizes=in_splits, group=group)
    return recv.view(world, batch, mine, n)


def scatter_table(peer_bases: Sequence[int], out_windows: Sequence[Tuple[int, int]], src_rank: int, batch: int, n: int,
                  count_max: int | None = None):
    """Destinations for mfhe_bconv_partial_scatter on rank `src_rank`.

    Every rank r owns a receive buffer (base address peer_bases[r], as seen from src_rank) laid out
    [world][batch][count_r][N] u64 -- slot `src_rank` of it is ours -- where count_r is rank r's output window.
    Returns (ptrs, batch_strides_in_words), one entry per output limb j: item b of limb j goes to
    ptrs[j] + 8 * (b * strides[j] + coefficient)."""
    ptrs, strides = [], []
    for r, (s, c) in enumerate(out_windows):
        for j in range(s, s + c):
            ptrs.append(int(peer_bases[r]) + 8 * ((src_rank * batch * c + (j - s)) * n))
            strides.append(c * n)
    return ptrs, strides


def emulate_exchange(partials, out_windows: Sequence[Tuple[int, int]], rank: int):
    """Single-process stand-in for exchange_partials: partials[r] is rank r's [batch][L_out][N]."""
    import torch
    s, c = out_windows[rank]
    return torch.stack([p[:, s:s + c, :] for p in partials]).contiguous()


class IpcComm:
    """mfhe_comm (csrc/comm.cu): the C-ABI form of the fused exchange -- receive buffers and barrier flags shared between
    the processes as CUDA IPC handles, no NCCL and no torch symmetric memory in the data path.  torch.distributed is used
    ONCE, to move the 192-byte handles (a C++ host would use its own transport)."""

    def __init__(self, rank: int, world: int, l_out: int, n: int, batch: int, group=None):
        import ctypes as C
        from . import lib, check
        L = lib()
        L.mfhe_bconv_sharded_recv_words.restype = C.c_size_t
        L.mfhe_bconv_sharded_recv_words.argtypes = [C.c_int, C.c_int, C.c_size_t, C.c_size_t]
        L.mfhe_comm_create.argtypes = [C.c_int, C.c_int, C.c_size_t, C.POINTER(C.c_void_p)]
        L.mfhe_comm_destroy.argtypes = [C.c_void_p]; L.mfhe_comm_destroy.restype = None
        L.mfhe_comm_export.argtypes = [C.c_void_p, C.c_void_p]
        L.mfhe_comm_import.argtypes = [C.c_void_p, C.c_void_p]
        L.mfhe_comm_barrier.argtypes = [C.c_void_p, C.c_void_p]
        L.mfhe_bconv_sharded_step.argtypes = [C.c_void_p] * 6 + [C.c_size_t, C.c_void_p]
        L.mfhe_comm_set_groups.argtypes = [C.c_void_p, C.c_int]
        self._L, self.rank, self.world = L, rank, world
        words = L.mfhe_bconv_sharded_recv_words(world, l_out, n, batch)
        h = C.c_void_p()
        check(L.mfhe_comm_create(rank, world, words, C.byref(h)))
        self._h = h
        if world > 1:
            import torch.distributed as dist
            mine = C.create_string_buffer(192)
            check(L.mfhe_comm_export(h, mine))
            gathered = [None] * world
            dist.all_gather_object(gathered, bytes(mine.raw), group=group)
            allh = C.create_string_buffer(b"".join(gathered), 192 * world)
            check(L.mfhe_comm_import(h, allh))
            dist.barrier(group=group)                     # every rank has opened every peer before the first store

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                self._L.mfhe_comm_destroy(h)
            except Exception:
                pass

    def barrier(self):
        from . import check, _stream_handle
        check(self._L.mfhe_comm_barrier(self._h, _stream_handle()))

    def set_groups(self, groups: int):
        """Batch slices per step: > 1 overlaps a slice's reduce + NTT (side stream) with the next slice's INTT + scatter."""
        from . import check
        check(self._L.mfhe_comm_set_groups(self._h, groups))

    def step(self, sh: "LimbShardedBconv", x_local, batch: int):
        """mfhe_bconv_sharded_step: the whole config-5 step of this rank in one C call."""
        import torch
        from . import check, _dev_ptr, _stream_handle
        pc = sh.p_windows[self.rank][1]
        out = torch.empty(batch * pc * sh.n, dtype=torch.int64, device=x_local.device)
        check(self._L.mfhe_bconv_sharded_step(self._h, sh.plan_q._h, sh.plan_p._h, sh.bconv._h, _dev_ptr(x_local), _dev_ptr(out),
                                              batch, _stream_handle()))
        return out.view(batch, pc, sh.n)


class LimbShardedBconv:
    """Config 5 step on one rank: INTT(own Q-limbs) -> partial base conversion -> exchange -> reduce
    -> NTT(own P-limbs).  x_local is [batch][q_count][N] in NTT form; the result is [batch][p_count][N]."""

    def __init__(self, logn: int, q_in, p_out, rank: int, world: int, group=None):
        from . import NttPlan
        from .scheme import BconvPlan
        self.logn, self.n = logn, 1 << logn
        self.rank, self.world, self.group = rank, world, group
        self.q = [int(v) for v in q_in]
        self.p = [int(v) for v in p_out]
        if world > len(self.q) or world > len(self.p):
            raise ValueError(f"limb sharding needs at least one limb of each base per rank: world = {world}, "
                             f"|Q| = {len(self.q)}, |P| = {len(self.p)}")
        self.q_windows = limb_windows(len(self.q), world)
        self.p_windows = limb_windows(len(self.p), world)
        self.plan_q = NttPlan(logn, self.q)
        self.plan_p = NttPlan(logn, self.p)
        self.bconv = BconvPlan(self.q, self.p)

    def local_partial(self, x_local, batch: int, rank: int | None = None):
        qs, qc = self.q_windows[self.rank if rank is None else rank]
        self.plan_q.inverse_limbs_(x_local, qs, qc)
        return self.bconv.partial(x_local, qs, qc, self.n, batch).view(batch, len(self.p), self.n)

    def finish(self, parts, batch: int, rank: int | None = None):
        ps, pc = self.p_windows[self.rank if rank is None else rank]
        out = self.bconv.reduce(parts.reshape(-1), parts.shape[0], ps, pc, self.n, batch)
        self.plan_p.forward_limbs_(out, ps, pc)
        return out.view(batch, pc, self.n)

    def step_unsharded(self, x, batch: int):
        """world == 1: the whole config-5 step on one GPU with no partial / reduce split: INTT over all Q limbs ->
        mfhe_bconv -> NTT over all P limbs (what Phantom does with nwt_2d_radix8_backward_inplace + bConv_BEHZ +
        nwt_2d_radix8_forward_inplace, one call per poly)."""
        if self.world != 1:
            raise RuntimeError("step_unsharded: single-rank form")
        import torch
        self.plan_q.inverse_(x)
        lin = len(self.q)
        # 16 or more input limbs and N a multiple of 128: one mfhe_bconv call runs on the tensor cores (bconv_umma.cuh,
        # 0.67 ms at 48 -> 16 limbs against 2 x 0.66 ms + reduce for the window form below)
        tensor = lin >= 16 and self.n % 128 == 0 and os.environ.get("MFHE_BCONV_PATH", "") != "c"
        win = 0 if tensor else next((w for w in (24, 16, 12, 8) if lin > w and lin % w == 0), 0)
        if win:
            # the register-resident, one-word-Barrett forms of the conversion kernel exist for windows of 2..24 input
            # limbs (bconv.cu); a wider base is converted window by window and the partial sums (already reduced mod
            # p_j) are added by mfhe_bconv_reduce -- same words as one mfhe_bconv call, 0.6 of its time at 48 limbs
            # (the kernel wants its input window packed, so the windows are taken poly by poly: [win][N] of one poly is
            # contiguous)
            from . import check, _dev_ptr, _stream_handle
            from .scheme import _L
            nw, lo = lin // win, len(self.p)
            parts = torch.empty(nw, batch, lo, self.n, dtype=torch.int64, device=x.device)
            xv = x.view(batch, lin, self.n)
            for b in range(batch):
                for w in range(nw):
                    check(_L().mfhe_bconv_partial(self.bconv._h, _dev_ptr(xv[b, w * win]), w * win, win, _dev_ptr(parts[w, b]),
                                                  self.n, 1, _stream_handle()))
            out = self.bconv.reduce(parts.view(-1), nw, 0, lo, self.n, batch)
        else:
            out = self.bconv.convert(x, self.n, batch)
        self.plan_p.forward_(out)
        return out.view(batch, len(self.p), self.n)

    def step(self, x_local, batch: int):
        partial = self.local_partial(x_local, batch)
        parts = exchange_partials(partial, self.p_windows, self.rank, self.world, self.group)
        return self.finish(parts, batch)

    # ---- pipelined NCCL form ------------------------------------------------------------------------
    def step_pipelined(self, x_local, batch: int, groups: int = 4):
        """Same result as `step`, the batch cut into `groups` slices: while NCCL moves slice g (on its own
        stream, async_op) the compute stream already runs INTT + partial sums of slice g+1 and, one slice
        later, reduce + NTT of slice g."""
        import torch
        import torch.distributed as dist
        groups = max(1, min(groups, batch))
        bounds = [batch * g // groups for g in range(groups + 1)]
        mine = self.p_windows[self.rank][1]
        x = x_local.view(batch, -1, self.n)
        out = torch.empty(batch, mine, self.n, dtype=x_local.dtype, device=x_local.device)
        inflight = []            # (work, recv, send, b0, nb)

        def drain():
            work, recv, send, b0, nb = inflight.pop(0)
            if work is not None:
                work.wait()                                     # compute stream waits for NCCL's stream
            out[b0:b0 + nb] = self.finish(recv.view(self.world, nb, mine, self.n), nb)

        for g in range(groups):
            b0, nb = bounds[g], bounds[g + 1] - bounds[g]
            if nb == 0:
                continue
            partial = self.local_partial(x[b0:b0 + nb], nb)
            send, in_splits = pack_for_exchange(partial, self.p_windows)
            recv = torch.empty(self.world * nb * mine * self.n, dtype=send.dtype, device=send.device)
            if self.world == 1:
                recv.copy_(send); work = None
            else:
                work = dist.all_to_all_single(recv, send, output_split_sizes=[nb * mine * self.n] * self.world,
                                              input_split_sizes=in_splits, group=self.group, async_op=True)
            inflight.append((work, recv, send, b0, nb))
            if len(inflight) > 1:
                drain()
        while inflight:
            drain()
        return out

    # ---- fused compute + exchange over peer memory -----------------------------------------------------
    def enable_p2p(self, batch: int):
        """Allocates the double-buffered symmetric receive buffers ([world][batch][count][N] each) and exchanges
        their peer addresses.  Needs an initialised NCCL process group with all ranks on one NVLink domain."""
        import torch
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        dev = torch.device("cuda", torch.cuda.current_device())
        cmax = max(c for _, c in self.p_windows)
        words = self.world * batch * cmax * self.n
        grp = self.group if self.group is not None else dist.group.WORLD
        self._p2p = []
        for _ in range(2):
            buf = symm.empty(words, dtype=torch.int64, device=dev)
            hdl = symm.rendezvous(buf, grp)
            ptrs, strides = scatter_table([int(v) for v in hdl.buffer_ptrs], self.p_windows, self.rank, batch, self.n)
            self._p2p.append((buf, hdl, ptrs, strides))
        self._p2p_batch, self._p2p_phase = batch, 0

    def step_p2p(self, x_local, batch: int, groups: int = 1):
        """INTT -> partial sums stored directly into the owners' receive buffers -> device barrier -> reduce -> NTT.

        groups > 1 cuts the batch into slices that alternate between two side streams, so that the NVLink-bound
        scatter kernel of one slice runs next to the integer-bound INTT of the following one."""
        import torch
        if getattr(self, "_p2p", None) is None or self._p2p_batch != batch:
            raise RuntimeError("step_p2p: call enable_p2p(batch) first")
        buf, hdl, ptrs, strides = self._p2p[self._p2p_phase]
        self._p2p_phase ^= 1
        qs, qc = self.q_windows[self.rank]
        mine = self.p_windows[self.rank][1]
        groups = max(1, min(groups, batch))
        if groups == 1:
            self.plan_q.inverse_limbs_(x_local, qs, qc)
            self.bconv.partial_scatter(x_local, qs, qc, ptrs, strides, self.n, batch)
        else:
            if getattr(self, "_side", None) is None:
                self._side = [torch.cuda.Stream(), torch.cuda.Stream()]
            cur = torch.cuda.current_stream()
            x = x_local.view(batch, qc, self.n)
            for g in range(groups):
                b0, b1 = batch * g // groups, batch * (g + 1) // groups
                if b1 == b0:
                    continue
                st = self._side[g & 1]
                st.wait_stream(cur)
                with torch.cuda.stream(st):
                    xs = x[b0:b1]
                    self.plan_q.inverse_limbs_(xs, qs, qc)
                    self.bconv.partial_scatter(xs, qs, qc, [p + 8 * b0 * sj for p, sj in zip(ptrs, strides)], strides,
                                               self.n, b1 - b0)
            for st in self._side:
                cur.wait_stream(st)
        hdl.barrier(channel=0)          # stream-ordered: every rank's stores are complete and visible after it
        parts = buf[: self.world * batch * mine * self.n].view(self.world, batch, mine, self.n)
        return self.finish(parts, batch)

    def step_p2p_stages(self, x_local, batch: int):
        """step_p2p (one stream group) with a CUDA event between the stages; returns (out, [ms_intt, ms_scatter,
        ms_barrier, ms_reduce_ntt]).  bench.py divides the bytes a rank sends by ms_scatter -- the scatter kernel's own
        duration -- for the achieved NVLink egress."""
        import torch
        if getattr(self, "_p2p", None) is None or self._p2p_batch != batch:
            raise RuntimeError("step_p2p_stages: call enable_p2p(batch) first")
        buf, hdl, ptrs, strides = self._p2p[self._p2p_phase]
        self._p2p_phase ^= 1
        qs, qc = self.q_windows[self.rank]
        mine = self.p_windows[self.rank][1]
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
        ev[0].record()
        self.plan_q.inverse_limbs_(x_local, qs, qc)
        ev[1].record()
        self.bconv.partial_scatter(x_local, qs, qc, ptrs, strides, self.n, batch)
        ev[2].record()
        hdl.barrier(channel=0)
        ev[3].record()
        parts = buf[: self.world * batch * mine * self.n].view(self.world, batch, mine, self.n)
        out = self.finish(parts, batch)
        ev[4].record()
        torch.cuda.synchronize()
        return out, [ev[i].elapsed_time(ev[i + 1]) for i in range(4)]
