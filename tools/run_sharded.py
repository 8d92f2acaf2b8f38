#!/usr/bin/env python
"""BASELINE config 5 under torchrun: limb-sharded INTT -> base conversion (all-to-all of partial sums over
NCCL) -> NTT.  Defaults: N = 2^17, 48 -> 16 limbs of 55-bit primes, batch 16 polys.

  python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/run_sharded.py [--check]

Prints one JSON line on rank 0: polys/s (device-timed, max over ranks), bytes each rank sends per step and
the achieved exchange bandwidth.  --check compares every rank's output with the CPU oracle (test-only use)."""
from __future__ import annotations

import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--logn", type=int, default=17)
    ap.add_argument("--lin", type=int, default=48)
    ap.add_argument("--lout", type=int, default=16)
    ap.add_argument("--bits", type=int, default=55)
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--check", action="store_true")
    ap.add_argument("--exchange", choices=["nccl", "pipelined", "p2p", "ipc"], default="nccl",
                    help="nccl: one all-to-all; pipelined: per batch group, overlapped; p2p: fused partial + peer stores")
    ap.add_argument("--groups", type=int, default=4)
    ap.add_argument("--p2p-groups", type=int, default=1, help="p2p: batch slices alternating between two streams")
    ap.add_argument("--ipc-groups", default="", help="ipc: comma list of batch-slice counts to time after the first pass (mfhe_comm_set_groups)")
    ap.add_argument("--also-p2p-groups", type=int, default=0,
                    help="p2p: after the main measurement, time the step again with this many stream groups (second JSON line)")
    a = ap.parse_args()
    import torch
    import torch.distributed as dist
    import oracle as O
    from matrix_fhe_gpu_b200 import sharded as S

    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    N = 1 << a.logn
    primes = [int(v) for v in O.get_primes(N, a.bits, a.lin + a.lout)]
    q, p = primes[:a.lin], primes[a.lin:]
    sh = S.LimbShardedBconv(a.logn, q, p, rank, world)
    qs, qc = sh.q_windows[rank]; ps, pc = sh.p_windows[rank]
    # every rank generates only its own limbs of the synthetic input (same values as synth_polys on the full array)
    x_full = None
    if a.check:
        x_full = O.synth_polys(42, a.batch, q, N)
        x_loc = np.ascontiguousarray(x_full[:, qs:qs + qc])
    else:
        idx = (np.arange(a.batch, dtype=np.uint64)[:, None, None] * np.uint64(a.lin * N)
               + (np.arange(qs, qs + qc, dtype=np.uint64)[None, :, None]) * np.uint64(N)
               + np.arange(N, dtype=np.uint64)[None, None, :] + np.uint64(42))
        x_loc = O.splitmix64(idx) % np.asarray(q[qs:qs + qc], dtype=np.uint64).reshape(1, -1, 1)
    x0 = torch.from_numpy(x_loc.view(np.int64)).to(dev)
    work = torch.empty_like(x0)

    if a.exchange == "p2p":
        if world < 2:
            raise SystemExit("--exchange p2p needs torchrun with >= 2 ranks")
        sh.enable_p2p(a.batch)

    ipc = None
    if a.exchange == "ipc":
        ipc = S.IpcComm(rank, world, a.lout, N, a.batch)      # mfhe_comm_*: CUDA IPC peers, C ABI only

    def step():
        work.copy_(x0)
        if ipc is not None:
            return ipc.step(sh, work, a.batch)
        if a.exchange == "p2p":
            return sh.step_p2p(work, a.batch, a.p2p_groups)
        if a.exchange == "pipelined":
            return sh.step_pipelined(work, a.batch, a.groups)
        return sh.step(work, a.batch)

    out = None
    for _ in range(max(a.warmup, 1)):
        out = step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev[0].record()
    for _ in range(a.steps):
        out = step()
    ev[1].record()
    torch.cuda.synchronize()
    t = torch.tensor([ev[0].elapsed_time(ev[1]) / a.steps], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    ok = True
    want = None
    if a.check:
        coeff = O.PhantomTables(a.logn, q).inverse(x_full)
        conv = np.stack([O.bconv(coeff[b], q, p) for b in range(a.batch)])
        want = O.PhantomTables(a.logn, p).forward(conv)[:, ps:ps + pc]
        ok = bool((out.cpu().numpy().view(np.uint64) == want).all())
        flag = torch.tensor([1 if ok else 0], device=dev)
        if world > 1:
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        ok = bool(flag.item())
    if rank == 0:
        sent = a.batch * (a.lout - pc) * N * 8
        print(json.dumps({"metric": "limb_sharded_intt_bconv_ntt_polys_per_s", "value": a.batch / (ms * 1e-3), "unit": "polys/s",
                          "n_gpus": world, "ms_per_step": ms, "config": {"N": N, "L_in": a.lin, "L_out": a.lout, "bits": a.bits,
                                                                         "batch": a.batch},
                          "exchange": {"kind": {"nccl": "all_to_all of partial sums (NCCL)",
                                                "pipelined": "all_to_all per batch group (NCCL, async) overlapped with the transforms, groups=%d" % a.groups,
                                                "ipc": "mfhe_bconv_sharded_step (C ABI): partial sums stored by the bconv kernel into the owners' CUDA-IPC receive buffers over NVLink, flag barrier",
                                                "p2p": "partial sums stored by the bconv kernel into the owners' symmetric-memory buffers over NVLink, one device barrier, stream groups=%d" % a.p2p_groups}[a.exchange], "bytes_sent_per_rank": sent,
                                       "note": "GB/s below is an upper bound on link use: whole step time, not the collective alone",
                                       "gbs_per_rank_over_step": sent / (ms * 1e-3) / 1e9},
                          "parity": ("ok" if ok else "FAILED") if a.check else "not checked"}))
        if a.check:
            print("parity ok" if ok else "parity FAILED")
    if a.exchange == "p2p" and a.also_p2p_groups > 0:
        a.p2p_groups = a.also_p2p_groups
        for _ in range(max(a.warmup, 1)):
            out = step()
        torch.cuda.synchronize(); dist.barrier()
        ev[0].record()
        for _ in range(a.steps):
            out = step()
        ev[1].record(); torch.cuda.synchronize()
        t = torch.tensor([ev[0].elapsed_time(ev[1]) / a.steps], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ok2 = True
        if a.check:                                      # the second form is parity-checked like the first
            ok2 = bool((out.cpu().numpy().view(np.uint64) == want).all())
            flag = torch.tensor([1 if ok2 else 0], device=dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            ok2 = bool(flag.item())
            ok = ok and ok2
        if rank == 0:
            print(json.dumps({"metric": "limb_sharded_intt_bconv_ntt_polys_per_s", "value": a.batch / (float(t.item()) * 1e-3),
                              "unit": "polys/s", "n_gpus": world, "ms_per_step": float(t.item()), "p2p_stream_groups": a.p2p_groups,
                              "parity": ("ok" if ok2 else "FAILED") if a.check else "not checked"}))
    if ipc is not None and a.ipc_groups:
        for g in [int(v) for v in a.ipc_groups.split(",")]:
            ipc.set_groups(g)
            for _ in range(max(a.warmup, 1)):
                out = step()
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            ev[0].record()
            for _ in range(a.steps):
                out = step()
            ev[1].record(); torch.cuda.synchronize()
            t = torch.tensor([ev[0].elapsed_time(ev[1]) / a.steps], device=dev)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            okg = True
            if a.check:
                okg = bool((out.cpu().numpy().view(np.uint64) == want).all())
                flag = torch.tensor([1 if okg else 0], device=dev)
                if world > 1:
                    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
                okg = bool(flag.item())
                ok = ok and okg
            if rank == 0:
                print(json.dumps({"metric": "limb_sharded_intt_bconv_ntt_polys_per_s", "value": a.batch / (float(t.item()) * 1e-3),
                                  "unit": "polys/s", "n_gpus": world, "ms_per_step": float(t.item()), "ipc_batch_slices": g,
                                  "parity": ("ok" if okg else "FAILED") if a.check else "not checked"}))
    if world > 1:
        dist.destroy_process_group()
    if not ok:
        sys.exit(1)


if __name__ == "__main__":
    main()
