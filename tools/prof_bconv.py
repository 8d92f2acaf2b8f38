#!/usr/bin/env python
"""One rank's share of the config-5 base conversion (N=2^17, 48 -> 16 limbs, batch 16, in_count = 48 / ranks) with
local destinations: the launch ncu looks at (tools: ncu --set full -k regex:bconv_kernel).  Prints CUDA-event times."""
import argparse, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

def main():
    ap = argparse.ArgumentParser(); ap.add_argument("--ranks", type=int, default=8); ap.add_argument("--reps", type=int, default=20)
    a = ap.parse_args()
    import torch
    import oracle as O
    from matrix_fhe_gpu_b200 import sharded as S
    logn, lin, lout, batch = 17, 48, 16, 16
    N = 1 << logn
    primes = [int(v) for v in O.get_primes(N, 55, lin + lout)]
    sh = S.LimbShardedBconv(logn, primes[:lin], primes[lin:], 0, a.ranks)
    qs, qc = sh.q_windows[0]
    dev = torch.device("cuda", 0)
    x = torch.randint(0, 1 << 54, (batch, qc, N), dtype=torch.int64, device=dev)
    cmax = max(c for _, c in sh.p_windows)
    bufs = [torch.empty(a.ranks * batch * cmax * N, dtype=torch.int64, device=dev) for _ in range(a.ranks)]
    ptrs, strides = S.scatter_table([b.data_ptr() for b in bufs], sh.p_windows, 0, batch, N)
    def t(fn):
        fn(); torch.cuda.synchronize()
        e = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        e[0].record()
        for _ in range(a.reps): fn()
        e[1].record(); torch.cuda.synchronize()
        return e[0].elapsed_time(e[1]) / a.reps
    ms_s = t(lambda: sh.bconv.partial_scatter(x, qs, qc, ptrs, strides, N, batch))
    ms_p = t(lambda: sh.bconv.partial(x, qs, qc, N, batch))
    ms_i = t(lambda: sh.plan_q.inverse_limbs_(x, qs, qc))
    parts = bufs[0][: a.ranks * batch * sh.p_windows[0][1] * N].view(a.ranks, batch, -1, N)
    ms_f = t(lambda: sh.finish(parts, batch))
    byts = 8 * batch * N * (qc + lout)
    macs = batch * N * qc * lout
    print(f"ranks={a.ranks} in_count={qc}: partial_scatter {ms_s:.4f} ms ({byts / ms_s / 1e6:.0f} GB/s algorithmic, "
          f"{macs / ms_s / 1e6:.1f} G modular MAC/s), partial {ms_p:.4f} ms, INTT({batch * qc} limb-polys) {ms_i:.4f} ms, "
          f"reduce+NTT {ms_f:.4f} ms")

if __name__ == "__main__":
    main()
