"""Device time of the config-2 shape (N = 2^16, L = 32, batch 64) on 50-bit primes, FP64-pipe kernels (or `int`):
   [MFHE_LIB_PATH=...] python tools/time_ntt50.py [int] [bits]"""
import statistics
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
import matrix_fhe_gpu_b200 as m
import oracle as O

bits = int(sys.argv[2]) if len(sys.argv) > 2 else 50
mods = O.coeff_modulus_create(1 << 16, bits, 32)
plan = m.NttPlan(16, mods)
if len(sys.argv) > 1 and sys.argv[1] == "int":
    plan.set_arith(0)
x = O.synth_polys(1, 64, mods, 1 << 16)
d = torch.from_numpy(x.view(np.int64)).cuda()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
tf, ti = [], []
for i in range(25):
    ev[0].record(); plan.forward_(d)
    ev[1].record(); plan.inverse_(d)
    ev[2].record(); torch.cuda.synchronize()
    if i >= 5:
        tf.append(ev[0].elapsed_time(ev[1])); ti.append(ev[1].elapsed_time(ev[2]))
assert (d[0].cpu().numpy().view(np.uint64) == x[0]).all() and (d[63].cpu().numpy().view(np.uint64) == x[63]).all()
f, i = statistics.median(tf), statistics.median(ti)
print(f"arith {plan.arith} bits {bits}: fwd {f:.3f} ms inv {i:.3f} ms  {2048 / ((f + i) * 1e-3):.0f} limb-polys/s")
