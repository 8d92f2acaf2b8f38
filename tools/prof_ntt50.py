"""One forward + inverse of the config-2 shape on 50-bit primes (FP64-pipe kernels) for an ncu capture:
   ncu --set full --clock-control none --import-source on -k regex:ntt_ -c 8 -o gpurun_out/ntt50 python tools/prof_ntt50.py [int]"""
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
import matrix_fhe_gpu_b200 as m
import oracle as O

bits = 50
mods = O.coeff_modulus_create(1 << 16, bits, 32)
plan = m.NttPlan(16, mods)
if len(sys.argv) > 1 and sys.argv[1] == "int":
    plan.set_arith(0)
x = O.synth_polys(1, 64, mods, 1 << 16)
d = torch.from_numpy(x.view(np.int64)).cuda()
for _ in range(2):
    plan.forward_(d)
    plan.inverse_(d)
torch.cuda.synchronize()
assert (d[0].cpu().numpy().view(np.uint64) == x[0]).all()
print("ok arith", plan.arith)
