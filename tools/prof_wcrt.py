"""Runs the W-CRT forward / inverse a few times at the native shape (for ncu).  python tools/prof_wcrt.py [L] [reps]"""
import sys
import torch
sys.path.insert(0, ".")
import matrix_fhe_gpu_b200.scheme as S
L = int(sys.argv[1]) if len(sys.argv) > 1 else 11
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
extra = [17194410241, 17194804993, 17195989249, 17196581377, 17196976129, 17197963009, 17199936769, 17201318401,
         17201515777, 17203884289, 17206844929, 17207437057, 17209608193]
g = S.Context(moduli=(S.REF_MODULI + extra)[:L])
a = g.sample_uniform()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
for _ in range(reps):
    ev[0].record(); b = g.wntt_forward_matrix(a); ev[1].record(); torch.cuda.synchronize()
    print("wcrt_forward ms", ev[0].elapsed_time(ev[1]))
    ev[0].record(); c = g.wntt_inverse_matrix(b); ev[1].record(); torch.cuda.synchronize()
    print("wcrt_inverse ms", ev[0].elapsed_time(ev[1]))
import os
if not os.environ.get("MFHE_NO_CHECK"):
    assert torch.equal(a, c), "W-CRT round trip"
    print("round trip ok")
