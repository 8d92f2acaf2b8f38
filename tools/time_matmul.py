"""Device time of mfhe_ct_matmul / mfhe_ct_matmul_decrypt at the native shape:  [MFHE_MATMUL_PATH=m] python tools/time_matmul.py [L]"""
import statistics
import sys

import torch

sys.path.insert(0, ".")
import matrix_fhe_gpu_b200.scheme as S

L = int(sys.argv[1]) if len(sys.argv) > 1 else 11
extra = [17194410241, 17194804993, 17195989249, 17196581377, 17196976129, 17197963009, 17199936769, 17201318401,
         17201515777, 17203884289, 17206844929, 17207437057, 17209608193]
g = S.Context(moduli=(S.REF_MODULI + extra)[:L])
a = g.sample_uniform()
cts = [torch.cat([a, a]) for _ in range(4)]
sk = g.generate_secret_key()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
tm, td = [], []
for i in range(12):
    ev[0].record(); prod = g.ct_matmul(*cts)
    ev[1].record(); g.ct_matmul_decrypt(prod, sk)
    ev[2].record(); torch.cuda.synchronize()
    if i >= 2:
        tm.append(ev[0].elapsed_time(ev[1])); td.append(ev[1].elapsed_time(ev[2]))
print(f"L={L}: ct_matmul {statistics.median(tm):.3f} ms, ct_matmul_decrypt {statistics.median(td):.3f} ms")
