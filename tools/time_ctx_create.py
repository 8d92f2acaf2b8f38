import time, sys
sys.path.insert(0, '.')
import torch
import matrix_fhe_gpu_b200.scheme as S
import bench
torch.cuda.init(); torch.zeros(1, device='cuda')
for L in (11, 24, 11):
    mods = (S.REF_MODULI + bench.L24_EXTRA)[:L]
    t = time.perf_counter(); g = S.Context(moduli=mods); torch.cuda.synchronize(); dt = time.perf_counter() - t
    print(f"ctx_create L={L}: {dt*1e3:.1f} ms")
    del g
