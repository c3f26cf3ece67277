"""Small driver for ncu: a few launches of the tcgen05 K1 kernel (default: the C3 shape).  argv: rows variant d k [lloyd]
(lloyd = 1: the labels-only form of the Lloyd iteration, b200km_assign_lloyd_f32)"""
import sys, os
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dask_ml_b200._engine import get_engine
eng = get_engine(); lib = eng.lib
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4_000_000
d = int(sys.argv[3]) if len(sys.argv) > 3 else 64
k = int(sys.argv[4]) if len(sys.argv) > 4 else 1024
var = int(sys.argv[2]) if len(sys.argv) > 2 else 2
lloyd = len(sys.argv) > 5 and sys.argv[5] == "1"
g = torch.Generator(device="cpu").manual_seed(0)
C = (torch.rand(k, d, generator=g, dtype=torch.float64) * 20 - 10)
X = eng.make_blobs(n, d, C.float(), 1.0, seed=1)
lib.b200km_set_option(1, var)     # before the pack: the tensor-core image differs between variants 2 and 3
pack = eng.pack(X[:k].double().contiguous())
lab, m2 = eng.assign(X, pack)
kw = {"lloyd": True} if lloyd else {}
for _ in range(4):
    eng.assign(X, pack, lab, m2, **kw)
torch.cuda.synchronize()
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    eng.assign(X, pack, lab, m2, **kw)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
print(f"variant {var} n={n} d={d} k={k}: {ms:.3f} ms per assign, {n / ms / 1e6:.3f} G rows/s, {2.0 * n * k * d / ms / 1e9:.1f} TFLOP/s (2Nkd)")
print("ok", int(lab.sum().item()), eng.flagged_count())
