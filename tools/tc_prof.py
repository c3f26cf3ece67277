"""Small driver for ncu: a few launches of the tcgen05 K1 kernel at the C3 shape."""
import sys, os
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dask_ml_b200._engine import get_engine
eng = get_engine(); lib = eng.lib
n, d, k = int(sys.argv[1]) if len(sys.argv) > 1 else 4_000_000, 64, 1024
var = int(sys.argv[2]) if len(sys.argv) > 2 else 2
g = torch.Generator(device="cpu").manual_seed(0)
C = (torch.rand(k, d, generator=g, dtype=torch.float64) * 20 - 10)
X = eng.make_blobs(n, d, C.float(), 1.0, seed=1)
pack = eng.pack(X[:k].double().contiguous())
lib.b200km_set_option(1, var)
lab, m2 = eng.assign(X, pack)
for _ in range(4):
    eng.assign(X, pack, lab, m2)
torch.cuda.synchronize()
print("ok", int(lab.sum().item()), eng.flagged_count())
