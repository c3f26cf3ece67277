"""k-means|| (init_scalable) at the C5 per-GPU shape: N = 5e8/8 rows, d = 64, k = 4096, l = 2."""
import sys, os, time, logging
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dask_ml_b200._engine import get_engine
import importlib
kmod = importlib.import_module("dask_ml_b200.cluster.k_means")
eng = get_engine()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 62_500_000
d, k = 64, 4096
g = torch.Generator(device="cpu").manual_seed(0)
C = (torch.rand(k, d, generator=g, dtype=torch.float64) * 20 - 10).float()
X = eng.make_blobs(n, d, C, 1.0, seed=5)
torch.cuda.synchronize()
# kernel-level: one K4 round (2 new candidates) and one K5 pass
mind2 = torch.empty(n, dtype=torch.float32, device=eng.device)
cost = eng.mindist_update(X, X[:1], mind2, init=True)
for m in (1, 2, 8):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(3): eng.mindist_update(X, X[5:5 + m], mind2, init=False)
    torch.cuda.synchronize(); t = (time.perf_counter() - t0) / 3
    print(f"K4 mindist_update m={m}: {t*1e3:.2f} ms  {n*(d*4+8)/t/1e9:.0f} GB/s algorithmic", flush=True)
torch.cuda.synchronize(); t0 = time.perf_counter()
for r in range(3): p = eng.bernoulli(mind2, 2.0 / float(cost.item()), 1, r, 0)
torch.cuda.synchronize(); t = (time.perf_counter() - t0) / 3
print(f"K5 bernoulli_pick: {t*1e3:.2f} ms  {n*4/t/1e9:.0f} GB/s  picked {p.numel()}", flush=True)
logging.basicConfig(level=logging.WARNING)
for rep in range(2):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    centers = kmod.init_scalable(X, k, 0, None, 2)
    torch.cuda.synchronize(); t = time.perf_counter() - t0
    print(f"init_scalable(k={k}, l=2) on {n} rows: {t:.3f} s, centers {centers.shape}", flush=True)
