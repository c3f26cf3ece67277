import sys, os, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dask_ml_b200._engine import get_engine
from dask_ml_b200.cluster import KMeans
eng = get_engine()
n, d, k = 4_194_304, 64, 1024
g = torch.Generator(device="cpu").manual_seed(0)
C = (torch.rand(k, d, generator=g, dtype=torch.float64) * 20 - 10).float()
X = eng.make_blobs(n, d, C, 1.0, seed=1)
host = torch.empty(n, d).pin_memory(); host.copy_(X)
init = X[:k].cpu().numpy()
for _ in range(2): KMeans(n_clusters=k, init=init, max_iter=1, tol=0.0).fit(host)
torch.cuda.synchronize()
start = torch.cuda.Event(enable_timing=True); start.record()
eng.trace = []
t0 = time.perf_counter()
KMeans(n_clusters=k, init=init, max_iter=1, tol=0.0).fit(host)
torch.cuda.synchronize()
print("wall %.2f ms" % ((time.perf_counter() - t0) * 1e3))
for name, ev in eng.trace:
    print("%8.2f  %s" % (start.elapsed_time(ev), name))
