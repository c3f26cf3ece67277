"""Lloyd iterations on wide rows (embedding-like shapes, outside the BASELINE configs).  argv: rows d k iters"""
import os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dask_ml_b200._engine import get_engine
from dask_ml_b200.cluster import KMeans
n, d, k = (int(a) for a in sys.argv[1:4]) if len(sys.argv) > 3 else (2_000_000, 768, 1000)
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 5
eng = get_engine()
cen = torch.rand(k, d, dtype=torch.float64) * 2 - 1
X = eng.make_blobs(n, d, cen.float(), 0.3, seed=1)
init = X[:k].double().cpu().numpy()
KMeans(n_clusters=k, init=init, max_iter=1, tol=0.0).fit(X); torch.cuda.synchronize()
eng.trace = []
t0 = time.perf_counter()
km = KMeans(n_clusters=k, init=init, max_iter=iters, tol=0.0).fit(X)
torch.cuda.synchronize(); t = time.perf_counter() - t0
print(f"n={n} d={d} k={k}: {iters} Lloyd iterations + final labelling in {t * 1e3:.1f} ms -> {t / (iters + 1) * 1e3:.1f} ms per pass, "
      f"{(iters + 1) * 2.0 * n * k * d / t / 1e12:.1f} TFLOP/s of 2Nkd over the whole fit; inertia {km.inertia_:.6g}")
