"""SpectralClustering tail: time of fit() and of its affinity kernel (b200km_pairwise_kernel_f32),
the only hand-written kernel on that path; the Nystrom algebra is small library GEMMs / an SVD of a 100 x 100 block.
argv: rows d n_components n_clusters"""
import os, sys, time
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dask_ml_b200._engine import get_engine
from dask_ml_b200.cluster import SpectralClustering
from dask_ml_b200.metrics import pairwise_kernels

n, d, l, k = (int(a) for a in sys.argv[1:5]) if len(sys.argv) > 4 else (2_000_000, 64, 100, 8)
eng = get_engine()
cen = torch.rand(k, d, dtype=torch.float64) * 4 - 2
X = eng.make_blobs(n, d, cen.float(), 0.5, seed=1)
torch.cuda.synchronize()
sc = SpectralClustering(n_clusters=k, n_components=l, gamma=1.0 / d, random_state=0, kmeans_params={"max_iter": 20})
sc.fit(X); torch.cuda.synchronize()
t0 = time.perf_counter(); sc.fit(X); torch.cuda.synchronize(); t_fit = time.perf_counter() - t0
# the affinity block B = rbf(X_keep, X_rest): l x (n - l)
Xk, Xr = X[:l].contiguous(), X[l:]
B = pairwise_kernels(Xk, Xr, metric="rbf", gamma=1.0 / d); torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    B = pairwise_kernels(Xk, Xr, metric="rbf", gamma=1.0 / d)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
alg = (n - l) * (d * 4 + l * 4)
print(f"SpectralClustering n={n} d={d} n_components={l} n_clusters={k}: fit {t_fit * 1e3:.1f} ms ({n / t_fit / 1e6:.1f} M rows/s); "
      f"affinity block rbf({l} x {n - l}): {ms:.3f} ms, {alg / ms / 1e6:.0f} GB/s of algorithmic bytes "
      f"(read X once + write the block), {2.0 * l * (n - l) * d / ms / 1e9:.1f} TFLOP/s")
