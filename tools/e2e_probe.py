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
def t(fn, reps=5):
    fn(); torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(reps): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / reps * 1e3
dst = torch.empty_like(X)
print("H2D whole  : %.2f ms" % t(lambda: dst.copy_(host, non_blocking=True)))
def chunked():
    step = n // 8
    for r0 in range(0, n, step): dst[r0:r0+step].copy_(host[r0:r0+step], non_blocking=True)
print("H2D 8 chunk: %.2f ms" % t(chunked))
print("fit resident: %.2f ms" % t(lambda: KMeans(n_clusters=k, init=init, max_iter=1, tol=0.0).fit(X)))
print("fit host    : %.2f ms" % t(lambda: KMeans(n_clusters=k, init=init, max_iter=1, tol=0.0).fit(host)))
eng.STREAM_MIN_BYTES = 1 << 60
print("fit host (no streaming): %.2f ms" % t(lambda: KMeans(n_clusters=k, init=init, max_iter=1, tol=0.0).fit(host)))
import cProfile, pstats
eng.STREAM_MIN_BYTES = 64 << 20
pr = cProfile.Profile(); pr.enable()
for _ in range(3): KMeans(n_clusters=k, init=init, max_iter=1, tol=0.0).fit(host)
pr.disable(); pstats.Stats(pr).sort_stats("cumulative").print_stats(18)
