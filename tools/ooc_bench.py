"""Out-of-core Lloyd iterations: X stays in page-locked host memory and is streamed through the device on every
pass (dask_ml_b200/_hostmatrix.py).  Reports rows/s and the PCIe rate against a plain pinned host->device copy of
the same buffer.  argv: rows d k [iters]"""
import os, sys, time
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["B200KM_OUT_OF_CORE"] = "1"
from dask_ml_b200._engine import get_engine
from dask_ml_b200.cluster import KMeans

n, d, k = (int(a) for a in sys.argv[1:4]) if len(sys.argv) > 3 else (16_000_000, 64, 1024)
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 3
eng = get_engine()
cen = torch.rand(k, d, dtype=torch.float64) * 20 - 10
Xd = eng.make_blobs(n, d, cen.float(), 1.0, seed=1)
host = torch.empty(n, d, dtype=torch.float32).pin_memory()
host.copy_(Xd); init = Xd[:k].double().cpu().numpy(); del Xd
torch.cuda.synchronize()
dst = torch.empty_like(host, device=eng.device)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
dst.copy_(host, non_blocking=True); torch.cuda.synchronize()
e0.record(); dst.copy_(host, non_blocking=True); e1.record(); torch.cuda.synchronize()
h2d = host.numel() * 4 / (e0.elapsed_time(e1) * 1e-3) / 1e9
del dst
KMeans(n_clusters=k, init=init, max_iter=1, tol=0.0).fit(host)          # warm-up (page-locks, allocates buffers)
t0 = time.perf_counter()
km = KMeans(n_clusters=k, init=init, max_iter=iters, tol=0.0).fit(host)
torch.cuda.synchronize(); t = time.perf_counter() - t0
passes = iters + 1                                                       # + the final labelling pass
gbs = passes * host.numel() * 4 / t / 1e9
print(f"out-of-core n={n} d={d} k={k}: {iters} iterations + final pass in {t:.3f} s -> {passes * n / t / 1e6:.1f} M rows/s per pass, "
      f"{gbs:.1f} GB/s over PCIe ({gbs / h2d:.2f} of the plain pinned copy's {h2d:.1f} GB/s); inertia {km.inertia_:.6g}")
