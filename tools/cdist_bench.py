"""K7: transform-shaped distance matrix (argv: rows d k): tensor-core kernel vs the FFMA tile kernel, GB/s of output."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dask_ml_b200._engine import get_engine
eng = get_engine()
n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 4_000_000
d = int(sys.argv[2]) if len(sys.argv) > 2 else 64
k = int(sys.argv[3]) if len(sys.argv) > 3 else 1024
g = torch.Generator(device="cpu").manual_seed(0)
C = (torch.rand(k, d, generator=g, dtype=torch.float64) * 20 - 10).float()
X = eng.make_blobs(n, d, C, 1.0, seed=1)
Y = X[:k].contiguous()
for tc in (True, False):
    eng.cdist_on_tensor_cores = tc
    out = eng.cdist(X, Y)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 5 if tc else 2
    e0.record()
    for _ in range(reps):
        out = eng.cdist(X, Y)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    print(f"{'tcgen05' if tc else 'ffma   '} n={n} d={d} k={k}: {ms:.3f} ms, {n * k * 4 / ms / 1e6:.0f} GB/s of output, "
          f"{2.0 * n * k * d / ms / 1e9:.1f} TFLOP/s (2nkd)   checksum {float(out[::997].double().sum()):.6e}")
    del out
