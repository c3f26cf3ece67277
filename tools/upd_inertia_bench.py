"""K2 at the BASELINE shapes: b200km_update_* (inertia from a distance vector) vs b200km_update_inertia_* (inertia
summed from the rows).  argv: [rows d k]"""
import os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dask_ml_b200._engine import get_engine
eng = get_engine()
shapes = [(int(float(sys.argv[1])), int(sys.argv[2]), int(sys.argv[3]))] if len(sys.argv) > 3 else \
    [(50_000_000, 32, 256), (20_000_000, 64, 1024)]
for n, d, k in shapes:
    X = torch.randn(n, d, device=eng.device)
    labels = torch.randint(0, k, (n,), device=eng.device, dtype=torch.int32)
    m2 = torch.rand(n, device=eng.device)
    C = torch.randn(k, d, device=eng.device, dtype=torch.float64)
    acc = torch.zeros(eng.acc_len(k, d), dtype=torch.float64, device=eng.device)
    for name, fn in (("update (distance vector)", lambda: eng.update(X, labels, m2, k, acc)),
                     ("update_inertia (from rows)", lambda: eng.update_inertia(X, labels, C, k, acc))):
        fn(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            fn()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        print(f"n={n} d={d} k={k} {name:28s}: {ms:.3f} ms  {n * (d * 4 + 4) / ms / 1e6:.0f} GB/s")
