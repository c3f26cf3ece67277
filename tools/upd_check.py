"""Timing of the K2 variants (owner table / chunk sort / atomics) on the BASELINE shapes."""
import sys, os, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dask_ml_b200._engine import get_engine
eng = get_engine(); lib = eng.lib
for n, d, k, dt in [(1_000_000, 16, 8, torch.float64), (20_000_000, 32, 256, torch.float32), (8_000_000, 64, 1024, torch.float32), (1_000_000, 128, 8192, torch.float32)]:
    X = torch.randn(n, d, device=eng.device, dtype=dt)
    labels = torch.randint(0, k, (n,), device=eng.device, dtype=torch.int32)
    m2 = torch.rand(n, device=eng.device, dtype=dt)
    ref = None
    for var in (1, 2, 3):
        lib.b200km_set_option(2, var)
        try:
            acc = eng.update(X, labels, m2, k)
            torch.cuda.synchronize(); t0 = time.perf_counter()
            for _ in range(3):
                acc.zero_(); eng.update(X, labels, m2, k, acc)
            torch.cuda.synchronize(); t = (time.perf_counter() - t0) / 3
            if ref is None: ref = acc.clone()
            err = float((acc - ref).abs().max() / ref.abs().max())
            print(f"n={n} d={d} k={k} {dt} variant {var}: {t*1e3:.3f} ms  {n*(d*X.element_size()+4)/t/1e9:.0f} GB/s  maxrel vs first {err:.1e}", flush=True)
        except Exception as e:
            print(f"n={n} d={d} k={k} variant {var}: {e}")
lib.b200km_set_option(2, 0)
