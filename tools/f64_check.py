"""Throughput of the float64 K1 paths."""
import sys, os, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dask_ml_b200._engine import get_engine
eng = get_engine()
for n, d, k in [(1_000_000, 16, 8), (2_000_000, 32, 256), (4_000_000, 64, 1024), (1_000_000, 128, 4096), (1_000_000, 256, 1024)]:
    X = torch.randn(n, d, device=eng.device, dtype=torch.float64)
    C = torch.randn(k, d, device=eng.device, dtype=torch.float64) * 3
    pack = eng.pack(C)
    res = {}
    for tc in (False, True):
        eng.f64_on_tensor_cores = tc
        lab, m2 = eng.assign(X, pack)
        torch.cuda.synchronize(); t0 = time.perf_counter()
        for _ in range(3): eng.assign(X, pack, lab, m2)
        torch.cuda.synchronize(); t = (time.perf_counter() - t0) / 3
        res[tc] = (lab.clone(), m2.clone())
        print(f"f64 n={n} d={d} k={k} {'tensor cores' if tc else 'float64 kernel'}: {t*1e3:.2f} ms  {n/t/1e6:.1f} Mpts/s  "
              f"{2*n*k*d/t/1e12:.2f} TFLOP/s of 2Nkd", flush=True)
    same = int((res[False][0] != res[True][0]).sum())
    rel = float(((res[False][1] - res[True][1]).abs() / res[False][1].clamp_min(1e-300)).max())
    print(f"   labels differing: {same}   max relative distance difference: {rel:.1e}", flush=True)
