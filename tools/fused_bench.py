"""The fused K1+K2 kernels at small k, d (BASELINE C1 and neighbours): the sorted-tile kernel
(B200KM_OPT_FUSED_VARIANT = 1) against the cp.async kernel (0 = auto), CUDA-event time per launch of
b200km_assign_update_*.  argv: [rows d k f32|f64]"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dask_ml_b200 import _lib
from dask_ml_b200._engine import get_engine
eng = get_engine()
if len(sys.argv) > 4:
    shapes = [(int(float(sys.argv[1])), int(sys.argv[2]), int(sys.argv[3]), sys.argv[4])]
else:
    shapes = [(1_000_000, 16, 8, "f64"), (10_000_000, 16, 8, "f64"), (10_000_000, 8, 16, "f64"),
              (10_000_000, 16, 8, "f32"), (10_000_000, 32, 16, "f32"), (10_000_000, 32, 8, "f32")]
REPS = 50
for n, d, k, ts in shapes:
    dt = torch.float64 if ts == "f64" else torch.float32
    X = torch.randn(n, d, device=eng.device, dtype=dt) + torch.randint(0, k, (n, 1), device=eng.device).to(dt) * 3
    C = X[:k].double().contiguous()
    labels = torch.empty(n, dtype=torch.int32, device=eng.device)
    m2 = torch.empty(n, dtype=dt, device=eng.device)
    acc = torch.zeros(eng.acc_len(k, d), dtype=torch.float64, device=eng.device)
    alg = n * (d * X.element_size() + 4)
    for variant, name in ((1, "sorted tile"), (0, "cp.async   ")):
        eng.lib.b200km_set_option(_lib.OPT_FUSED_VARIANT, variant)
        for _ in range(3):
            eng.assign_update(X, C, labels, m2, acc)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(REPS):
            eng.assign_update(X, C, labels, m2, acc)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / REPS
        print(f"n={n} d={d} k={k} {ts} {name}: {ms * 1e3:8.1f} us  {alg / ms / 1e6:6.0f} GB/s (algorithmic)", flush=True)
    eng.lib.b200km_set_option(_lib.OPT_FUSED_VARIANT, 0)
