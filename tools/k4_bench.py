"""K4 per pass over X: B200KM_OPT_K4_VARIANT = 1 (register-path kernel, 8 candidates per launch; K1 route from 48
candidates on) against auto (TMA ring up to 4 candidates, one thread per row for 3 .. 8 per launch, one tensor-core pass for 25 .. 128).  argv: [rows d]"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dask_ml_b200 import _lib
from dask_ml_b200._engine import get_engine
eng = get_engine()
n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 60_000_000
d = int(sys.argv[2]) if len(sys.argv) > 2 else 64
X = torch.randn(n, d, device=eng.device)
mind2 = torch.full((n,), 1e30, device=eng.device)
for m in (1, 2, 3, 4, 5, 8, 16, 40, 128):
    cand = X[100 : 100 + m].contiguous()
    for variant, name in ((1, "registers / K1 route "), (0, "auto (ring, rows, TC) ")):
        eng.lib.b200km_set_option(_lib.OPT_K4_VARIANT, variant)
        for _ in range(2):
            eng.mindist_update(X, cand, mind2, init=False)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            eng.mindist_update(X, cand, mind2, init=False)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        print(f"n={n} d={d} m={m} {name}: {ms:7.3f} ms  {n * (d * 4 + 8) / ms / 1e6:6.0f} GB/s", flush=True)
eng.lib.b200km_set_option(_lib.OPT_K4_VARIANT, 0)
