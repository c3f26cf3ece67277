"""Accuracy of K1's winner distance on data far from the origin (||x|| >> distance to the centre).
Compares mind2 / labels of every float32 K1 variant with float64 numpy on the same rows."""
import sys, os
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dask_ml_b200._engine import get_engine
eng = get_engine(); lib = eng.lib
rng = np.random.default_rng(0)
for (n, d, k, off) in [(200_000, 32, 64, 0.0), (200_000, 32, 64, 1000.0), (200_000, 64, 512, 300.0), (100_000, 128, 300, 1000.0)]:
    C = rng.uniform(-10, 10, size=(k, d))
    Xh = (C[rng.integers(0, k, n)] + rng.normal(size=(n, d)) + off).astype(np.float32)
    cent = Xh[:k].astype(np.float64)
    X64 = Xh.astype(np.float64)
    D = ((X64 ** 2).sum(1)[:, None] - 2.0 * X64 @ cent.T + (cent ** 2).sum(1)[None, :])
    lab_ref = D.argmin(1)
    m_ref = ((X64 - cent[lab_ref]) ** 2).sum(1)
    X = torch.from_numpy(Xh).cuda()
    for var, name in [(1, "ffma"), (2, "tcgen05"), (0, "auto")]:
        lib.b200km_set_option(1, var)
        pack = eng.pack(torch.from_numpy(cent).cuda())
        lab, m2 = eng.assign(X, pack)
        torch.cuda.synchronize()
        lab = lab.cpu().numpy(); m2 = m2.double().cpu().numpy()
        nz = m_ref > 0
        rel = np.abs(m2[nz] - m_ref[nz]) / m_ref[nz]
        print(f"n={n} d={d} k={k} offset={off:7.1f} {name:8s}: label mismatches {int((lab != lab_ref).sum()):4d}  "
              f"mind2 rel err max {rel.max():.2e} mean {rel.mean():.2e}  inertia rel err {abs(m2.sum() - m_ref.sum()) / m_ref.sum():.2e}  flagged {eng.flagged_count()}")
    lib.b200km_set_option(1, 0)
