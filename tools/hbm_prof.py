"""Driver for ncu: one launch each of the HBM-bound kernels at BASELINE shapes."""
import sys, os
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dask_ml_b200._engine import get_engine
eng = get_engine()
g = torch.Generator(device="cpu").manual_seed(0)
# C3 shape: K2 chunk-sort kernel, K4 / K5
n, d, k = 8_000_000, 64, 1024
C = (torch.rand(k, d, generator=g, dtype=torch.float64) * 20 - 10).float()
X = eng.make_blobs(n, d, C, 1.0, seed=1)
pack = eng.pack(X[:k].double().contiguous())
lab, m2 = eng.assign(X, pack)
for _ in range(2):
    acc = eng.update(X, lab, m2, k)
mind2 = torch.empty(n, device=eng.device)
for _ in range(2):
    cost = eng.mindist_update(X, X[:1], mind2, init=True)
    cost = eng.mindist_update(X, X[7:9], mind2, init=False)
    cost = eng.mindist_update(X, X[20:26], mind2, init=False)     # six new candidates in one launch
    cost = eng.mindist_update(X, X[30:33], mind2, init=False)     # three
    eng.bernoulli(mind2, 2.0 / float(cost.item()), 1, 0, 0)
# C2 shape: K2
n2, d2, k2 = 20_000_000, 32, 256
X2 = torch.randn(n2, d2, device=eng.device); l2 = torch.randint(0, k2, (n2,), device=eng.device, dtype=torch.int32)
C2c = torch.randn(k2, d2, device=eng.device, dtype=torch.float64)
acc2 = torch.zeros(eng.acc_len(k2, d2), dtype=torch.float64, device=eng.device)
for _ in range(2):
    eng.update(X2, l2, None, k2)
    eng.update_inertia(X2, l2, C2c, k2, acc2)     # the form the Lloyd iteration uses since round 2
# C1 shape: fused K1+K2
n1, d1, k1 = 1_000_000, 16, 8
X1 = torch.randn(n1, d1, device=eng.device, dtype=torch.float64)
c1 = X1[:k1].clone(); la = torch.empty(n1, dtype=torch.int32, device=eng.device); mm = torch.empty(n1, dtype=torch.float64, device=eng.device)
a1 = torch.zeros(eng.acc_len(k1, d1), dtype=torch.float64, device=eng.device)
for _ in range(2):
    eng.assign_update(X1, c1, la, mm, a1)
torch.cuda.synchronize()
print("ok")
