"""Small driver for ncu: a few launches of one K2 variant.  argv: rows d k variant"""
import sys, os
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dask_ml_b200._engine import get_engine
eng = get_engine(); lib = eng.lib
n, d, k, var = (int(a) for a in sys.argv[1:5])
X = torch.randn(n, d, device=eng.device)
labels = torch.randint(0, k, (n,), device=eng.device, dtype=torch.int32)
m2 = torch.rand(n, device=eng.device)
lib.b200km_set_option(2, var)
acc = eng.update(X, labels, m2, k)
for _ in range(3):
    acc.zero_(); eng.update(X, labels, m2, k, acc)
torch.cuda.synchronize()
print("ok", float(acc.sum()))
