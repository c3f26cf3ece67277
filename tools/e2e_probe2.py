import sys, os, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dask_ml_b200._engine import get_engine
eng = get_engine()
n, d, k = 4_194_304, 64, 1024
g = torch.Generator(device="cpu").manual_seed(0)
C = (torch.rand(k, d, generator=g, dtype=torch.float64) * 20 - 10).float()
X = eng.make_blobs(n, d, C, 1.0, seed=1)
host = torch.empty(n, d).pin_memory(); host.copy_(X)
pack = eng.pack(X[:k].double().contiguous())
labels = torch.empty(n, dtype=torch.int32, device=eng.device); mind2 = torch.empty(n, device=eng.device)
acc = torch.zeros(eng.acc_len(k, d), dtype=torch.float64, device=eng.device)
for trial in range(3):
    torch.cuda.synchronize()
    cur = torch.cuda.current_stream()
    start = torch.cuda.Event(enable_timing=True); start.record(cur)
    Xd = eng.to_device_streamed(host)
    cev, kev = [], []
    for r0, r1, ev in Xd._b200_pending:
        cur.wait_event(ev)
        eng.assign(Xd[r0:r1], pack, labels[r0:r1], mind2[r0:r1])
        eng.update(Xd[r0:r1], labels[r0:r1], mind2[r0:r1], k, acc)
        e = torch.cuda.Event(enable_timing=True); e.record(cur); kev.append(e)
    torch.cuda.synchronize()
    print("trial", trial, "compute done after chunk i (ms):", [round(start.elapsed_time(e), 2) for e in kev])
