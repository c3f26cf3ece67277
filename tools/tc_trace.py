"""Timeline of the small-k tcgen05 kernel's first steps on CTA 0 (library built with make EXTRA=-DKM_TC_TRACE)."""
import sys, os, ctypes
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dask_ml_b200._engine import get_engine
eng = get_engine(); lib = eng.lib
n, d, k = 4_000_000, int(sys.argv[1]) if len(sys.argv) > 1 else 32, int(sys.argv[2]) if len(sys.argv) > 2 else 256
g = torch.Generator(device="cpu").manual_seed(0)
C = (torch.rand(k, d, generator=g, dtype=torch.float64) * 20 - 10)
X = eng.make_blobs(n, d, C.float(), 1.0, seed=1)
pack = eng.pack(X[:k].double().contiguous())
lab, m2 = eng.assign(X, pack)
kw = {"lloyd": True} if (len(sys.argv) > 5 and sys.argv[5] == "1") else {}
for _ in range(3):
    eng.assign(X, pack, lab, m2, **kw)
torch.cuda.synchronize()
buf = (ctypes.c_longlong * (4 * 256 * 8))()
raw = ctypes.CDLL(lib._name)
raw.b200km_debug_read_trace(buf, 0)
names = {0: "mma issue", 1: "iter start", 3: "stage+issue done", 2: "a_full seen", 4: "turn seen", 5: "track done",
         6: "output done"}
ev = []
for role in range(4):
    for step in range(256):
        for e in names:
            t = buf[(role * 256 + step) * 8 + e]
            if t:
                ev.append((t, role, step, e))
ev.sort()
lo, hi = (int(sys.argv[3]), int(sys.argv[4])) if len(sys.argv) > 4 else (20, 24)
sel = [x for x in ev if lo <= x[2] < hi]
t0 = sel[0][0]
for t, role, step, e in sel:
    who = f"MMA(g{role - 2})" if role >= 2 else f"group{role}"
    print(f"{t - t0:9d}  {who:8s} step {step:4d}  {names[e]}")
