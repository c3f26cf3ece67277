"""Where does a k-means|| round go?  init_scalable on a C5-shaped shard (argv: rows [k]); per entry point: launches,
total and median GPU time between the event recorded after the previous launch and the one after this launch
(so host latency in front of a launch is charged to it), plus the wall time."""
import os, sys, time, collections, statistics
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import importlib
from dask_ml_b200._engine import get_engine
kmod = importlib.import_module("dask_ml_b200.cluster.k_means")
eng = get_engine()
n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 62_500_000
k = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
d = 64
g = torch.Generator(device="cpu").manual_seed(0)
C = (torch.rand(k, d, generator=g, dtype=torch.float64) * 20 - 10).float()
X = eng.make_blobs(n, d, C, 1.0, seed=77)
for rep in range(2):
    rs = kmod.PhiloxRandomState(5)
    eng.trace = []
    torch.cuda.synchronize(); t0 = time.perf_counter()
    e0 = torch.cuda.Event(enable_timing=True); e0.record()
    kmod.init_scalable(X, k, rs, None, 2)
    torch.cuda.synchronize(); wall = time.perf_counter() - t0
    tr, eng.trace = eng.trace, None
agg = collections.defaultdict(list)
prev = e0
for name, ev in tr:
    agg[name].append(prev.elapsed_time(ev)); prev = ev
print(f"rows {n} k {k}: wall {wall*1e3:.1f} ms, rounds {len(rs.picks)}, gpu-timeline total {e0.elapsed_time(prev):.1f} ms")
for name, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
    print(f"  {name:34s} x{len(v):3d}  total {sum(v):8.2f} ms  median {statistics.median(v):7.3f}  max {max(v):7.3f}")
k4 = agg.get("b200km_mindist_update_f32", [])
if k4:   # one K4 launch per round that found new candidates: its time against the number of candidates it took
    m_of = [len(set(int(v) for v in p)) for p in rs.picks]
    print("  K4 launches (ms):", " ".join(f"{t:.1f}" for t in k4))
    print("  picks per round :", " ".join(str(m) for m in m_of))
