"""How fast do N ranks upload pinned host memory at the same time?  (SCALE e2e efficiency: the fit of a host matrix is
bound by this copy.)  Launch with torchrun, one rank per GPU:

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/h2d_concurrency.py [--affinity]

Every rank copies a pinned 1 GiB buffer to its GPU: first each rank ALONE (the others wait at a barrier), then all
ranks TOGETHER; GB/s per rank, CUDA-event timed.  --affinity first binds the process to the CPUs NVML names as
closest to its GPU (what `numactl` / a launcher would do), so the pinned pages are allocated on that NUMA node."""
import os
import sys
import time

import torch
import torch.distributed as dist

rank = int(os.environ.get("RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
note = ""
if "--affinity" in sys.argv:
    try:
        import pynvml

        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local)
        pynvml.nvmlDeviceSetCpuAffinity(h)
        note = f"affinity={sorted(os.sched_getaffinity(0))[:4]}..({len(os.sched_getaffinity(0))} cpus)"
    except Exception as e:  # pragma: no cover
        note = f"affinity failed: {e}"
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
nbytes = 1 << 30
host = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
host.fill_(1)
dev = torch.empty(nbytes, dtype=torch.uint8, device="cuda")


def copy_rate(reps=5):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    dev.copy_(host, non_blocking=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        dev.copy_(host, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    return reps * nbytes / (e0.elapsed_time(e1) * 1e-3) / 1e9


def barrier():
    if world > 1:
        dist.barrier()


alone = 0.0
for r in range(world):
    barrier()
    if r == rank:
        alone = copy_rate()
barrier()
together = copy_rate()
barrier()
res = torch.tensor([alone, together], dtype=torch.float64, device="cuda")
if world > 1:
    allr = [torch.zeros_like(res) for _ in range(world)]
    dist.all_gather(allr, res)
else:
    allr = [res]
if rank == 0:
    print(f"world={world} {note}")
    for r, t in enumerate(allr):
        print(f"  rank {r}: alone {t[0].item():6.1f} GB/s   all {world} together {t[1].item():6.1f} GB/s")
    print(f"  aggregate together: {sum(t[1].item() for t in allr):.1f} GB/s; cpus={os.cpu_count()}")
if world > 1:
    dist.destroy_process_group()
