#!/usr/bin/env python
"""bench.py -- Lloyd-iteration throughput of the B200 k-means path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c3|c2|c1|c4] [--data blobs|overlap|uniform]
                    [--impl reference]

N > 1 is launched by the driver as ``python -m torch.distributed.run --nproc-per-node N ...``
(one rank per GPU, NCCL).  Rank 0 prints ONE JSON line.

A *step* is one Lloyd iteration over the rank's resident row shard: pack centres -> K1 fused
distance/argmin (+ fp64 re-decision of flagged rows) -> K2 sums/counts/inertia -> one allreduce of
the packed [k*d | k | 1] float64 buffer -> K3 finalize -> 8-byte shift read (the reference's one
``compute()`` per iteration, dask_ml/cluster/k_means.py:602).

Workloads (BASELINE.json configs).  The metric is quoted on C3 (N=1e9, d=64, k=1024, float32):
256 GB of X does not fit one 180 GB B200, so the default per-GPU workload is the C3 shard of the
8-GPU run (N/8 = 1.25e8 rows per GPU; WEAK scaling: at --gpus 8 the job IS C3).

The default line also carries ``other_configs``: short runs of C1, C2 and the C4 shard (N = 1 only) and, for every
N, a C5 record -- ``init_scalable`` (k-means||) alone on 5e8 x 64 rows shared out over the ranks -- plus, for N > 1,
``checks.shard_invariant``: the N-rank fit of a small problem against a one-rank fit of the same rows.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

# OpenBLAS sizes its thread tables from the machine's core count at load time; on boxes with more cores than its
# compiled-in limit it falls back to an "auxiliary array" path that has aborted the CPU legs here ("BLAS : Bad
# memory unallocation", "corrupted size vs. prev_size").  The CPU baseline runs one single-threaded BLAS call per
# row block anyway (the thread pool provides the parallelism), so pin BLAS to one thread BEFORE numpy is imported.
os.environ.setdefault("OPENBLAS_NUM_THREADS", "1")
os.environ.setdefault("OMP_NUM_THREADS", "1")

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

WORKLOADS = {
    # name: (rows per GPU, d, k, dtype, description)
    "c3": (125_000_000, 64, 1024, "float32", "C3 shard: N=1e9/8=1.25e8 rows per GPU, d=64, k=1024, float32 blobs"),
    "c2": (100_000_000, 32, 256, "float32", "C2: N=1e8 rows per GPU, d=32, k=256, float32 blobs"),
    "c1": (1_000_000, 16, 8, "float64", "C1: N=1e6 rows per GPU, d=16, k=8, float64 blobs"),
    "c4": (25_000_000, 128, 8192, "float32", "C4 shard: N=2e8/8=2.5e7 rows per GPU, d=128, k=8192, float32 blobs"),
}
# DRAM traffic of the dominant kernel per row, from the committed `ncu --set full` captures
# (dram__bytes_read.sum + dram__bytes_write.sum of one launch / rows of that launch); scaled by the rows of a
# bench launch to fill roofline.traffic.  Source files under profiles/.
NCU_TRAFFIC_BYTES_PER_ROW = {
    # filled only from captures of the CURRENT build (profiles/r2_*); absent -> roofline.traffic is null
    ("c3", "tc"): ((1.025004e9 + 33.045504e6) / 4_000_000, "profiles/r2_k1_tcgen05_c3_ncu_full_raw.csv"),
}
C5_ROWS_TOTAL = 500_000_000   # BASELINE config 5: k-means|| init only, N = 5e8, d = 64, k = 4096, l = 2
METRIC = "lloyd_points_per_sec"
UNIT = "points/s"


def _peaks():
    """(HBM GB/s, dense bf16 TFLOP/s burst, sustained, source) -- driver-measured, else the recipe's fallback."""
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return (float(p["hbm_gbs"]), float(p["bf16_tflops"]), float(p.get("bf16_tflops_sustained", p["bf16_tflops"])),
                "measured (MEASURED_PEAKS.json)")
    except Exception:
        return 6650.0, 1590.0, 1400.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clock / throttle-reason sampling DURING the timed region (profiling recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu_index = gpu_index
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                 "-i", str(self.gpu_index)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except Exception:
            self.proc.kill()
            out = ""
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# =================================================================================================
# reference arm: the reference's CPU path (oracle restatement -- dask is not installable, see
# DESIGN.md) on the host cores, same metric / config, bounded sample
# =================================================================================================
def _host_sample(n, d, k, dtype, seed=0):
    rs = np.random.RandomState(seed)
    centers = rs.uniform(-10, 10, size=(k, d))
    lab = rs.randint(0, k, size=n)
    X = centers[lab]
    X += rs.standard_normal((n, d))
    return np.ascontiguousarray(X.astype(dtype))


def cpu_reference_fit_rate(X, k, steps, warmup, cores):
    """points/s of ``KMeans(init=ndarray, max_iter=1, tol=0).fit(X)`` restated on the CPU: one Lloyd
    iteration + the reference's final labelling pass, one row block per core in a thread pool
    (what KMeans._check_array + the threaded scheduler do, reference k_means.py:192-193)."""
    from threadpoolctl import threadpool_limits

    from oracle import kmeans_oracle as orc

    chunks = orc.row_chunks(len(X), ncpu=cores)
    init = X[:k].copy()
    times = []
    # one single-threaded sklearn/BLAS call per row block, `cores` blocks in flight (dask's threaded
    # scheduler); without the limit sklearn's OpenMP x the pool oversubscribes and OpenBLAS aborts
    with threadpool_limits(limits=1):
        for it in range(warmup + steps):
            t0 = time.perf_counter()
            orc.KMeansOracle(n_clusters=k, init=init, max_iter=1, tol=0.0, chunks=chunks, n_threads=cores).fit(X)
            dt = time.perf_counter() - t0
            if it >= warmup:
                times.append(dt)
    return len(X) / (sum(times) / len(times)), sum(times) / len(times)


def cpu_lloyd_step_rate(X, k, reps, cores):
    """points/s of ONE Lloyd iteration (argmin_min + _centers_dense + bincount + divide) on the CPU."""
    from concurrent.futures import ThreadPoolExecutor

    from threadpoolctl import threadpool_limits

    from oracle import kmeans_oracle as orc

    chunks = orc.row_chunks(len(X), ncpu=cores)
    centers = X[:k].astype(np.float64)
    with threadpool_limits(limits=1), ThreadPoolExecutor(cores) as pool:
        orc.lloyd_step(X[: max(k, len(X) // 8)], centers, k, (max(k, len(X) // 8),), pool)  # numba/BLAS warm-up
        t0 = time.perf_counter()
        for _ in range(reps):
            _, _, centers, _ = orc.lloyd_step(X, centers, k, chunks, pool)
        dt = (time.perf_counter() - t0) / reps
    return len(X) / dt, dt


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n_full, d, k, dtype, desc = WORKLOADS[args.workload]
    cores = os.cpu_count() or 1
    n = args.ref_rows or {"c3": 2_000_000, "c2": 4_000_000, "c1": 1_000_000, "c4": 200_000}[args.workload]
    X = _host_sample(n, d, k, dtype)
    steps, warmup = max(1, min(args.steps, 3)), max(1, min(args.warmup, 1))
    rate, sec = cpu_reference_fit_rate(X, k, steps, warmup, cores)
    sample = (f"{n} rows of the same blobs (d={d}, k={k}, {dtype}); each step = KMeans(init=ndarray, max_iter=1, "
              f"tol=0).fit(X) restated on the CPU (oracle/kmeans_oracle.py: sklearn ArgKmin + numba scatter per row "
              f"block, {cores} threads); dask is not installable here so this is the restatement, not dask itself")
    line = {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": desc, "sample_rows": n, "d": d, "k": k},
        "cpu_baseline": {"value": rate, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# =================================================================================================
# our arm
# =================================================================================================
def _blob_centres(torch, k, d, data, tdtype):
    """Centres of the synthetic blobs.  ``blobs``: k_true = k well-separated blobs (sigma 1 in a +-10 box, the
    make_blobs defaults); ``overlap``: k_true = k/16, so every blob is shared out between ~16 of the k centres and
    neighbouring centres sit inside one Gaussian; ``uniform``: one N(0, I) cloud, no cluster structure at all."""
    g = torch.Generator(device="cpu").manual_seed(0)
    if data == "uniform":
        return torch.zeros(1, d, dtype=tdtype)
    k_true = k if data == "blobs" else max(1, k // 16)
    return (torch.rand(k_true, d, generator=g, dtype=torch.float64) * 20 - 10).to(tdtype)


def _fma_peak(torch, eng, dtype_s):
    """FP32 / FP64 FMA pipe peak measured here with the probe kernel (MEASURED_PEAKS.json has HBM + bf16 only)."""
    tdtype = torch.float32 if dtype_s == "float32" else torch.float64
    sink = torch.zeros(8, dtype=tdtype, device=eng.device)
    probe = "b200km_probe_ffma" if dtype_s == "float32" else "b200km_probe_dfma"
    iters = 4000 if dtype_s == "float32" else 1000
    pe0, pe1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 0.0
    for _ in range(4):
        pe0.record()
        eng._call(probe, sink.data_ptr(), 148 * 2, 1024, iters, eng._stream())
        pe1.record()
        torch.cuda.synchronize()
        best = max(best, 2.0 * 64 * iters * 148 * 2 * 1024 / (pe0.elapsed_time(pe1) * 1e-3) / 1e12)
    return best


def _roofline(torch, eng, args, workload, n_local, d, k, dtype_s, k1_avg_ms, ms_per_step, fused):
    """Roofline of the dominant kernel (K1) from its live CUDA-event time: algorithmic flops 2*N*k*d (or bytes
    N*(d*sizeof(T)+4) in the HBM-bound regime) / time, against the measured peak."""
    from dask_ml_b200 import _lib

    hbm_peak, bf16_peak, bf16_sust, peak_src = _peaks()
    esize = 4 if dtype_s == "float32" else 8
    flops = 2.0 * n_local * k * d
    bytes_alg = n_local * (d * esize + 4)
    fused_name = ("assign_update_bulk_kernel (K1+K2 fused: cp.async staging, scores and per-cluster sums on the FP64 "
                  "tensor cores)" if d * esize in (64, 128) else "assign_update_small_kernel (K1+K2 fused)")
    roof = {"kernel": fused_name if fused else "assign (K1)", "traffic": None}
    variant = eng.lib.b200km_get_option(_lib.OPT_K1_VARIANT)
    k1_is_tc = dtype_s == "float32" and k > 16 and d <= 128 and variant != _lib.K1_FFMA
    k1_is_ffma = dtype_s == "float32" and k > 16 and d <= 128 and not k1_is_tc
    fma_peak = _fma_peak(torch, eng, dtype_s)
    if k1_is_tc:
        # tcgen05 kind::tf32 runs at half the bf16 rate; the 3xTF32 scheme executes (3d+8)/d TF32 flops per
        # algorithmic flop (hi.hi + hi.lo + lo.hi, plus the K-step that adds ||c||^2)
        kpad = (d + 7) // 8 * 8
        overhead = (3.0 * kpad + 8.0) / d
        ach = flops / (k1_avg_ms * 1e-3) / 1e12
        peak = bf16_peak / 2.0 / overhead
        small = k <= 256 and d <= 32 and args.k1 == "auto"
        roof.update({"kernel": ("assign_tc_small_kernel (K1, tcgen05 3xTF32, resident centre image, two row groups)"
                                if small else "assign_tc_f32_kernel (K1, tcgen05 3xTF32, TMEM accumulators)"),
                     "bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
                     "peak_source": f"{peak_src}: dense bf16 burst {bf16_peak:.0f} TFLOP/s / 2 (TF32 rate) / "
                                    f"{overhead:.3f} (TF32 flops executed per algorithmic 2Nkd flop)",
                     "executed_tf32_tflops": ach * overhead,
                     "frac_of_sustained_peak": ach / (bf16_sust / 2.0 / overhead),
                     "fp32_ffma_pipe_peak_tflops": fma_peak,
                     "hbm_frac": bytes_alg / (k1_avg_ms * 1e-3) / 1e9 / hbm_peak})
    elif k1_is_ffma:
        ach = flops / (k1_avg_ms * 1e-3) / 1e12
        roof.update({"kernel": "assign_tiled_f32_kernel (K1, FP32 FFMA)", "bound": "fp32_ffma_pipe", "achieved": ach,
                     "peak": fma_peak, "unit": "TFLOP/s", "frac": ach / fma_peak,
                     "peak_source": "FFMA probe kernel (b200km_probe_ffma) timed in this run; nominal 148 SM x 128 "
                                    "x 2 x 1.965 GHz = 74.5",
                     "hbm_frac": bytes_alg / (k1_avg_ms * 1e-3) / 1e9 / hbm_peak})
    else:
        ach = bytes_alg / (k1_avg_ms * 1e-3) / 1e9
        roof.update({"bound": "hbm", "achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak,
                     "peak_source": peak_src, "fma_pipe_peak_tflops": fma_peak})
    # traffic: only from an `ncu --set full` capture of THIS build and workload (profiles/), else null
    tkey = (workload, "tc" if k1_is_tc else ("ffma" if k1_is_ffma else "other"))
    if tkey in NCU_TRAFFIC_BYTES_PER_ROW and args.data == "blobs":
        bpr, src = NCU_TRAFFIC_BYTES_PER_ROW[tkey]
        roof["traffic"] = bpr * n_local
        roof["traffic_source"] = f"{src}: {bpr:.1f} B/row measured by ncu on one launch of this kernel, x rows"
    roof["k1_ms_per_launch"] = k1_avg_ms
    roof["k1_share_of_step"] = k1_avg_ms / ms_per_step
    roof["alg_bytes_per_launch"] = bytes_alg
    roof["alg_flops_per_launch"] = flops
    return roof


def lloyd_bench(torch, dist, eng, comm, args, workload, steps, warmup, local_rank, rows=0):
    """Times ``steps`` Lloyd iterations of one workload on resident device data (CUDA events, max over ranks).
    Returns (record dict, X, init) -- the caller drops X when it is done with it."""
    world, rank = comm.world, comm.rank
    dev = eng.device
    n_local, d, k, dtype_s, desc = WORKLOADS[workload]
    if rows:
        n_local = rows
    tdtype = torch.float32 if dtype_s == "float32" else torch.float64
    esize = 4 if dtype_s == "float32" else 8

    # ---- synthetic blobs, generated on the device per shard (make_blobs distribution, keyed by global row) ----
    true_centers = _blob_centres(torch, k, d, args.data, tdtype)
    X = eng.make_blobs(n_local, d, true_centers, 1.0, seed=1234, row_offset=rank * n_local, dtype=tdtype)
    init = X[:k].to(torch.float64).contiguous().clone()
    if world > 1:
        comm.broadcast_(init, 0)  # fixed init = first k rows of the GLOBAL array
    torch.cuda.synchronize()

    acc = torch.zeros(eng.acc_len(k, d), dtype=torch.float64, device=dev)
    labels = torch.empty(n_local, dtype=torch.int32, device=dev)
    mind2 = torch.empty(n_local, dtype=tdtype, device=dev)
    c_a, c_b = init.clone(), torch.empty_like(init)
    shift_t = torch.zeros(1, dtype=torch.float64, device=dev)
    state = {"pack": None}
    fused = eng.can_fuse(k, d)
    split = not fused and eng.lloyd_split(X, k, d)

    one_launch = fused and world == 1
    ticket = torch.zeros(1, dtype=torch.int32, device=dev)

    def step(c_old, c_new):
        if one_launch:   # single GPU, tiny k and d: the whole iteration is one launch (same branch as _kmeans_single_lloyd)
            shift, state["inertia"] = eng.lloyd_small(X, c_old, labels, mind2, acc, ticket, c_new)
            return shift
        acc.zero_()
        if fused:   # HBM-bound regime: K1+K2 in one pass (same branch as _kmeans_single_lloyd)
            eng.assign_update(X, c_old, labels, mind2, acc)
        elif split:  # small k, d: labels from K1, inertia summed by K2 (same branch as _kmeans_single_lloyd)
            state["pack"] = eng.pack(c_old, state["pack"])
            eng.assign(X, state["pack"], labels, mind2, lloyd=True)
            eng.update_inertia(X, labels, c_old, k, acc)
        else:
            state["pack"] = eng.pack(c_old, state["pack"])
            eng.assign(X, state["pack"], labels, mind2)
            eng.update(X, labels, mind2, k, acc)
        comm.allreduce_sum_(acc)
        eng.finalize(acc, k, d, c_old, c_new, shift_t)
        return float(shift_t.item())

    if one_launch:
        # as _kmeans_single_lloyd runs it: iteration j + 1 is queued behind iteration j before j's shift is read (three
        # centre buffers, two label / distance buffers in rotation), every iteration's result is read on the host
        cbuf = [c_a, c_b, torch.empty_like(c_a)]
        lbuf = [labels, torch.empty_like(labels)]
        mbuf = [mind2, torch.empty_like(mind2)]
        state["q"] = 0

        def run(nsteps):
            q0 = state["q"]

            def launch(q):
                return eng.lloyd_small_launch(X, cbuf[q % 3], lbuf[q & 1], mbuf[q & 1], acc, ticket, cbuf[(q + 1) % 3],
                                              slot=q & 1)

            tag = launch(q0)
            for q in range(q0, q0 + nsteps):
                nxt = launch(q + 1) if q + 1 < q0 + nsteps else None
                _shift, state["inertia"] = eng.lloyd_small_wait(tag, slot=q & 1)
                tag = nxt
            state["q"] = q0 + nsteps
    else:
        def run(nsteps):
            nonlocal c_a, c_b
            for _ in range(nsteps):
                step(c_a, c_b)
                c_a, c_b = c_b, c_a

    run(warmup)

    # ---- timed region -----------------------------------------------------------------------
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    eng.k1_events = []
    launches0 = eng.launches
    flagged_steps = []
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    run(steps)
    ev1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms_total = ev0.elapsed_time(ev1)
    t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    clocks = sampler.stop() if rank == 0 else None
    launches = eng.launches - launches0
    k1_ms = [a.elapsed_time(b) for a, b in eng.k1_events]
    eng.k1_events = None
    flagged = eng.flagged_count() if not fused and dtype_s == "float32" else 0
    if one_launch:   # acc comes back zeroed: the counts are checked through one separate pass
        inertia = float(state["inertia"])
        chk = torch.zeros_like(acc)
        eng.assign_update(X, cbuf[state["q"] % 3], labels, mind2, chk)
        counts_ok = int(round(float(chk[k * d : k * d + k].sum().item()))) == n_local * world
    else:
        inertia = float(acc[k * d + k].item())
        counts_ok = int(round(float(acc[k * d : k * d + k].sum().item()))) == n_local * world
    ms_per_step = ms_total / steps
    k1_avg_ms = sum(k1_ms) / max(1, len(k1_ms))
    rec = {
        "workload": desc, "data": args.data, "rows_per_gpu": n_local, "rows_total": n_local * world, "d": d, "k": k,
        "dtype": "f32" if dtype_s == "float32" else "f64", "steps": steps, "warmup": warmup,
        "value": n_local * world / (ms_per_step * 1e-3), "unit": UNIT, "ms_per_step": ms_per_step,
        "k1_ms": k1_avg_ms, "gpu_launches": launches, "clocks": clocks,
        "checks": {"counts_sum_to_n": counts_ok, "inertia": inertia, "rows_refined_fp64_last_iter": flagged,
                   "flagged_fraction": flagged / n_local},
    }
    if rank == 0:
        rec["roofline"] = _roofline(torch, eng, args, workload, n_local, d, k, dtype_s, k1_avg_ms, ms_per_step, fused)
    del acc, labels, mind2
    return rec, X, init


def e2e_bench(torch, dist, eng, comm, args, X, init, d, k, dtype_s):
    """The user's call with HOST buffers: H2D + fit + D2H inside the timed region; every rank fits ITS pinned host
    batch (the fit's collectives need all ranks); wall clock, max over ranks."""
    from dask_ml_b200.cluster import KMeans

    world = comm.world
    tdtype = X.dtype
    esize = X.element_size()
    n_e = min(X.shape[0], args.e2e_rows)
    host = torch.empty(n_e, d, dtype=tdtype).pin_memory()
    host.copy_(X[:n_e])
    init_h = init.cpu().numpy().astype(np.float32 if dtype_s == "float32" else np.float64)
    torch.cuda.synchronize()
    for _ in range(2):
        KMeans(n_clusters=k, init=init_h, max_iter=1, tol=0.0).fit(host)
    es = max(3, min(args.steps, 5))
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(es):
        km = KMeans(n_clusters=k, init=init_h, max_iter=1, tol=0.0).fit(host)
        _ = km.inertia_
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / es
    tt = torch.tensor([dt], dtype=torch.float64, device=eng.device)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    dt = float(tt.item())
    return {"value": n_e * world / dt, "unit": UNIT,
            "h2d_bytes_per_step": n_e * d * esize + k * d * 8,
            "d2h_bytes_per_step": k * d * esize + 8 + 8,
            "rows_per_step": n_e * world, "ms_per_step": dt * 1e3,
            "call": "KMeans(n_clusters=k, init=ndarray, max_iter=1, tol=0).fit(pinned host tensor) on every rank: "
                    "H2D copy, finite check, 1 Lloyd iteration, the final (non-squared) labelling pass, D2H of "
                    "centres and inertia (bytes are per rank)"}


def c5_bench(torch, dist, eng, comm, args):
    """BASELINE config 5: k-means|| initialisation only (init_scalable, oversampling_factor 2, k = 4096, d = 64) on
    N = 5e8 rows shared out over the ranks (STRONG scaling: the same global rows whatever the number of GPUs, so the
    rows picked -- Philox draws keyed by global row -- must be the same; their hash is in the record).  One GPU holds
    the 128 GB if its memory allows; otherwise the row count is cut to what fits and said so."""
    import hashlib
    import importlib

    kmod = importlib.import_module("dask_ml_b200.cluster.k_means")
    world, rank = comm.world, comm.rank
    n_total, d, k = C5_ROWS_TOTAL if not args.c5_rows else args.c5_rows, 64, 4096
    torch.cuda.empty_cache()
    free, _tot = torch.cuda.mem_get_info()
    per_row = d * 4 + 16
    fit_rows = int(0.9 * free // per_row) * world
    ft = torch.tensor([fit_rows], dtype=torch.float64, device=eng.device)
    if world > 1:
        dist.all_reduce(ft, op=dist.ReduceOp.MIN)
    n_total = min(n_total, int(ft.item()))
    n_local = n_total // world
    n_total = n_local * world
    true_centers = _blob_centres(torch, k, d, "blobs", torch.float32)
    X = eng.make_blobs(n_local, d, true_centers, 1.0, seed=77, row_offset=rank * n_local, dtype=torch.float32)
    torch.cuda.synchronize()
    out = {}
    for rep in range(2):     # the first call also pays first-use costs (buffers, module loads); report the second
        rs = kmod.PhiloxRandomState(5)
        eng.trace = []
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        e0 = torch.cuda.Event(enable_timing=True)
        e0.record()
        C = kmod.init_scalable(X, k, rs, None, 2)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        trace, eng.trace = eng.trace, None
        tt = torch.tensor([dt], dtype=torch.float64, device=eng.device)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt = float(tt.item())
        # per-launch time of K4 (b200km_mindist_update_f32): gap between the events recorded after consecutive launches
        k4 = []
        prev = e0
        for name, ev in trace:
            if name.startswith("b200km_mindist_update"):
                k4.append(prev.elapsed_time(ev))
            prev = ev
        rounds = len(rs.picks)
        picked = np.unique(np.concatenate([np.zeros(1, np.int64)] + [p.astype(np.int64) for p in rs.picks]))
        out = {"seconds": dt, "rounds": rounds, "candidates": int(len(picked)),
               "picked_sha1": hashlib.sha1(picked.tobytes()).hexdigest(),
               "centres_sha1": hashlib.sha1(np.ascontiguousarray(C).tobytes()).hexdigest(),
               "ms_per_round": dt * 1e3 / max(1, rounds)}
        if k4:
            k4_med = statistics.median(k4)
            hbm_peak = _peaks()[0]
            bytes_round = n_local * (d * 4 + 2 * 4)
            out.update({"k4_ms_median": k4_med, "k4_launches": len(k4),
                        "k4_hbm_gbs": bytes_round / (k4_med * 1e-3) / 1e9,
                        "k4_hbm_frac": bytes_round / (k4_med * 1e-3) / 1e9 / hbm_peak,
                        "k4_alg_bytes_per_launch": bytes_round,
                        "k4_peak_note": "peak = MEASURED_PEAKS.json's copy bandwidth (half reads, half writes); K4 is 97 % "
                                        "reads (TMA ring), which this HBM serves faster than a copy: the fraction can exceed 1"})
    out.update({"workload": f"C5: init_scalable only, oversampling_factor=2, k={k}, N={n_total} rows x d={d} float32 "
                            f"blobs over {world} GPU(s) ({n_local} rows each), random_state=5",
                "rows_total": n_total, "rows_per_gpu": n_local, "k": k, "d": d, "scaling": "strong",
                "exchange": "per round: 1 scalar allreduce (cost) + 2 variable-length allgathers (picked indices, their rows)"})
    del X
    torch.cuda.empty_cache()
    return out


def shard_invariance_check(torch, dist, eng, comm):
    """N ranks on N shards against ONE rank on the same global rows, on the real kernels and NCCL: every rank fits its
    shard of a 2e6 x 64, k = 1024 problem (3 Lloyd iterations, fixed init), rank 0 also fits the whole array alone
    (``_comm.local_only``).  The only difference is the order of the float64 sums (the allreduce replaces the
    reference's block-order chain, dask_ml/cluster/k_means.py:595-598)."""
    from dask_ml_b200 import _comm
    from dask_ml_b200.cluster import KMeans

    world, rank = comm.world, comm.rank
    n, d, k = 2_000_000, 64, 1024
    n_local = n // world
    n = n_local * world
    true_centers = _blob_centres(torch, k, d, "blobs", torch.float32)
    Xl = eng.make_blobs(n_local, d, true_centers, 1.0, seed=99, row_offset=rank * n_local, dtype=torch.float32)
    init = Xl[:k].to(torch.float64).contiguous().clone()
    comm.broadcast_(init, 0)
    init_h = init.cpu().numpy().astype(np.float32)
    km = KMeans(n_clusters=k, init=init_h, max_iter=3, tol=0.0).fit(Xl)
    whole_labels = torch.empty(n, dtype=torch.int32, device=eng.device)
    ref = torch.zeros(k * d + 1, dtype=torch.float64, device=eng.device)
    if rank == 0:
        Xw = eng.make_blobs(n, d, true_centers, 1.0, seed=99, row_offset=0, dtype=torch.float32)
        with _comm.local_only():
            km1 = KMeans(n_clusters=k, init=init_h, max_iter=3, tol=0.0).fit(Xw)
        whole_labels.copy_(km1.labels_.tensor if hasattr(km1.labels_, "tensor") else torch.as_tensor(km1.labels_.compute()))
        ref[: k * d] = torch.from_numpy(km1.cluster_centers_.astype(np.float64).reshape(-1)).to(eng.device)
        ref[k * d] = km1.inertia_
        del Xw
    comm.broadcast_(whole_labels, 0)
    comm.broadcast_(ref, 0)
    mine = km.labels_.tensor if hasattr(km.labels_, "tensor") else torch.as_tensor(km.labels_.compute()).to(eng.device)
    diff = (mine != whole_labels[rank * n_local : (rank + 1) * n_local]).sum().to(torch.float64).reshape(1)
    comm.allreduce_sum_(diff)
    c1 = ref[: k * d].cpu().numpy().reshape(k, d)
    cn = km.cluster_centers_.astype(np.float64)
    scale = np.abs(c1).max()
    inertia1 = float(ref[k * d].item())
    return {"rows": n, "d": d, "k": k, "iterations": 3, "labels_differ": int(diff.item()),
            "labels_equal": int(diff.item()) == 0,
            "centers_max_rel": float(np.abs(cn - c1).max() / scale),
            "inertia_rel": abs(km.inertia_ - inertia1) / abs(inertia1),
            "note": "centres are compared after the cast to X's dtype (float32), as the estimator publishes them"}


def run_ours(args):
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    from dask_ml_b200 import _comm, _lib
    from dask_ml_b200._engine import get_engine

    eng = get_engine()
    eng.lib.b200km_set_option(_lib.OPT_K1_VARIANT,
                              {"auto": _lib.K1_AUTO, "ffma": _lib.K1_FFMA, "tc": _lib.K1_TCGEN05}[args.k1])
    comm = _comm.default_comm()
    n_local, d, k, dtype_s, desc = WORKLOADS[args.workload]
    esize = 4 if dtype_s == "float32" else 8

    rec, X, init = lloyd_bench(torch, dist, eng, comm, args, args.workload, args.steps, args.warmup, local_rank, args.rows)
    n_local = rec["rows_per_gpu"]
    e2e = None if args.no_e2e else e2e_bench(torch, dist, eng, comm, args, X, init, d, k, dtype_s)

    # ---- CPU baseline on a bounded sample (rank 0, N=1 only) -------------------------------------
    cpu = None
    if world == 1 and not args.no_cpu:
        cores = os.cpu_count() or 1
        n_s = args.cpu_rows or {"c3": 4_000_000, "c2": 8_000_000, "c1": 1_000_000, "c4": 400_000}[args.workload]
        n_s = min(n_s, n_local)
        Xs = X[:n_s].cpu().numpy()
        rate, sec = cpu_lloyd_step_rate(Xs, k, 3, cores)
        cpu = {"value": rate, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": f"first {n_s} rows of the same device-generated blobs; 1 warm-up + 3 timed Lloyd iterations "
                         f"of oracle/kmeans_oracle.py (sklearn ArgKmin float64 + numba scatter, one row block per "
                         f"core, {cores} threads; the oracle is pinned to the unmodified reference modules by "
                         f"tests/test_reference_pin.py); {sec * 1e3:.1f} ms per iteration",
               "gflops": 2.0 * n_s * k * d / sec / 1e9}
    del X
    torch.cuda.empty_cache()

    # ---- the other BASELINE configs, so that every one of them is in front of the driver ----------------
    others = {}
    checks = rec["checks"]
    if args.workload == "c3" and not args.no_others and not args.rows:
        if world == 1:
            for w in ("c1", "c2", "c4"):
                o_steps = {"c1": 50, "c2": 5, "c4": 3}[w]
                r, Xo, _ = lloyd_bench(torch, dist, eng, comm, args, w, o_steps, 3, local_rank)
                del Xo
                torch.cuda.empty_cache()
                if "roofline" in r:
                    r["roofline_frac"] = r["roofline"]["frac"]
                others[w] = r
        others["c5"] = c5_bench(torch, dist, eng, comm, args)
        if world > 1:
            checks["shard_invariant"] = shard_invariance_check(torch, dist, eng, comm)

    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    line = {
        "metric": METRIC, "value": rec["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": rec["ms_per_step"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": rec["dtype"], "data": "synthetic",
        "config": {"workload": desc, "rows_per_gpu": n_local, "rows_total": n_local * world, "d": d, "k": k,
                   "init": "first k rows", "blobs": args.data,
                   "l2": f"inputs {n_local * d * esize / 1e9:.1f} GB per GPU >> 126 MB L2 (no flush needed)",
                   "parallelism": f"rows sharded over {world} GPU(s), 1 allreduce/iter"},
        "iters_per_sec": 1e3 / rec["ms_per_step"],
        "roofline": rec["roofline"], "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": rec["gpu_launches"],
        "clocks": rec["clocks"], "checks": checks, "other_configs": others or None,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--rows", type=int, default=0, help="override rows per GPU")
    ap.add_argument("--e2e-rows", type=int, default=4_194_304)
    ap.add_argument("--cpu-rows", type=int, default=0)
    ap.add_argument("--ref-rows", type=int, default=0)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-others", action="store_true", help="skip the other_configs records (c1, c2, c4, c5)")
    ap.add_argument("--c5-rows", type=int, default=0, help="override the TOTAL rows of the c5 record")
    ap.add_argument("--data", default="blobs", choices=["blobs", "overlap", "uniform"],
                    help="blobs: k_true = k separated blobs; overlap: k_true = k/16; uniform: one N(0, I) cloud")
    ap.add_argument("--k1", default="auto", choices=["auto", "ffma", "tc"], help="K1 kernel variant (float32)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
