"""Two GPUs, NCCL: what the gloo tests (tests/test_dist_gloo.py, CPU engine) cannot show -- the real kernels under a
real process group.  Skipped on a one-GPU box.

  * a fit on the NON-current device (engine calls make their device current);
  * a NaN in one rank's shard raises the reference's ValueError on every rank (reference k_means.py:195-201);
  * sharded fits equal the one-rank fit of the same rows: fixed init, k-means||, k-means++ (same draws), and the
    candidate weights."""
import os
import socket
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _two_gpus():
    try:
        import torch

        return torch.cuda.is_available() and torch.cuda.device_count() >= 2
    except Exception:
        return False


pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not _two_gpus(), reason="needs two GPUs")]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _data():
    rs = np.random.RandomState(0)
    C = rs.uniform(-10, 10, size=(40, 24))
    return (C[rs.randint(0, 40, 60001)] + rs.standard_normal((60001, 24))).astype(np.float32)


def _run(rank, world, port, out):
    for p in (ROOT, HERE):
        if p not in sys.path:
            sys.path.insert(0, p)
    import importlib

    import torch
    import torch.distributed as dist

    torch.cuda.set_device(rank)
    if world > 1:
        dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world,
                                device_id=torch.device("cuda", rank))
    from dask_ml_b200.cluster import KMeans

    kmod = importlib.import_module("dask_ml_b200.cluster.k_means")
    X = _data()
    n = len(X)
    lo, hi = (rank * n) // world, ((rank + 1) * n) // world
    Xl = torch.from_numpy(X[lo:hi]).cuda()
    res = {}
    km = KMeans(n_clusters=40, init=X[:40].copy(), max_iter=4, tol=0.0).fit(Xl)
    res["lloyd"] = (km.cluster_centers_, km.inertia_, km.labels_.compute(), lo, hi)
    km2 = KMeans(n_clusters=40, init="k-means||", random_state=3, max_iter=2).fit(Xl)
    res["par"] = (km2.cluster_centers_, km2.inertia_)
    km3 = KMeans(n_clusters=12, init="k-means++", random_state=5, max_iter=2).fit(Xl)
    res["pp"] = (km3.cluster_centers_, km3.inertia_)
    res["weights"] = kmod.candidate_weights(Xl, X[[0, 5, 9, 100, 2000, 50000]])
    Xbad = Xl.clone()
    if rank == world - 1:
        Xbad[3, 2] = float("nan")
    for tag, init in (("nan_fixed", X[:40].copy()), ("nan_par", "k-means||")):
        try:
            KMeans(n_clusters=40, init=init, random_state=0, max_iter=2).fit(Xbad)
            res[tag] = "no error"
        except ValueError as e:
            res[tag] = str(e)
    torch.save(res, os.path.join(out, f"r{rank}_of{world}.pt"))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


@pytest.mark.timeout(600)
def test_two_gpus_match_one(tmp_path):
    import torch
    import torch.multiprocessing as mp

    out = str(tmp_path)
    mp.spawn(_run, args=(1, 0, out), nprocs=1, join=True)
    mp.spawn(_run, args=(2, _free_port(), out), nprocs=2, join=True)
    one = torch.load(os.path.join(out, "r0_of1.pt"), weights_only=False)
    for i in range(2):
        r = torch.load(os.path.join(out, f"r{i}_of2.pt"), weights_only=False)
        c, inertia, labels, lo, hi = r["lloyd"]
        np.testing.assert_array_equal(c, one["lloyd"][0])                    # published float32 centres: identical
        np.testing.assert_allclose(inertia, one["lloyd"][1], rtol=1e-12)
        np.testing.assert_array_equal(labels, one["lloyd"][2][lo:hi])
        np.testing.assert_allclose(r["par"][0], one["par"][0], rtol=1e-6, atol=1e-6)
        np.testing.assert_allclose(r["par"][1], one["par"][1], rtol=1e-9)
        np.testing.assert_allclose(r["pp"][0], one["pp"][0], rtol=1e-6, atol=1e-6)   # same draws -> same rows picked
        np.testing.assert_allclose(r["pp"][1], one["pp"][1], rtol=1e-9)
        np.testing.assert_array_equal(r["weights"], one["weights"])
        assert r["nan_fixed"].startswith("Input contains NaN") and r["nan_par"].startswith("Input contains NaN")


def test_fit_on_the_non_current_device():
    """cuda:0 current, X on cuda:1: every entry point keeps per-device state and launches on the current device, so
    the engine makes its own device current around each call."""
    import torch

    from dask_ml_b200.cluster import KMeans

    X = _data()[:20000]
    torch.cuda.set_device(0)
    a = KMeans(n_clusters=40, init=X[:40].copy(), max_iter=3, tol=0.0).fit(torch.from_numpy(X).to("cuda:0"))
    b = KMeans(n_clusters=40, init=X[:40].copy(), max_iter=3, tol=0.0).fit(torch.from_numpy(X).to("cuda:1"))
    assert torch.cuda.current_device() == 0
    np.testing.assert_array_equal(a.labels_.compute(), b.labels_.compute())
    np.testing.assert_array_equal(a.cluster_centers_, b.cluster_centers_)
