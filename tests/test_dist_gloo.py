"""world_size = 2 over gloo, on CPU: the N > 1 path of the host drivers.

The CUDA kernels cannot run here, so the engine is swapped for tests/_cpu_engine.NumpyEngine (oracle
arithmetic on CPU tensors); everything else is the product code: row sharding and global row offsets,
the ONE packed allreduce per Lloyd iteration, redundant finalize, the variable-length allgather of
k-means|| picks and rows, the seed agreement, sharding-independent Philox draws."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _data():
    rs = np.random.RandomState(0)
    C = rs.uniform(-10, 10, size=(12, 6))
    X = (C[rs.randint(0, 12, 3001)] + rs.standard_normal((3001, 6))).astype(np.float32)
    return X


def _run(rank, world, port, out):
    for p in (ROOT, HERE):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch.distributed as dist

    from _cpu_engine import NumpyEngine
    from dask_ml_b200 import _comm, _engine
    from dask_ml_b200.cluster import KMeans
    import importlib

    kmod = importlib.import_module("dask_ml_b200.cluster.k_means")
    _engine.set_engine_override(NumpyEngine())
    if world > 1:
        dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    X = _data()
    n = len(X)
    lo, hi = (rank * n) // world + (7 if 0 < rank else 0), ((rank + 1) * n) // world + (7 if rank + 1 < world else 0)
    Xl = torch.from_numpy(X[lo:hi])  # uneven shards on purpose
    comm = _comm.default_comm()
    assert comm.world == world and comm.row_range(len(Xl)) == (lo, n)
    res = {}
    # Lloyd with a fixed init
    km = KMeans(n_clusters=12, init=X[:12].copy(), max_iter=5, tol=0.0).fit(Xl)
    res["lloyd"] = (km.cluster_centers_, km.inertia_, km.n_iter_, km.labels_.compute(), lo, hi)
    # k-means|| with the device-style Philox draws (keyed by GLOBAL row -> same picks for any sharding)
    km2 = KMeans(n_clusters=12, init="k-means||", random_state=3, max_iter=3).fit(Xl)
    res["par"] = (km2.cluster_centers_, km2.inertia_)
    # random_state=None: ranks must agree on a seed
    km3 = KMeans(n_clusters=5, init="k-means||", random_state=None, max_iter=2).fit(Xl)
    res["none"] = km3.cluster_centers_
    res["cost"] = kmod.evaluate_cost(Xl, X[[0, 5, 9]])
    res["picks"] = kmod._sample_points(Xl, X[[0, 5, 9]], 40, kmod.PhiloxRandomState(11))
    res["weights"] = kmod.candidate_weights(Xl, X[[0, 5, 9, 100, 2000]])
    # a NaN in the LAST rank's shard only: every rank must raise the reference's ValueError (reference :195-201) --
    # a rank raising alone would leave the others in the next collective
    Xbad = Xl.clone()
    if rank == world - 1:
        Xbad[3, 2] = float("nan")
    for tag, init in (("nan_fixed", X[:12].copy()), ("nan_par", "k-means||")):
        try:
            KMeans(n_clusters=12, init=init, random_state=0, max_iter=2).fit(Xbad)
            res[tag] = "no error"
        except ValueError as e:
            res[tag] = str(e)
    torch.save(res, os.path.join(out, f"r{rank}_of{world}.pt"))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_ranks_match_one(tmp_path):
    out = str(tmp_path)
    _run(0, 1, 0, out)                       # single process, NullComm
    from dask_ml_b200 import _engine

    _engine.set_engine_override(None)
    mp.spawn(_run, args=(2, _free_port(), out), nprocs=2, join=True)
    one = torch.load(os.path.join(out, "r0_of1.pt"), weights_only=False)
    r = [torch.load(os.path.join(out, f"r{i}_of2.pt"), weights_only=False) for i in range(2)]
    for i in range(2):
        c, inertia, n_iter, labels, lo, hi = r[i]["lloyd"]
        np.testing.assert_allclose(c, one["lloyd"][0], rtol=1e-6, atol=1e-6)
        np.testing.assert_allclose(inertia, one["lloyd"][1], rtol=1e-9)
        assert n_iter == one["lloyd"][2]
        np.testing.assert_array_equal(labels, one["lloyd"][3][lo:hi])   # each rank holds ITS rows' labels
        np.testing.assert_allclose(r[i]["par"][0], one["par"][0], rtol=1e-6, atol=1e-6)
        np.testing.assert_allclose(r[i]["par"][1], one["par"][1], rtol=1e-9)
        np.testing.assert_allclose(r[i]["cost"], one["cost"], rtol=1e-9)
        np.testing.assert_array_equal(r[i]["picks"], one["picks"])
        np.testing.assert_array_equal(r[i]["weights"], one["weights"])    # candidate weights: one allreduce of the counts
        assert r[i]["weights"].sum() == 3001
        assert r[i]["nan_fixed"].startswith("Input contains NaN") and r[i]["nan_par"].startswith("Input contains NaN")
    np.testing.assert_array_equal(r[0]["none"], r[1]["none"])           # same seed on both ranks
    np.testing.assert_allclose(r[0]["lloyd"][0], r[1]["lloyd"][0], rtol=0, atol=0)  # bit-identical centres
