"""-m gpu: BASELINE.json's FULL single-GPU sizes, checked through size-independent properties (the oracle
cannot finish these sizes): conservation (counts sum to N, per-cluster sums add up to the column sums),
inertia = sum of the per-row minima, idempotence, shard invariance, Lloyd monotonicity, plus exact oracle
labels on a random row sample."""
import numpy as np
import numpy.testing as npt
import pytest

from oracle import kmeans_oracle as orc

pytestmark = pytest.mark.gpu


def _run(n, d, k, dtype_name, sample=20000):
    import torch
    from dask_ml_b200._engine import get_engine

    torch.cuda.set_device(0)
    eng = get_engine()
    tdt = getattr(torch, dtype_name)
    g = torch.Generator(device="cpu").manual_seed(0)
    true_c = (torch.rand(k, d, generator=g, dtype=torch.float64) * 20 - 10).to(tdt)
    X = eng.make_blobs(n, d, true_c, 1.0, seed=77, dtype=tdt)
    c0 = X[:k].double().contiguous()
    pack = eng.pack(c0)
    labels, mind2 = eng.assign(X, pack)
    # idempotence
    labels2, _ = eng.assign(X, pack)
    assert torch.equal(labels, labels2)
    # shard invariance: the two halves labelled separately
    h = n // 2 + 13
    la, _ = eng.assign(X[:h], pack)
    lb, _ = eng.assign(X[h:], pack)
    assert torch.equal(torch.cat([la, lb]), labels)
    assert int(labels.min()) >= 0 and int(labels.max()) < k
    # conservation
    acc = eng.update(X, labels, mind2, k)
    sums = acc[: k * d].view(k, d)
    counts = acc[k * d : k * d + k]
    assert int(round(float(counts.sum()))) == n
    colsum = torch.zeros(d, dtype=torch.float64, device=eng.device)
    step = 1 << 22
    for r0 in range(0, n, step):   # torch only as the checker, in float64, blockwise to bound memory
        colsum += X[r0 : r0 + step].double().sum(0)
    npt.assert_allclose(sums.sum(0).cpu().numpy(), colsum.cpu().numpy(), rtol=1e-9, atol=1e-6 * n ** 0.5)
    inertia = float(acc[-1])
    npt.assert_allclose(inertia, float(eng.reduce_sum(mind2)), rtol=1e-12)
    # exact labels / distances on a random sample, against the float64 oracle
    idx = np.random.RandomState(1).choice(n, sample, replace=False)
    idx.sort()
    Xs = X[torch.from_numpy(idx).to(eng.device)].cpu().numpy()
    lab_ref, d_ref = orc.pairwise_distances_argmin_min(Xs, c0.cpu().numpy(), metric_kwargs={"squared": True})
    got = labels[torch.from_numpy(idx).to(eng.device)].cpu().numpy()
    npt.assert_array_equal(got, lab_ref)
    tol = 1e-5 if dtype_name == "float32" else 1e-10
    npt.assert_allclose(mind2[torch.from_numpy(idx).to(eng.device)].cpu().numpy(), d_ref, rtol=3 * tol, atol=tol * d_ref.max())
    # one Lloyd step cannot increase the inertia
    new_c, shift = eng.finalize(acc, k, d, c0)
    pack2 = eng.pack(new_c)
    l2, m2 = eng.assign(X, pack2)
    inertia2 = float(eng.reduce_sum(m2))
    assert inertia2 <= inertia * (1 + 1e-9)
    assert float(shift) > 0


def test_c2_full_size():   # KMeans k=256 on N=1e8, d=32 float32 (12.8 GB)
    _run(100_000_000, 32, 256, "float32")


def test_c3_shard_full_size():   # the per-GPU shard of N=1e9, d=64, k=1024 float32 (32 GB)
    _run(125_000_000, 64, 1024, "float32", sample=8000)


def test_c1_full_size():   # KMeans k=8 on N=1e6, d=16 float64
    _run(1_000_000, 16, 8, "float64")


def test_c4_shard():   # the per-GPU shard of N=2e8, d=128, k=8192 float32 (12.8 GB)
    _run(25_000_000, 128, 8192, "float32", sample=3000)
