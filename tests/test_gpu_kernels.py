"""-m gpu: every C-ABI entry point against the CPU oracle on seeded inputs.

Bars: labels / indices bit-exact (float32 X included: K1 + fp64 re-decision must reproduce the
float64 ArgKmin labels; rows whose two nearest float64 distances agree to 1e-6 relative are
reported and excluded); float values within 1e-5 relative (float32) / 1e-10 (float64)."""
import numpy as np
import numpy.testing as npt
import pytest
import sklearn.metrics

from conftest import blobs
from oracle import kmeans_oracle as orc

pytestmark = pytest.mark.gpu

RTOL = {np.float32: 1e-5, np.float64: 1e-10}


@pytest.fixture(scope="module")
def eng():
    import torch
    from dask_ml_b200._engine import get_engine

    torch.cuda.set_device(0)
    return get_engine()


def _near_tie_mask(X, C, rel=1e-6):
    """Rows whose two smallest float64 squared distances tie within `rel` (north_star's exemption)."""
    D = sklearn.metrics.pairwise_distances(X.astype(np.float64), C.astype(np.float64), metric="sqeuclidean")
    if D.shape[1] < 2:
        return np.zeros(len(X), bool)
    part = np.partition(D, 1, axis=1)
    return (part[:, 1] - part[:, 0]) <= rel * np.maximum(part[:, 1], 1e-300)


def _check_assign(eng, X, C, expect_flags=None):
    import torch

    Xd = eng.to_device(X)
    pack = eng.pack(eng.to_device(C, torch.float64))
    labels, mind2 = eng.assign(Xd, pack)
    lab_ref, d_ref = orc.pairwise_distances_argmin_min(X, C.astype(np.float64), metric_kwargs={"squared": True})
    got = labels.cpu().numpy()
    diff = got != lab_ref
    if diff.any():
        ties = _near_tie_mask(X, C)
        assert not (diff & ~ties).any(), f"{(diff & ~ties).sum()} labels differ outside near-ties ({diff.sum()} total)"
    dt = X.dtype.type
    npt.assert_allclose(mind2.cpu().numpy().astype(np.float64), d_ref, rtol=RTOL[dt] * 3, atol=RTOL[dt] * d_ref.max())
    return got, eng.flagged_count() if X.dtype == np.float32 else 0


@pytest.mark.parametrize("variant", [1, 2], ids=["ffma", "tcgen05"])
@pytest.mark.parametrize("n,d,k", [(5000, 64, 1024), (3001, 32, 256), (777, 20, 150), (1000, 128, 300),
                                    (513, 7, 33), (128, 4, 17), (4096, 16, 129), (1, 8, 40), (300, 36, 128),
                                    (40000, 64, 257), (129, 64, 600), (20000, 48, 2000), (2500, 13, 70),
                                    (3000, 100, 700), (2000, 128, 4100), (700, 66, 90)])
def test_assign_f32_tiled(eng, n, d, k, variant):
    """Both K1 float32 kernels (FP32 FFMA tiles; tcgen05 3xTF32 with TMEM accumulators, d <= 64)."""
    X, C = blobs(n, d, k, seed=n + d + k, dtype=np.float32)
    init = (C + 0.3 * np.random.RandomState(1).standard_normal(C.shape))
    eng.lib.b200km_set_option(1, variant)
    try:
        _check_assign(eng, X, init)
    finally:
        eng.lib.b200km_set_option(1, 0)


@pytest.mark.parametrize("n,d,k", [(3000, 64, 512), (2000, 32, 100)])
def test_assign_variants_agree_without_refine(eng, n, d, k):
    """The two float32 kernels must flag every row on which they could disagree: after removing the
    flagged rows of either kernel their raw (un-refined) labels are identical."""
    import torch

    X, C = blobs(n, d, k, seed=11, dtype=np.float32)
    C = C + 0.5 * np.random.RandomState(2).standard_normal(C.shape)   # split blobs -> many near ties
    Xd = eng.to_device(X)
    pack = eng.pack(eng.to_device(C, torch.float64))
    out = {}
    for variant in (1, 2):
        eng.lib.b200km_set_option(1, variant)
        try:
            labels = torch.empty(n, dtype=torch.int32, device=eng.device)
            flags = torch.zeros(n, dtype=torch.int32, device=eng.device)
            cnt = torch.zeros(1, dtype=torch.int32, device=eng.device)
            eng._call("b200km_assign_f32", Xd.data_ptr(), n, d, d, pack.buf.data_ptr(), k, labels.data_ptr(), None,
                      flags.data_ptr(), cnt.data_ptr(), 1, eng._stream())
            c = int(cnt.item())
            out[variant] = (labels.cpu().numpy(), set(flags[:c].cpu().numpy().tolist()))
        finally:
            eng.lib.b200km_set_option(1, 0)
    flagged = out[1][1] | out[2][1]
    keep = np.array([i not in flagged for i in range(n)])
    ref, _ = orc.pairwise_distances_argmin_min(X, C)
    npt.assert_array_equal(out[1][0][keep], ref[keep])
    npt.assert_array_equal(out[2][0][keep], ref[keep])


@pytest.mark.parametrize("n,d,k", [(4000, 16, 8), (999, 3, 2), (2000, 32, 16), (500, 200, 40), (700, 64, 5), (50, 2, 1)])
def test_assign_f32_rowblock_paths(eng, n, d, k):
    X, C = blobs(n, d, k, seed=7 * n + k, dtype=np.float32)
    _check_assign(eng, X, C + 0.1)


@pytest.mark.parametrize("n,d,k", [(4000, 16, 8), (1500, 5, 3), (2000, 40, 100), (333, 130, 20), (1, 1, 1)])
def test_assign_f64(eng, n, d, k):
    X, C = blobs(n, d, k, seed=n + k, dtype=np.float64)
    _check_assign(eng, X, C + 0.05)


def test_assign_f32_near_ties_are_refined(eng):
    """Centres that are float32-indistinguishable but float64-distinct, exact duplicates, and rows
    equidistant from two centres: the fp64 re-decision must give the float64 answer (lowest index
    on exact ties, as sklearn's ArgKmin: strict '<' while scanning)."""
    rs = np.random.RandomState(5)
    d, k = 24, 64
    C = rs.uniform(-10, 10, size=(k, d))
    C[10] = C[3]                      # exact duplicate -> index 3 must win
    C[20] = C[7] * (1 + 1e-9)         # differs only below float32 resolution
    X = np.vstack([C[3] + 0.01 * rs.standard_normal((200, d)), C[7] + 0.01 * rs.standard_normal((200, d)),
                   0.5 * (C[30] + C[31]) + 1e-4 * rs.standard_normal((200, d)),
                   rs.uniform(-10, 10, size=(600, d))]).astype(np.float32)
    got, nflag = _check_assign(eng, X, C)
    assert nflag >= 400                # the constructed ties were caught
    assert not (got[:200] == 10).any()  # duplicate resolved to the lower index
    lab_ref, _ = orc.pairwise_distances_argmin_min(X, C)
    # this case has no sklearn-level near ties beyond the duplicates: demand exact equality there
    ok = ~np.isin(lab_ref, [3, 10])
    npt.assert_array_equal(got[ok & ~_near_tie_mask(X, C, 1e-12)], lab_ref[ok & ~_near_tie_mask(X, C, 1e-12)])


def test_assign_strided_rows(eng):
    import torch

    X, C = blobs(1000, 24, 130, seed=3, dtype=np.float32)
    big = torch.zeros(1000, 40, dtype=torch.float32, device=eng.device)
    big[:, :24] = eng.to_device(X)
    pack = eng.pack(eng.to_device(C, torch.float64))
    lab, _ = eng.assign(big[:, :24], pack)
    lab_ref, _ = orc.pairwise_distances_argmin_min(X, C)
    npt.assert_array_equal(lab.cpu().numpy(), lab_ref)


@pytest.mark.parametrize("n,d,k,dtype", [(5000, 16, 8, np.float64), (6000, 32, 256, np.float32),
                                         (3000, 64, 1024, np.float32), (2047, 5, 3, np.float64),
                                         (1000, 128, 700, np.float32), (900, 33, 40, np.float64),
                                         (40000, 64, 1024, np.float32), (70001, 24, 3000, np.float64),
                                         (16384 * 3 + 5, 100, 1500, np.float32), (33000, 130, 8192, np.float32)])
def test_update_and_finalize(eng, n, d, k, dtype):
    import torch

    X, C = blobs(n, d, k, seed=k, dtype=dtype)
    labels = np.random.RandomState(0).randint(0, k, size=n).astype(np.int32)
    labels[labels == k - 1] = 0  # leave one cluster empty
    mind2 = np.random.RandomState(1).uniform(0, 9, size=n).astype(dtype)
    acc = eng.update(eng.to_device(X), eng.to_device(labels), eng.to_device(mind2), k)
    a = acc.cpu().numpy()
    sums = orc.centers_dense(X, labels, k)
    counts = np.bincount(labels, minlength=k)
    scale = np.abs(sums).max()
    npt.assert_allclose(a[: k * d].reshape(k, d), sums, rtol=RTOL[dtype], atol=RTOL[dtype] * scale)
    npt.assert_array_equal(a[k * d : k * d + k], counts)
    npt.assert_allclose(a[-1], mind2.astype(np.float64).sum(), rtol=1e-10)
    # sqrt flavour (non-squared inertia)
    acc2 = eng.update(eng.to_device(X), eng.to_device(labels), eng.to_device(mind2), k, sqrt=True)
    npt.assert_allclose(acc2[-1].item(), np.sqrt(mind2.astype(np.float64)).sum(), rtol=1e-10)
    old = eng.to_device(C, torch.float64)
    new, shift = eng.finalize(acc, k, d, old)
    ref_new = sums / np.maximum(counts, 1)[:, None]
    npt.assert_allclose(new.cpu().numpy(), ref_new, rtol=RTOL[dtype], atol=RTOL[dtype] * np.abs(ref_new).max())
    assert (new.cpu().numpy()[k - 1] == 0).all()  # empty cluster -> origin
    npt.assert_allclose(shift.item(), ((C - ref_new) ** 2).sum(), rtol=1e-6)


@pytest.mark.parametrize("n,d,k,dtype", [(6000, 32, 256, np.float32), (50000, 32, 200, np.float32),
                                         (3000, 64, 1024, np.float32), (2047, 5, 3, np.float64),
                                         (40000, 24, 40, np.float64), (33000, 130, 8192, np.float32),
                                         (20000, 16, 64, np.float32), (300, 8, 100, np.float32)])
def test_update_inertia(eng, n, d, k, dtype):
    """b200km_update_inertia_*: the sums and counts of b200km_update_* and, in the inertia slot, the float64 sum of
    ||x - c_label||^2 taken from the rows themselves (reference dask_ml/cluster/k_means.py:616 on the squared pass)."""
    import torch

    X, C = blobs(n, d, k, seed=k + 1, dtype=dtype)
    labels = np.random.RandomState(0).randint(0, k, size=n).astype(np.int32)
    if k > 100:
        labels[: n // 2] = 7          # one dominant cluster: the long-run path of the chunk-sort kernel
    C64 = (C + 1000.0).astype(np.float64) if d == 24 else C.astype(np.float64)   # data far from the centres too
    Xd, ld, cd = eng.to_device(X), eng.to_device(labels), eng.to_device(C64, torch.float64)
    acc = torch.zeros(eng.acc_len(k, d), dtype=torch.float64, device=eng.device)
    eng.update_inertia(Xd, ld, cd, k, acc)
    a = acc.cpu().numpy()
    ref = eng.update(Xd, ld, None, k).cpu().numpy()
    npt.assert_allclose(a[: k * d], ref[: k * d], rtol=1e-12, atol=1e-9 * np.abs(ref[: k * d]).max())
    npt.assert_array_equal(a[k * d : k * d + k], np.bincount(labels, minlength=k))
    want = ((X.astype(np.float64) - C64[labels]) ** 2).sum()
    npt.assert_allclose(a[k * d + k], want, rtol=1e-12)


@pytest.mark.parametrize("n,d,k", [(20000, 32, 256), (5001, 32, 200), (30000, 16, 64), (7000, 7, 33), (4000, 32, 17),
                                    (20000, 64, 1024), (6000, 128, 1500), (9000, 48, 300), (5000, 100, 257)])
def test_assign_lloyd_labels_match_full_assign(eng, n, d, k):
    """b200km_assign_lloyd_f32 (labels + near-tie flags only) gives the labels of b200km_assign_f32 after the
    float64 re-decision -- with both the pruned and the full re-decision -- on separated blobs and on one
    Gaussian cloud split between many centres (lots of near ties)."""
    import torch

    for seed, k_true in ((1, k), (2, 1)):
        X, _ = blobs(n, d, k_true, seed=seed, dtype=np.float32)
        C = X[:k].astype(np.float64) if k_true > 1 else np.random.RandomState(5).standard_normal((k, d))
        Xd = eng.to_device(X)
        pack = eng.pack(eng.to_device(C, torch.float64))
        want, d2 = eng.assign(Xd, pack)
        for pruned in (True, False):
            eng.pruned_refine = pruned
            try:
                scratch = torch.full((n,), float("nan"), dtype=torch.float32, device=eng.device)
                got, _ = eng.assign(Xd, pack, mind2=scratch, lloyd=True)
            finally:
                eng.pruned_refine = True
            assert int((got != want).sum().item()) == 0
        # and against the float64 oracle outside float64's own near ties
        lab_o, d2_o = orc.pairwise_distances_argmin_min(X.astype(np.float64), C, metric_kwargs={"squared": True})
        bad = want.cpu().numpy() != lab_o
        assert bad.sum() <= _near_tie_mask(X, C).sum() and not (bad & ~_near_tie_mask(X, C)).any()


@pytest.mark.parametrize("n,d,k,offset", [(5000, 64, 1024, 0.0), (3001, 32, 256, 0.0), (777, 7, 5, 0.0), (260, 64, 129, 0.0),
                                           (4096, 48, 300, 100.0), (10000, 16, 40, -1000.0), (300, 1, 3, 0.0),
                                           (2000, 61, 1000, 5.0), (3000, 128, 700, 0.0), (1001, 100, 130, 20.0),
                                           (500, 65, 64, 0.0)])
def test_cdist_on_tensor_cores(eng, n, d, k, offset):
    """K7's tcgen05 kernel (float32, d <= 128: b200km_cdist_pack_f32) against float64 numpy: squared distances,
    distances and the rbf kernel, incl. ragged n / k / d, strided rows and data far from the origin (rows and centres
    are centred on the mean of the centres, so the error scales with the spread of the data); and against the FFMA
    tile kernel it replaces for these shapes (the reference evaluates the float32 expanded form, atol 1e-4 in its
    own tests: reference tests/metrics/test_metrics.py:15-19, 42-71)."""
    import torch

    rs = np.random.RandomState(n + d + k)
    cen = rs.uniform(-10, 10, size=(max(2, k // 8), d))
    X = (cen[rs.randint(0, len(cen), n)] + rs.standard_normal((n, d)) + offset).astype(np.float32)
    Y = (cen[rs.randint(0, len(cen), k)] + 0.5 * rs.standard_normal((k, d)) + offset).astype(np.float32)
    D2 = sklearn.metrics.pairwise_distances(X.astype(np.float64), Y.astype(np.float64), metric="sqeuclidean")
    big = torch.zeros(n, d + 3, dtype=torch.float32, device=eng.device)    # ldx > d
    big[:, :d] = eng.to_device(X)
    Xd, Yd = big[:, :d], eng.to_device(Y)
    launches0 = eng.launches
    got2 = eng.cdist(Xd, Yd, squared=True).cpu().numpy()
    assert eng.launches - launches0 == 4          # three pack kernels + the tensor-core kernel: not the FFMA tile kernel
    # error model: fp32 accumulation of 3xTF32 products on CENTRED data -- a few 1e-7 of (||x'||^2 + ||c'||^2)
    mu = Y.astype(np.float64).mean(0)
    scale = ((X - mu) ** 2).sum(1)[:, None] + ((Y - mu) ** 2).sum(1)[None, :]
    assert (np.abs(got2 - D2) <= 4e-6 * scale + 1e-30).all()
    got = eng.cdist(Xd, Yd).cpu().numpy()
    # (the square root of a squared distance that is mostly rounding noise -- a row on top of a centre -- is off by
    #  up to the root of that noise, for the reference's float32 expanded form just the same)
    assert (np.abs(got - np.sqrt(D2)) <= 1e-4 * np.sqrt(D2) + np.sqrt(4e-6 * scale) + 1e-6).all()
    gamma = 1.0 / max(D2.mean(), 1e-9)
    npt.assert_allclose(eng.pairwise_kernel(Xd, Yd, "rbf", gamma=gamma).cpu().numpy(), np.exp(-gamma * D2), atol=2e-5)
    eng.cdist_on_tensor_cores = False
    try:
        ffma = eng.cdist(Xd, Yd, squared=True).cpu().numpy()
    finally:
        eng.cdist_on_tensor_cores = True
    npt.assert_allclose(got2, ffma, rtol=1e-3, atol=1e-4 * max(1.0, (offset * offset * d)))


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_mindist_many_candidates_goes_through_k1(eng, dtype):
    """From 48 new candidates on (a large oversampling factor) the running minimum is updated by ONE K1 pass against all
    of them + b200km_mindist_merge_* instead of m / 8 passes of K4: same minima, same cost."""
    import torch

    X, _ = blobs(30000, 20, 9, seed=5, dtype=dtype)
    Xd = eng.to_device(X)
    tol = 3e-5 if dtype == np.float32 else 1e-10
    mind2 = torch.empty(len(X), dtype=Xd.dtype, device=eng.device)
    eng.mindist_update(Xd, Xd[:1], mind2, init=True)
    cand = X[np.random.RandomState(1).choice(len(X), 300, replace=False)]
    launches0 = eng.launches
    cost = eng.mindist_update(Xd, eng.to_device(cand), mind2, init=False).item()
    assert eng.launches - launches0 < 20                       # not 300 / 8 passes of K4
    allc = np.vstack([X[:1], cand])
    d2 = sklearn.metrics.pairwise_distances(X.astype(np.float64), allc.astype(np.float64), metric="sqeuclidean").min(1)
    npt.assert_allclose(mind2.cpu().numpy(), d2, rtol=tol * 3, atol=tol * d2.max())
    npt.assert_allclose(cost, d2.sum(), rtol=tol)
    # init form (no old minima)
    cost0 = eng.mindist_update(Xd, eng.to_device(cand), mind2, init=True).item()
    d20 = sklearn.metrics.pairwise_distances(X.astype(np.float64), cand.astype(np.float64), metric="sqeuclidean").min(1)
    npt.assert_allclose(cost0, d20.sum(), rtol=tol)


def test_cdist_on_tensor_cores_long_second_operand(eng):
    """SpectralClustering's B block: 100 kept rows against many rows.  The long side becomes the streamed operand
    (roles swapped, the result handed back as the transposed view of what was written): same values."""
    rs = np.random.RandomState(0)
    A = rs.standard_normal((100, 20)).astype(np.float32)
    B = rs.standard_normal((70001, 20)).astype(np.float32)
    D2 = sklearn.metrics.pairwise_distances(A.astype(np.float64), B.astype(np.float64), metric="sqeuclidean")
    got = eng.pairwise_kernel(eng.to_device(A), eng.to_device(B), "rbf", gamma=0.05)
    assert tuple(got.shape) == (100, 70001)
    npt.assert_allclose(got.cpu().numpy(), np.exp(-0.05 * D2), atol=2e-5)
    npt.assert_allclose(eng.cdist(eng.to_device(A), eng.to_device(B), squared=True).cpu().numpy(), D2, rtol=1e-4, atol=1e-3)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("n,d", [(5000, 64), (3001, 3), (1000, 100), (64, 16), (2000, 33)])
def test_mindist_and_cost(eng, n, d, dtype):
    X, _ = blobs(n, d, 5, seed=n + d, dtype=dtype)
    Xd = eng.to_device(X)
    import torch

    mind2 = torch.empty(n, dtype=Xd.dtype, device=eng.device)
    cost = eng.mindist_update(Xd, Xd[:1], mind2, init=True).item()
    ref = orc.evaluate_cost(X, X[:1])
    tol = 3e-5 if dtype == np.float32 else 1e-10
    npt.assert_allclose(cost, ref, rtol=tol)
    cand = X[[5, 17, 40, 41, 42, 43, 44, 45, 46, 47, 48][: min(11, n - 1)]]
    cost2 = eng.mindist_update(Xd, eng.to_device(cand), mind2, init=False).item()
    allc = np.vstack([X[:1], cand])
    npt.assert_allclose(cost2, orc.evaluate_cost(X, allc), rtol=tol)
    d2 = sklearn.metrics.pairwise_distances(X.astype(np.float64), allc.astype(np.float64), metric="sqeuclidean").min(1)
    npt.assert_allclose(mind2.cpu().numpy(), d2, rtol=tol * 3, atol=tol * d2.max())


@pytest.mark.parametrize("n,d", [(200001, 64), (160000, 32), (180017, 128), (155000, 256), (170000, 48), (160000, 512)])
def test_mindist_tma_ring_equals_register_path(eng, n, d):
    """K4 fed by the TMA ring (B200KM_OPT_K4_VARIANT = 2: wherever it fits, from 1024 rows per SM on) against the
    register-path kernel (= 1): the same per-row arithmetic, so the running minima are equal bit for bit; the
    float64 cost differs only in the order of the additions.  One or two new candidates (more take the row kernel
    whatever the size), first pass and update."""
    import torch

    from dask_ml_b200 import _lib

    rs = np.random.RandomState(n + d)
    X = (rs.standard_normal((n, d)) * 3 + rs.uniform(-10, 10, size=(1, d))).astype(np.float32)
    Xd = eng.to_device(X)
    out = {}
    try:
        for variant in (1, 2):
            assert eng.lib.b200km_set_option(_lib.OPT_K4_VARIANT, variant) == 0
            mind2 = torch.empty(n, dtype=torch.float32, device=eng.device)
            costs = [eng.mindist_update(Xd, Xd[:1], mind2, init=True).item()]
            first = 1
            for m in (2, 1, 2, 1, 2, 2):
                cand = Xd[first * 37 : first * 37 + m].contiguous()
                costs.append(eng.mindist_update(Xd, cand, mind2, init=False).item())
                first += m
            out[variant] = (mind2.cpu().numpy(), np.array(costs))
    finally:
        eng.lib.b200km_set_option(_lib.OPT_K4_VARIANT, 0)
    npt.assert_array_equal(out[2][0], out[1][0])
    npt.assert_allclose(out[2][1], out[1][1], rtol=1e-13)
    picked = np.concatenate([[0]] + [np.arange(f * 37, f * 37 + m) for f, m in
                                     zip(np.cumsum([1, 2, 1, 2, 1, 2]), (2, 1, 2, 1, 2, 2))])
    d2 = sklearn.metrics.pairwise_distances(X[:20000].astype(np.float64), X[picked].astype(np.float64),
                                            metric="sqeuclidean").min(1)
    npt.assert_allclose(out[2][0][:20000], d2, rtol=1e-4, atol=3e-5 * d2.max())


@pytest.mark.parametrize("n,d,m,pad", [(50000, 64, 5, 0), (20001, 64, 8, 0), (30000, 32, 20, 0), (30000, 32, 25, 0), (9000, 128, 128, 0),
                                       (7000, 7, 47, 0), (12345, 100, 33, 4), (300, 64, 6, 0), (4000, 16, 129, 0)])
def test_mindist_many_candidates_on_tensor_cores(eng, n, d, m, pad):
    """K4 against 25 .. 128 new candidates in one tensor-core pass (b200km_mindist_tc_f32: K7's kernel with a min-merge
    epilogue) against float64 numpy, first pass and update, far from the origin; fewer / more candidates take the other
    routes and meet the same bounds."""
    import torch

    rs = np.random.RandomState(n + m)
    C = rs.uniform(-10, 10, size=(12, d)) + 300.0
    X = (C[rs.randint(0, 12, size=n)] + rs.standard_normal((n, d))).astype(np.float32)
    full = eng.to_device(np.concatenate([X, np.zeros((n, pad), np.float32)], axis=1)) if pad else eng.to_device(X)
    Xd = full[:, :d]
    cand_rows = rs.choice(n, size=m + 3, replace=False)
    mind2 = torch.empty(n, dtype=torch.float32, device=eng.device)
    launches0 = eng.launches
    cost0 = eng.mindist_update(Xd, Xd[cand_rows[:3]].contiguous(), mind2, init=True).item()
    eng.trace = []
    cost1 = eng.mindist_update(Xd, Xd[cand_rows[3:]].contiguous(), mind2, init=False).item()
    names = [nm for nm, _ in eng.trace]
    eng.trace = None
    assert ("b200km_mindist_tc_f32" in names) == (25 <= m <= 128)
    X64 = X.astype(np.float64)
    d2a = sklearn.metrics.pairwise_distances(X64, X64[cand_rows[:3]], metric="sqeuclidean").min(1)
    d2b = np.minimum(d2a, sklearn.metrics.pairwise_distances(X64, X64[cand_rows[3:]], metric="sqeuclidean").min(1))
    npt.assert_allclose(cost0, d2a.sum(), rtol=3e-5)
    npt.assert_allclose(cost1, d2b.sum(), rtol=3e-5)
    # centred expanded form: absolute error ~2^-21 of the spread squared (NOT of the offset 300)
    spread2 = ((X64 - X64.mean(0)) ** 2).sum(1).max()
    npt.assert_allclose(mind2.cpu().numpy(), d2b, rtol=1e-4, atol=4e-6 * spread2)
    # the candidates themselves: distance (numerically) zero
    assert mind2.cpu().numpy()[cand_rows].max() <= 4e-6 * spread2


@pytest.mark.parametrize("n,d,pad", [(200001, 64, 0), (100, 64, 0), (33, 32, 0), (50017, 128, 0), (70000, 32, 4), (9999, 64, 8)])
def test_mindist_row_kernel(eng, n, d, pad):
    """K4 with 3 .. 8 new candidates: one thread per row behind warp-private cp.async buffers.  Against float64 numpy
    (direct-form float32 accuracy), and: the result of a row does not depend on where the row sits -- the same rows
    handed over as a shifted slice give the same minima bit for bit (what makes the picks independent of the sharding)."""
    import torch

    rs = np.random.RandomState(n + d)
    X = (rs.standard_normal((n, d)) * 3 + rs.uniform(-10, 10, size=(1, d))).astype(np.float32)
    full = eng.to_device(np.concatenate([X, np.zeros((n, pad), np.float32)], axis=1)) if pad else eng.to_device(X)
    Xd = full[:, :d]
    mind2 = torch.empty(n, dtype=torch.float32, device=eng.device)
    eng.mindist_update(Xd, Xd[:1].contiguous(), mind2, init=True)
    cands = []
    first = 1
    eng.trace = []
    for m in (3, 8, 5, 4, 7, 6):
        c = np.arange(first * 3, first * 3 + m) % n
        cands.append(c)
        eng.mindist_update(Xd, Xd[torch.as_tensor(c, device=eng.device)].contiguous(), mind2, init=False)
        first += m
    names = [nm for nm, _ in eng.trace]
    eng.trace = None
    assert names.count("b200km_mindist_update_f32") == 6
    allc = np.concatenate([[0]] + cands)
    X64 = X.astype(np.float64)
    d2 = sklearn.metrics.pairwise_distances(X64[:20000], X64[allc], metric="sqeuclidean").min(1)
    got = mind2.cpu().numpy()
    npt.assert_allclose(got[:20000], d2, rtol=2e-5, atol=1e-5 * d2.max())
    if n > 100:
        off = 37                                  # a slice that starts mid-warp: other lanes, other CTAs, same answers
        m2 = torch.empty(n - off, dtype=torch.float32, device=eng.device)
        eng.mindist_update(Xd[off:], Xd[:1].contiguous(), m2, init=True)
        for c in cands:
            eng.mindist_update(Xd[off:], Xd[torch.as_tensor(c, device=eng.device)].contiguous(), m2, init=False)
        npt.assert_array_equal(m2.cpu().numpy(), got[off:])


def test_bernoulli_matches_oracle_philox(eng):
    import torch

    rs = np.random.RandomState(0)
    n = 200000
    mind2 = rs.uniform(0, 4, size=n).astype(np.float32)
    cost = float(mind2.astype(np.float64).sum())
    md = eng.to_device(mind2)
    for seed, rnd, off, l in [(0, 0, 0, 2), (123456789012345, 3, 10**10, 50), (7, 11, 12345, 1000)]:
        got = np.sort(eng.bernoulli(md, l / cost, seed, rnd, off).cpu().numpy())
        u = orc.philox_uniform_f64(seed, rnd, off, n)
        p = (l / cost) * mind2.astype(np.float64)
        ref = off + np.where(p > u)[0]
        npt.assert_array_equal(got, ref)
    # injected draws
    u = rs.uniform(size=n)
    got = np.sort(eng.bernoulli(md, 30 / cost, 0, 0, 5, draws=eng.to_device(u)).cpu().numpy())
    npt.assert_array_equal(got, 5 + np.where((30 / cost) * mind2.astype(np.float64) > u)[0])
    # capacity overflow is retried transparently
    got = eng.bernoulli(md, 1e9, 1, 0, 0, capacity=16)
    assert got.numel() == (mind2 > 0).sum()


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_gather_cdist_norms_finite(eng, dtype):
    import torch

    rs = np.random.RandomState(2)
    X = rs.standard_normal((1000, 37)).astype(dtype)
    Y = rs.standard_normal((70, 37)).astype(dtype)
    Xd, Yd = eng.to_device(X), eng.to_device(Y)
    idx = np.array([999, 0, 5, 5, 321], dtype=np.int64)
    npt.assert_array_equal(eng.gather(Xd, torch.from_numpy(idx)).cpu().numpy(), X[idx])
    tol = 2e-4 if dtype == np.float32 else 1e-10
    npt.assert_allclose(eng.cdist(Xd, Yd).cpu().numpy(), orc.euclidean_distances(X, Y), rtol=tol, atol=tol)
    npt.assert_allclose(eng.cdist(Xd, Yd, squared=True).cpu().numpy(), orc.euclidean_distances(X, Y, squared=True),
                        rtol=tol, atol=tol)
    npt.assert_allclose(eng.row_norms(Xd, squared=True).cpu().numpy(), orc.row_norms(X, squared=True), rtol=tol)
    npt.assert_allclose(eng.row_norms(Xd).cpu().numpy(), orc.row_norms(X), rtol=tol)
    assert not eng.has_nonfinite(Xd)
    for bad in (np.nan, np.inf, -np.inf):
        Z = X.copy()
        Z[777, 36] = bad
        assert eng.has_nonfinite(eng.to_device(Z))


def test_make_blobs_statistics(eng):
    import torch

    k, d, n = 16, 8, 400000
    centers = torch.from_numpy(np.random.RandomState(0).uniform(-10, 10, size=(k, d)).astype(np.float32))
    X = eng.make_blobs(n, d, centers, 1.0, seed=3)
    pack = eng.pack(centers.double().to(eng.device))
    lab, d2 = eng.assign(X, pack)
    cnt = np.bincount(lab.cpu().numpy(), minlength=k)
    assert abs(cnt - n / k).max() < 6 * np.sqrt(n / k)
    assert abs(d2.double().mean().item() - d) < 0.05  # chi^2_d mean
    # sharding invariance: rows generated with an offset equal the same rows of the full array
    Y = eng.make_blobs(1000, d, centers, 1.0, seed=3, row_offset=5000)
    assert torch.equal(Y, X[5000:6000])


@pytest.mark.parametrize("n,d,k,dtype", [(4000, 16, 8, np.float64), (999, 3, 2, np.float32), (70000, 32, 16, np.float32),
                                         (1, 1, 1, np.float64), (257, 5, 3, np.float64), (100000, 16, 8, np.float32)])
def test_assign_update_fused(eng, n, d, k, dtype):
    """K1+K2 in one pass (HBM-bound regime, k <= 16, d <= 32) against the oracle's two passes."""
    import torch

    X, C = blobs(n, d, k, seed=n + k, dtype=dtype)
    C = C + 0.2
    Xd = eng.to_device(X)
    Cd = eng.to_device(C, torch.float64)
    labels = torch.full((n,), -1, dtype=torch.int32, device=eng.device)
    mind2 = torch.zeros(n, dtype=Xd.dtype, device=eng.device)
    acc = torch.zeros(eng.acc_len(k, d), dtype=torch.float64, device=eng.device)
    assert eng.can_fuse(k, d)
    eng.assign_update(Xd, Cd, labels, mind2, acc)
    lab_ref, d_ref = orc.pairwise_distances_argmin_min(X, C, metric_kwargs={"squared": True})
    got = labels.cpu().numpy()
    diff = got != lab_ref
    assert not (diff & ~_near_tie_mask(X, C, 1e-12)).any()
    tol = RTOL[dtype]
    npt.assert_allclose(mind2.cpu().numpy().astype(np.float64), d_ref, rtol=3 * tol, atol=tol * max(d_ref.max(), 1e-30))
    a = acc.cpu().numpy()
    sums = orc.centers_dense(X, got.astype(np.int32), k)
    npt.assert_allclose(a[: k * d].reshape(k, d), sums, rtol=1e-12, atol=1e-9 * max(np.abs(sums).max(), 1.0))
    npt.assert_array_equal(a[k * d : k * d + k], np.bincount(got, minlength=k))
    npt.assert_allclose(a[-1], mind2.double().sum().item(), rtol=1e-6 if dtype == np.float32 else 1e-12)
    with pytest.raises(ValueError):
        eng.assign_update(eng.to_device(np.zeros((4, 40), dtype)), eng.to_device(np.zeros((2, 40)), torch.float64),
                          labels[:4], mind2[:4], acc)


@pytest.mark.parametrize("n,d,k,dtype,pad", [(100001, 16, 8, np.float64, 0), (5000, 16, 16, np.float64, 0),
                                             (70001, 8, 5, np.float64, 0), (33333, 8, 13, np.float64, 2),
                                             (90000, 16, 3, np.float32, 0), (1234, 16, 16, np.float32, 4),
                                             (300001, 32, 8, np.float32, 0), (65537, 32, 16, np.float32, 8),
                                             (255, 16, 8, np.float64, 0), (1, 16, 1, np.float64, 0)])
def test_fused_kernels_agree(eng, n, d, k, dtype, pad):
    """The cp.async fused kernel (rows of 64 / 128 bytes, 16-byte aligned) against the sorted-tile one
    (B200KM_OPT_FUSED_VARIANT = 1): same labels (up to exact near-ties), distances to the last bits, same counts;
    the float64 sums differ only in the order of the additions.  ``pad`` > 0: rows with a leading dimension of d + pad."""
    import torch

    from dask_ml_b200 import _lib

    X, C = blobs(n, d, k, seed=3 * n + k, dtype=dtype)
    C = C + 0.1
    full = eng.to_device(np.concatenate([X, np.zeros((n, pad), dtype)], axis=1)) if pad else eng.to_device(X)
    Xd = full[:, :d]
    assert Xd.stride(0) == d + pad
    Cd = eng.to_device(C, torch.float64)
    out = {}
    try:
        for variant in (1, 0):
            assert eng.lib.b200km_set_option(_lib.OPT_FUSED_VARIANT, variant) == 0
            labels = torch.full((n,), -1, dtype=torch.int32, device=eng.device)
            mind2 = torch.zeros(n, dtype=Xd.dtype, device=eng.device)
            acc = torch.zeros(eng.acc_len(k, d), dtype=torch.float64, device=eng.device)
            eng.assign_update(Xd, Cd, labels, mind2, acc)
            out[variant] = (labels.cpu().numpy(), mind2.cpu().numpy(), acc.cpu().numpy())
    finally:
        eng.lib.b200km_set_option(_lib.OPT_FUSED_VARIANT, 0)
    # (the cp.async kernel decides on centred expanded-form float64 scores from the FP64 tensor cores and rebuilds the
    #  winner's distance in direct form; the sorted-tile kernel is direct form throughout: exact near-ties may differ)
    same = out[0][0] == out[1][0]
    assert not (~same & ~_near_tie_mask(X, C, 1e-12)).any()
    npt.assert_allclose(out[0][1][same], out[1][1][same], rtol=1e-13 if dtype == np.float64 else 2e-7)
    a0, a1 = out[0][2], out[1][2]
    if same.all():
        npt.assert_array_equal(a0[k * d : k * d + k], a1[k * d : k * d + k])
        npt.assert_allclose(a0[: k * d], a1[: k * d], rtol=1e-13, atol=1e-11 * max(np.abs(a1[: k * d]).max(), 1.0))
        npt.assert_allclose(a0[-1], a1[-1], rtol=1e-13)
    npt.assert_array_equal(a0[k * d : k * d + k], np.bincount(out[0][0], minlength=k))
    lab_ref, _ = orc.pairwise_distances_argmin_min(X, C, metric_kwargs={"squared": True})
    assert not ((out[0][0] != lab_ref) & ~_near_tie_mask(X, C, 1e-12)).any()
    sums = orc.centers_dense(X, out[0][0].astype(np.int32), k)
    npt.assert_allclose(a0[: k * d].reshape(k, d), sums, rtol=1e-12, atol=1e-9 * max(np.abs(sums).max(), 1.0))


def test_fused_kernel_random_shapes_fuzz(eng):
    """The cp.async / FP64-tensor-core fused kernel on random shapes of its domain: every row width it takes, k from 1
    to 16, sizes around the tile and warp boundaries, data far from the origin, duplicated centres (lowest index must win
    an exact tie) -- labels against float64 direct-form distances, sums / counts / inertia against numpy."""
    import torch

    rs = np.random.RandomState(1234)
    for trial in range(40):
        dtype = [np.float32, np.float64][trial & 1]
        d = int(rs.choice([16, 32] if dtype == np.float32 else [8, 16]))
        k = int(rs.randint(1, 17))
        n = int(rs.choice([1, 31, 32, 33, 255, 256, 257, 511, 1000, 4097, 20001, 66000]))
        offset = float(rs.choice([0.0, 0.0, 100.0, -3e4]))
        scale = float(rs.choice([1.0, 1e-3, 50.0]))
        C = rs.uniform(-10, 10, size=(k, d)) * scale + offset
        if k >= 4:
            C[k - 1] = C[1]                       # an exact duplicate: index 1 must win every tie with index k - 1
        X = (C[rs.randint(0, k, size=n)] + rs.standard_normal((n, d)) * scale).astype(dtype)
        Xd, Cd = eng.to_device(X), eng.to_device(C, torch.float64)
        labels = torch.full((n,), -7, dtype=torch.int32, device=eng.device)
        mind2 = torch.zeros(n, dtype=Xd.dtype, device=eng.device)
        acc = torch.zeros(eng.acc_len(k, d), dtype=torch.float64, device=eng.device)
        eng.assign_update(Xd, Cd, labels, mind2, acc)
        got = labels.cpu().numpy()
        D = ((X.astype(np.float64)[:, None, :] - C[None, :, :]) ** 2).sum(-1)
        ref = D.argmin(1)
        best = D[np.arange(n), ref]
        mine = D[np.arange(n), got]
        # a different label only where the two distances agree to rounding (relative to the centred magnitudes)
        tol = 1e-11 * (np.abs(X.astype(np.float64) - C.mean(0)).max() ** 2 * d + 1e-300)
        assert (np.abs(mine - best) <= tol).all(), (trial, dtype, n, d, k, offset, scale)
        if k >= 4:
            assert not (got == k - 1).any(), (trial, "duplicate centre won a tie")
        npt.assert_allclose(mind2.cpu().numpy().astype(np.float64), mine, rtol=1e-12 if dtype == np.float64 else 2e-7,
                            atol=1e-300)
        a = acc.cpu().numpy()
        sums = np.zeros((k, d))
        np.add.at(sums, got, X.astype(np.float64))
        npt.assert_allclose(a[: k * d].reshape(k, d), sums, rtol=1e-12, atol=1e-12 * max(np.abs(sums).max(), 1e-300))
        npt.assert_array_equal(a[k * d : k * d + k], np.bincount(got, minlength=k))
        npt.assert_allclose(a[-1], mind2.double().sum().item(), rtol=1e-6 if dtype == np.float32 else 1e-12)


@pytest.mark.parametrize("dtype,d,k", [(np.float64, 16, 8), (np.float32, 32, 16), (np.float64, 5, 3)])
def test_lloyd_small_one_launch_matches_separate_steps(eng, dtype, d, k):
    """b200km_lloyd_small_* (one launch: assign + sums + divide + shift, result through page-locked memory) against
    assign_update + finalize, on the shapes of both fused kernels; acc and the ticket come back zeroed."""
    import torch

    n = 50001
    X, C = blobs(n, d, k, seed=11 + d, dtype=dtype)
    Xd = eng.to_device(X)
    Cd = eng.to_device(C + 0.3, torch.float64)
    labels = torch.empty(n, dtype=torch.int32, device=eng.device)
    mind2 = torch.empty(n, dtype=Xd.dtype, device=eng.device)
    acc = torch.zeros(eng.acc_len(k, d) + 1, dtype=torch.float64, device=eng.device)
    ticket = torch.zeros(1, dtype=torch.int32, device=eng.device)
    new = torch.empty_like(Cd)
    shift, inertia = eng.lloyd_small(Xd, Cd, labels, mind2, acc, ticket, new)
    assert int(ticket.item()) == 0 and float(acc.abs().max().item()) == 0.0
    l2 = torch.empty_like(labels)
    m2 = torch.empty_like(mind2)
    acc2 = torch.zeros(eng.acc_len(k, d), dtype=torch.float64, device=eng.device)
    eng.assign_update(Xd, Cd, l2, m2, acc2)
    inertia2 = float(acc2[-1].item())
    new2, shift2 = eng.finalize(acc2, k, d, Cd)
    assert torch.equal(labels, l2) and torch.equal(mind2, m2)
    npt.assert_allclose(new.cpu().numpy(), new2.cpu().numpy(), rtol=1e-13, atol=1e-13)
    npt.assert_allclose(shift, float(shift2.item()), rtol=1e-10)
    npt.assert_allclose(inertia, inertia2, rtol=1e-13)


def test_datasets_make_blobs_api(eng):
    """dask_ml_b200.datasets.make_blobs: reference signature (dask_ml/datasets.py:78-205), device output."""
    from dask_ml_b200.datasets import make_blobs

    X, y = make_blobs(n_samples=60000, n_features=5, centers=4, cluster_std=0.5, random_state=0, chunks=1000)
    assert X.shape == (60000, 5) and X.dtype == np.float64 and y.shape == (60000,) and y.dtype == np.int64
    Xh, yh = X.compute(), y.compute()
    assert set(np.unique(yh)) == {0, 1, 2, 3}
    c = np.random.RandomState(0).uniform(-10, 10, size=(4, 5))      # the centres sklearn would draw first
    for j in range(4):
        npt.assert_allclose(Xh[yh == j].mean(0), c[j], atol=0.02)
        assert abs((Xh[yh == j] - c[j]).std() - 0.5) < 0.01
    X2, y2 = make_blobs(n_samples=100, n_features=5, centers=4, cluster_std=0.5, random_state=0, row_offset=500)
    npt.assert_array_equal(X2.compute(), Xh[500:600])                # shard invariance
    npt.assert_array_equal(y2.compute(), yh[500:600])
    Xc, _ = make_blobs(n_samples=10, centers=np.array([[-7.0, -7.0], [0.0, 0.0], [7.0, 7.0]]), cluster_std=0.1,
                       random_state=0, dtype="f4")
    assert Xc.shape == (10, 2) and Xc.dtype == np.float32


def test_assign_random_shapes_fuzz(eng):
    """Many random (n, d, k) shapes through every K1 path (tcgen05, FFMA, row-block fp64) against the oracle."""
    import torch

    rs = np.random.RandomState(123)
    for trial in range(36):
        d = int(rs.choice([1, 2, 3, 5, 8, 12, 16, 24, 31, 40, 64, 65, 96, 128, 130, 200]))
        k = int(rs.choice([1, 2, 7, 16, 17, 40, 100, 255, 256, 257, 513, 1000]))
        n = int(rs.choice([1, 5, 127, 128, 129, 255, 256, 257, 1000, 4097]))
        dtype = np.float32 if trial % 3 else np.float64
        X, C = blobs(n, d, k, seed=1000 + trial, dtype=dtype)
        C = C + 0.3 * rs.standard_normal(C.shape)
        variant = 1 + trial % 3 if dtype == np.float32 else 0
        eng.lib.b200km_set_option(1, variant)
        try:
            _check_assign(eng, X, C)
        finally:
            eng.lib.b200km_set_option(1, 0)


@pytest.mark.parametrize("n,d,k", [(150001, 32, 256), (40000, 32, 200), (77777, 16, 100), (60000, 7, 17),
                                    (130, 24, 33), (100000, 13, 256), (38017, 32, 64), (1, 4, 20),
                                    (300000, 8, 255)])
def test_assign_f32_small_k_kernel(eng, n, d, k):
    """The small-k tcgen05 kernel (auto variant, k <= 256 and d <= 32: resident centre image, two row groups)
    against the oracle, and against the general tcgen05 kernel on the same rows (several steps per CTA and group,
    ragged tails, d not a multiple of 4)."""
    import torch

    X, C = blobs(n, d, k, seed=7 * n + d + k, dtype=np.float32)
    init = (C + 0.3 * np.random.RandomState(2).standard_normal(C.shape))
    assert eng.lib.b200km_get_option(1) == 0
    got, _ = _check_assign(eng, X, init)
    eng.lib.b200km_set_option(1, 2)
    try:
        Xd = eng.to_device(X)
        pack = eng.pack(eng.to_device(init, torch.float64))
        lab2, m2 = eng.assign(Xd, pack)
    finally:
        eng.lib.b200km_set_option(1, 0)
    npt.assert_array_equal(got, lab2.cpu().numpy())


def test_assign_f32_small_k_strided(eng):
    """Row stride > d (a column slice of a wider matrix) through the small-k kernel's scalar-load path."""
    import torch

    rs = np.random.RandomState(5)
    wide = rs.standard_normal((50000, 37)).astype(np.float32) * 3
    Xd = eng.to_device(wide)[:, 3:30]                    # d = 27, ldx = 37, base pointer not 16-byte aligned
    C = wide[:150, 3:30].astype(np.float64) + 0.1
    pack = eng.pack(eng.to_device(C, torch.float64))
    labels, mind2 = eng.assign(Xd, pack)
    lab_ref, d_ref = orc.pairwise_distances_argmin_min(wide[:, 3:30], C, metric_kwargs={"squared": True})
    got = labels.cpu().numpy()
    diff = got != lab_ref
    assert not (diff & ~_near_tie_mask(wide[:, 3:30], C)).any()
    npt.assert_allclose(mind2.cpu().numpy().astype(np.float64), d_ref, rtol=3e-5, atol=1e-5 * d_ref.max())


@pytest.mark.parametrize("variant", [0, 1, 2], ids=["auto", "ffma", "tcgen05"])
@pytest.mark.parametrize("n,d,k,offset", [(20000, 32, 64, 1000.0), (20000, 64, 300, 300.0)])
def test_assign_f32_far_from_origin(eng, n, d, k, offset, variant):
    """Data far from the origin: the float32 kernels work on x - mu and c - mu (mu = mean of the centres), so the
    labels stay exact WITHOUT sending most rows through the float64 re-decision, and the winner distance keeps
    float32 accuracy."""
    import torch

    rs = np.random.RandomState(11)
    cen = rs.uniform(-10, 10, size=(k, d))
    X = (cen[rs.randint(0, k, n)] + rs.standard_normal((n, d)) + offset).astype(np.float32)
    C = X[:k].astype(np.float64)
    X64 = X.astype(np.float64)
    D = (X64 ** 2).sum(1)[:, None] - 2.0 * X64 @ C.T + (C ** 2).sum(1)[None, :]
    lab_ref = D.argmin(1)
    m_ref = ((X64 - C[lab_ref]) ** 2).sum(1)
    eng.lib.b200km_set_option(1, variant)
    try:
        pack = eng.pack(eng.to_device(C, torch.float64))
        labels, mind2 = eng.assign(eng.to_device(X), pack)
        flagged = eng.flagged_count()
    finally:
        eng.lib.b200km_set_option(1, 0)
    assert flagged < 0.02 * n, f"{flagged} of {n} rows flagged: the rounding bound follows the offset"
    got = labels.cpu().numpy()
    diff = got != lab_ref
    assert not (diff & ~_near_tie_mask(X, C, rel=1e-9)).any()
    m = mind2.cpu().numpy().astype(np.float64)
    nz = m_ref > 0
    assert (np.abs(m[nz] - m_ref[nz]) / m_ref[nz]).max() < 2e-6


def _assign_both_refines(eng, X, C):
    import torch

    out = []
    Xd = eng.to_device(X)
    pack = eng.pack(eng.to_device(C, torch.float64))
    for pruned in (True, False):
        eng.pruned_refine = pruned
        try:
            labels, mind2 = eng.assign(Xd, pack)
            out.append((labels.cpu().numpy(), mind2.cpu().numpy(), eng.flagged_count()))
        finally:
            eng.pruned_refine = True
    return out


@pytest.mark.parametrize("n,d,k", [(60000, 64, 1024), (50000, 32, 256), (30000, 128, 700), (20000, 20, 90),
                                    (5000, 64, 300), (4096, 8, 40), (20000, 256, 600), (9000, 333, 100),
                                    (6000, 1024, 300), (30000, 136, 1024)])
def test_refine_pruned_equals_full_scan(eng, n, d, k):
    """The pruned float64 re-decision (tensor-core candidate pass + float64 on the listed centres) gives the labels
    and distances of the full float64 scan bit for bit; centres = data rows, so near-ties are plentiful."""
    X, _ = blobs(n, d, max(4, k // 8), seed=n + k, dtype=np.float32)
    C = X[:k].astype(np.float64)
    (lp, mp, fp), (lf, mf, ff) = _assign_both_refines(eng, X, C)
    assert fp == ff and fp > 0
    npt.assert_array_equal(lp, lf)
    npt.assert_array_equal(mp, mf)
    lab_ref, _ = orc.pairwise_distances_argmin_min(X, C, metric_kwargs={"squared": True})
    diff = lp != lab_ref
    assert not (diff & ~_near_tie_mask(X, C)).any()


def test_refine_pruned_duplicate_centres_overflow(eng):
    """More than six centres tie for a row (duplicates, as when several empty clusters are parked at the origin):
    those rows overflow the candidate list and take the full scan; the lowest index must win."""
    rs = np.random.RandomState(3)
    n, d, k = 30000, 16, 64
    X = rs.standard_normal((n, d)).astype(np.float32)
    C = rs.standard_normal((k, d)) * 3
    C[5:17] = 0.0                                           # twelve centres at the origin
    (lp, mp, fp), (lf, mf, ff) = _assign_both_refines(eng, X, C)
    npt.assert_array_equal(lp, lf)
    npt.assert_array_equal(mp, mf)
    assert not np.isin(lp, np.arange(6, 17)).any()          # of the duplicates only index 5 can be a label
    assert (lp == 5).sum() > 100


@pytest.mark.parametrize("scratch", [1, 4])
def test_refine_pruned_long_candidate_lists(eng, scratch):
    """Heavily overlapping clusters: every blob is shared out between 16 centres that differ by 1e-3 sigma, so a flagged
    row lists ~16 columns.  With the n-entry flag buffer of the C ABI's base form (scratch = 1) the lists hold what the
    buffer allows for the number of rows flagged and longer ones overflow to the full scan; with the engine's 4 n
    entries they hold up to 32.  Either way: the labels and distances of the full float64 scan, bit for bit."""
    rs = np.random.RandomState(11)
    n, d, k = 40000, 64, 512
    blobs_c = rs.uniform(-10, 10, size=(k // 16, d))
    X = (blobs_c[rs.randint(0, k // 16, n)] + rs.standard_normal((n, d))).astype(np.float32)
    C = np.repeat(blobs_c, 16, axis=0) + 1e-3 * rs.standard_normal((k, d))
    old = eng.FLAG_SCRATCH
    eng.FLAG_SCRATCH = scratch
    try:
        (lp, mp, fp), (lf, mf, ff) = _assign_both_refines(eng, X, C)
    finally:
        eng.FLAG_SCRATCH = old
    assert fp == ff and fp > 0.5 * n
    npt.assert_array_equal(lp, lf)
    npt.assert_array_equal(mp, mf)
    lab_ref, _ = orc.pairwise_distances_argmin_min(X, C, metric_kwargs={"squared": True})
    assert not ((lp != lab_ref) & ~_near_tie_mask(X, C)).any()


def test_refine_pruned_falls_back_when_most_rows_are_flagged(eng):
    """All centres equal: every row is flagged, more than the scratch layout holds, so the device-side gate runs the
    full scan instead (and the pruned kernels return at once)."""
    rs = np.random.RandomState(4)
    n, d, k = 20000, 32, 24
    X = rs.standard_normal((n, d)).astype(np.float32)
    C = np.tile(rs.standard_normal((1, d)), (k, 1))
    (lp, mp, fp), (lf, mf, ff) = _assign_both_refines(eng, X, C)
    assert fp == n
    npt.assert_array_equal(lp, lf)
    assert (lp == 0).all()
    npt.assert_array_equal(mp, mf)


@pytest.mark.parametrize("n,d,k", [(3000, 129, 300), (5000, 192, 1024), (4097, 256, 513), (2000, 333, 100),
                                    (1500, 500, 257), (1000, 1024, 64), (600, 2048, 40), (129, 160, 17),
                                    (40000, 136, 256)])
def test_assign_f32_wide_rows_kernel(eng, n, d, k):
    """Rows wider than 128 features: the tcgen05 kernel that cuts K into 64-feature panels (auto variant), against
    the oracle and against the float64 row-block kernel it replaced for these shapes (variant 1)."""
    import torch

    X, C = blobs(n, d, max(2, k // 4), seed=3 * n + d + k, dtype=np.float32)
    init = X[np.random.RandomState(7).randint(0, n, k)].astype(np.float64) + \
        0.2 * np.random.RandomState(8).standard_normal((k, d))
    got, flagged = _check_assign(eng, X, init)
    eng.lib.b200km_set_option(1, 1)
    try:
        pack = eng.pack(eng.to_device(init, torch.float64))
        lab1, _ = eng.assign(eng.to_device(X), pack)
    finally:
        eng.lib.b200km_set_option(1, 0)
    diff = got != lab1.cpu().numpy()
    assert not (diff & ~_near_tie_mask(X, init)).any()


def test_assign_f32_wide_rows_strided_and_far(eng):
    """Wide rows taken as a column slice (row stride > d, unaligned base) of data far from the origin."""
    import torch

    rs = np.random.RandomState(9)
    wide = (rs.standard_normal((6000, 210)) + 400.0).astype(np.float32)
    Xd = eng.to_device(wide)[:, 5:204]                  # d = 199, ldx = 210
    Xh = wide[:, 5:204]
    C = Xh[:90].astype(np.float64) + 0.3
    pack = eng.pack(eng.to_device(C, torch.float64))
    labels, mind2 = eng.assign(Xd, pack)
    X64 = Xh.astype(np.float64)
    D = np.stack([((X64 - C[j]) ** 2).sum(1) for j in range(90)], axis=1)
    lab_ref = D.argmin(1)
    got = labels.cpu().numpy()
    diff = got != lab_ref
    assert not (diff & ~_near_tie_mask(Xh, C, rel=1e-9)).any()
    npt.assert_allclose(mind2.cpu().numpy(), D[np.arange(6000), got], rtol=2e-5)
    assert eng.flagged_count() < 0.05 * 6000


@pytest.mark.parametrize("n,d,k,offset", [(30000, 64, 1024, 0.0), (20000, 32, 200, 50.0), (9000, 128, 600, -300.0),
                                           (5000, 256, 300, 10.0), (4097, 13, 40, 0.0), (2000, 500, 64, 1000.0),
                                           (150001, 16, 257, 0.0)])
def test_assign_f64_on_tensor_cores_equals_float64_kernels(eng, n, d, k, offset):
    """float64 rows decided on the tensor cores (centred, rounded once, flagged rows re-decided in float64) against
    the all-float64 kernels and the oracle: same labels, distances to float64 accuracy."""
    import torch

    rs = np.random.RandomState(n + d + k)
    cen = rs.uniform(-10, 10, size=(max(2, k // 4), d))
    X = cen[rs.randint(0, len(cen), n)] + rs.standard_normal((n, d)) + offset
    C = X[rs.randint(0, n, k)] + 0.2 * rs.standard_normal((k, d))
    Xd = eng.to_device(X)
    pack = eng.pack(eng.to_device(C, torch.float64))
    out = {}
    for tc in (True, False):
        eng.f64_on_tensor_cores = tc
        try:
            lab, m2 = eng.assign(Xd, pack)
            out[tc] = (lab.cpu().numpy(), m2.cpu().numpy())
        finally:
            eng.f64_on_tensor_cores = True
    diff = out[True][0] != out[False][0]
    if diff.any():   # only where float64 itself cannot separate the two centres
        D = np.stack([((X - C[j]) ** 2).sum(1) for j in range(k)], axis=1) if n * k <= 3e7 else None
        assert D is not None, f"{diff.sum()} labels differ"
        r = np.arange(n)
        gap = np.abs(D[r, out[True][0]] - D[r, out[False][0]])
        assert (gap[diff] <= 1e-9 * D[r, out[False][0]][diff] + 1e-13 * ((X ** 2).sum(1)[diff] + (C ** 2).sum(1).max())).all()
    same = ~diff
    npt.assert_allclose(out[True][1][same], out[False][1][same], rtol=1e-12, atol=1e-300)
    # against the oracle (scikit-learn's float64 ArgKmin, expanded form ||x||^2 - 2 x.c + ||c||^2): its squared
    # distances carry a cancellation error |e_i| <= 2 (d + 2) 2^-52 (||x_i||^2 + ||c||^2), so a label may differ
    # only where the two best float64 distances are closer than that noise (or than north_star's 1e-6-relative tie),
    # and a distance by 1e-10 relative plus that noise
    lab_ref, d_ref = orc.pairwise_distances_argmin_min(X, C, metric_kwargs={"squared": True})
    e = 2.0 * (d + 2) * 2.0**-52 * ((X ** 2).sum(1) + (C ** 2).sum(1).max())
    got_lab, got_d2 = out[True]
    bad = got_lab != lab_ref
    if bad.any():
        rows = np.where(bad)[0]
        Db = sklearn.metrics.pairwise_distances(X[rows], C, metric="sqeuclidean")
        part = np.partition(Db, 1, axis=1)
        gap = part[:, 1] - part[:, 0]
        assert (gap <= np.maximum(4 * e[rows], 1e-6 * part[:, 1])).all(), f"{bad.sum()} labels differ from the oracle"
    ok = ~bad
    assert (np.abs(got_d2[ok] - d_ref[ok]) <= 1e-10 * d_ref[ok] + e[ok]).all()
