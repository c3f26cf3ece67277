"""Pins the CPU oracle (oracle/kmeans_oracle.py) by replaying the reference's OWN tests for the
path against it -- their expectations are computed live from scikit-learn:

  reference tests/metrics/test_metrics.py:15-71   pairwise_distances / argmin_min / euclidean
  reference tests/test_kmeans.py:36-39, 64-94, 184-197   row_norms, 3-blob fit, dtypes

plus the committed golden vectors (tests/golden/*.npz, made by tests/golden/make_golden.py FROM THE REFERENCE run
through oracle/reference_loader.py; tests/test_reference_pin.py compares oracle and reference directly).
Runs on CPU (no GPU needed)."""
import glob
import os

import numpy as np
import numpy.testing as npt
import pytest
import sklearn
import sklearn.metrics
from sklearn.cluster import KMeans as SKKMeans

from oracle import kmeans_oracle as orc

HERE = os.path.dirname(os.path.abspath(__file__))


def test_philox_known_answers():
    # Random123 kat_vectors for philox4x32-10
    z = np.zeros((1, 4), np.uint32)
    assert [int(v) for v in orc.philox4x32_10(z, (0, 0))[0]] == [0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8]
    f = np.full((1, 4), 0xFFFFFFFF, np.uint32)
    assert [int(v) for v in orc.philox4x32_10(f, (0xFFFFFFFF, 0xFFFFFFFF))[0]] == [
        0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD]
    p = np.array([[0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344]], np.uint32)
    assert [int(v) for v in orc.philox4x32_10(p, (0xA4093822, 0x299F31D0))[0]] == [
        0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1]
    u = orc.philox_uniform_f64(7, 3, 10, 1000)
    assert u.min() >= 0 and u.max() < 1 and abs(u.mean() - 0.5) < 0.05


def test_row_norms(X_blobs):  # reference tests/test_kmeans.py:36-39
    result = orc.row_norms(X_blobs, squared=True, chunks=(500, 500))
    expected = sklearn.utils.extmath.row_norms(X_blobs, squared=True)
    npt.assert_array_equal(result, expected)


def test_pairwise_distances(X_blobs):  # reference tests/metrics/test_metrics.py:15-19
    centers = X_blobs[::100]
    result = orc.pairwise_distances(X_blobs, centers, chunks=(500, 500))
    expected = sklearn.metrics.pairwise_distances(X_blobs, centers)
    npt.assert_allclose(result, expected, atol=1e-4)


def test_pairwise_distances_argmin_min(X_blobs):  # reference tests/metrics/test_metrics.py:22-39
    centers = X_blobs[::100]
    working_memory = float(80 * 500) / 2**20
    with sklearn.config_context(working_memory=working_memory):
        a_, b_ = sklearn.metrics.pairwise_distances_argmin_min(X_blobs, centers)
        a, b = orc.pairwise_distances_argmin_min(X_blobs, centers, chunks=(500, 500))
    npt.assert_array_equal(a, a_)
    npt.assert_array_equal(b, b_)
    assert a.dtype == np.int64 and b.dtype == np.float64


def test_euclidean_distances():  # reference tests/metrics/test_metrics.py:42-58
    rs = np.random.RandomState(0)
    X, Y = rs.uniform(size=(100, 4)), rs.uniform(size=(100, 4))
    npt.assert_allclose(orc.euclidean_distances(X, Y), sklearn.metrics.euclidean_distances(X, Y))
    xn = (X**2).sum(axis=1)[:, np.newaxis]
    npt.assert_allclose(orc.euclidean_distances(X, Y, X_norm_squared=xn),
                        sklearn.metrics.euclidean_distances(X, Y, X_norm_squared=xn))
    yn = (Y**2).sum(axis=1)[np.newaxis, :]
    npt.assert_allclose(orc.euclidean_distances(X, Y, Y_norm_squared=yn),
                        sklearn.metrics.euclidean_distances(X, Y, Y_norm_squared=yn))


def test_euclidean_distances_same():  # reference tests/metrics/test_metrics.py:61-71
    X = np.random.RandomState(0).uniform(size=(100, 4))
    npt.assert_allclose(orc.euclidean_distances(X, X), sklearn.metrics.euclidean_distances(X, X), atol=1e-4)
    npt.assert_allclose(orc.euclidean_distances(X), sklearn.metrics.euclidean_distances(X), atol=1e-4)


def _replace(a, old, new):
    arr = np.empty(a.max() + 1, dtype=new.dtype)
    arr[old] = new
    return arr[a]


def test_basic(Xl_blobs_easy):  # reference tests/test_kmeans.py:64-94
    X = Xl_blobs_easy
    a = orc.KMeansOracle(n_clusters=3, random_state=0, chunks=(50, 50))
    b = SKKMeans(n_clusters=3, random_state=0, n_init="auto")
    a.fit(X)
    b.fit(X)
    assert abs(a.inertia_ - b.inertia_) < 0.01
    a_order = np.argsort(a.cluster_centers_, 0)[:, 0]
    b_order = np.argsort(b.cluster_centers_, 0)[:, 0]
    npt.assert_allclose(a.cluster_centers_[a_order], b.cluster_centers_[b_order], rtol=1e-3)
    b_labels = _replace(b.labels_, [0, 1, 2], a_order[b_order]).astype(b.labels_.dtype)
    npt.assert_array_equal(a.labels_, b_labels)
    assert a.n_iter_
    b.cluster_centers_ = b.cluster_centers_[b_order]
    a.cluster_centers_ = a.cluster_centers_[a_order]
    npt.assert_allclose(a.transform(X), b.transform(X), rtol=1e-3)
    npt.assert_array_equal(a.predict(X), b.predict(X))


def test_dtypes():  # reference tests/test_kmeans.py:184-197
    X = np.random.RandomState(0).uniform(size=(100, 2))
    X2 = X.astype("f4")
    pairs = [(X, X), (X2, X2), (X, X2), (X2, X)]
    for xx, yy in pairs:
        a = SKKMeans(n_init="auto", random_state=0)
        b = orc.KMeansOracle(random_state=0)
        a.fit(xx)
        b.fit(xx)
        assert a.cluster_centers_.dtype == b.cluster_centers_.dtype
        assert a.labels_.dtype == b.labels_.dtype
        assert a.transform(xx).dtype == b.transform(xx).dtype
        assert a.transform(yy).dtype == b.transform(yy).dtype


def test_too_small():  # reference tests/test_kmeans.py:48-51 -- "found fewer than k" branch
    X = np.random.RandomState(0).uniform(size=(20, 2))
    km = orc.KMeansOracle(random_state=0, chunks=(10, 10)).fit(X)
    assert km.cluster_centers_.shape == (8, 2)


def test_quirks():
    """The behaviours SURVEY section 0 lists, on the restatement itself."""
    rs = np.random.RandomState(3)
    X = rs.standard_normal((300, 3))
    init = X[:4].copy()
    # capped run: inertia is the sum of NON-squared distances (reference :612-616)
    lab, inertia, C, n_iter = orc.kmeans_single_lloyd(X, 4, init, max_iter=1, tol=0.0, chunks=(100, 200))
    a, dmin = sklearn.metrics.pairwise_distances_argmin_min(X, C.astype(np.float64))
    assert n_iter == 1 and np.isclose(inertia, dmin.sum()) and not np.isclose(inertia, (dmin**2).sum())
    # empty cluster -> origin (reference :598-601)
    far = np.vstack([init[:3], [[1e3, 1e3, 1e3]]])
    _, _, C2, _ = orc.kmeans_single_lloyd(X, 4, far, max_iter=1, tol=0.0, chunks=(300,))
    npt.assert_array_equal(C2[3], 0.0)
    # converged run returns the OLD centres and squared inertia (reference :608-610)
    lab3, inertia3, C3, it3 = orc.kmeans_single_lloyd(X, 4, init, max_iter=300, tol=0.0, chunks=(300,))
    a3, d3 = sklearn.metrics.pairwise_distances_argmin_min(X, C3, metric_kwargs={"squared": True})
    npt.assert_array_equal(lab3, a3)
    assert np.isclose(inertia3, d3.sum())


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(HERE, "golden", "*.npz"))))
def test_golden_vectors(path):
    """The committed fixtures -- the REFERENCE's outputs -- must be what the oracle produces, here and on the GPU box
    (where the reference tree is absent; also guards against drift of sklearn / numba / numpy versions)."""
    g = np.load(path, allow_pickle=False)
    kind = str(g["kind"])
    if kind == "lloyd":
        chunks = tuple(int(c) for c in g["chunks"])
        lab, inertia, C, n_iter = orc.kmeans_single_lloyd(
            g["X"], int(g["k"]), g["init"], max_iter=int(g["max_iter"]), tol=float(g["tol"]), chunks=chunks)
        npt.assert_array_equal(lab, g["labels"])
        npt.assert_allclose(C, g["centers"], rtol=1e-12, atol=1e-12)
        npt.assert_allclose(inertia, float(g["inertia"]), rtol=1e-12)
        assert n_iter == int(g["n_iter"])
    elif kind == "init":
        trace = []
        rs = orc.PhiloxDraws(int(g["seed"]))
        C = orc.init_scalable(g["X"], int(g["k"]), rs, None, float(g["l"]), "auto", tuple(int(c) for c in g["chunks"]),
                              supplement=lambda n, need: g["supplement"], trace=trace)
        picked = sorted(set(v for t in trace if "picked" in t for v in t["picked"]))
        npt.assert_array_equal(np.array(picked, dtype=np.int64), g["picked"])
        npt.assert_allclose(C, g["centers"], rtol=1e-12)
    else:
        raise AssertionError(kind)
