"""a host matrix that stays on the host and is streamed through the device on every pass
(k-means|| rounds, Lloyd iterations, final labelling).  Forced here with B200KM_OUT_OF_CORE=1 and tiny blocks so
that every pass has many blocks and a ragged tail; results must equal the resident run."""
import numpy as np
import numpy.testing as npt
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture()
def ooc(monkeypatch):
    monkeypatch.setenv("B200KM_OUT_OF_CORE", "1")
    monkeypatch.setenv("B200KM_CHUNK_BYTES", str(37 * 1000))   # a few hundred rows per block, ragged


def _data(n=20011, d=24, k=12, dtype=np.float32, seed=0):
    rs = np.random.RandomState(seed)
    cen = rs.uniform(-8, 8, size=(k, d))
    X = (cen[rs.randint(0, k, n)] + rs.standard_normal((n, d))).astype(dtype)
    return X, cen


def _fit(X, **kw):
    from dask_ml_b200.cluster import KMeans

    return KMeans(**kw).fit(X)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("init", ["array", "k-means||", "random", "k-means++"])
def test_streamed_fit_equals_resident_fit(monkeypatch, init, dtype):
    X, cen = _data(dtype=dtype)
    kw = dict(n_clusters=12, random_state=3, max_iter=8, tol=0.0)
    kw["init"] = (X[:12].astype(np.float64) + 0.05) if init == "array" else init
    monkeypatch.setenv("B200KM_OUT_OF_CORE", "0")
    a = _fit(X, **kw)
    monkeypatch.setenv("B200KM_OUT_OF_CORE", "1")
    monkeypatch.setenv("B200KM_CHUNK_BYTES", str(37 * 1000))
    b = _fit(X, **kw)
    assert a.n_iter_ == b.n_iter_
    npt.assert_array_equal(a.labels_.compute(), b.labels_.compute())
    npt.assert_allclose(a.cluster_centers_, b.cluster_centers_, rtol=1e-6 if dtype == np.float32 else 1e-11)
    npt.assert_allclose(a.inertia_, b.inertia_, rtol=1e-9)
    npt.assert_array_equal(np.asarray(b.predict(X).compute()), np.asarray(a.predict(X).compute()))


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_streamed_fit_tensor_core_path(monkeypatch, dtype):
    """k > 16: K1 on the tensor cores + re-decision + K2, block by block (float64: the centred / rounded route)."""
    X, cen = _data(n=30011, d=24, k=40, seed=5, dtype=dtype)
    kw = dict(n_clusters=40, random_state=1, max_iter=5, tol=0.0, init=X[:40].astype(np.float64) + 0.05)
    monkeypatch.setenv("B200KM_OUT_OF_CORE", "0")
    a = _fit(X, **kw)
    monkeypatch.setenv("B200KM_OUT_OF_CORE", "1")
    monkeypatch.setenv("B200KM_CHUNK_BYTES", str(512 * 1024))
    b = _fit(X, **kw)
    npt.assert_array_equal(a.labels_.compute(), b.labels_.compute())
    npt.assert_allclose(a.cluster_centers_, b.cluster_centers_, rtol=1e-6 if dtype == np.float32 else 1e-11)
    npt.assert_allclose(a.inertia_, b.inertia_, rtol=1e-9)


def test_streamed_fit_matches_oracle(ooc):
    from oracle import kmeans_oracle as orc

    X, cen = _data(n=9001, d=10, k=6)
    init = X[:6].astype(np.float64) + 0.1
    km = _fit(X, n_clusters=6, init=init, max_iter=20, tol=1e-4, random_state=0)
    ref = orc.KMeansOracle(n_clusters=6, init=init, max_iter=20, tol=1e-4, random_state=0).fit(X)
    assert km.n_iter_ == ref.n_iter_
    npt.assert_array_equal(km.labels_.compute(), ref.labels_)
    npt.assert_allclose(km.cluster_centers_, ref.cluster_centers_, rtol=1e-5)
    npt.assert_allclose(km.inertia_, ref.inertia_, rtol=1e-5)


def test_streamed_fit_finds_nan_on_the_first_pass(ooc):
    X, _ = _data(n=5000, d=8, k=4)
    X[4321, 3] = np.nan
    with pytest.raises(ValueError, match="NaN|infinity|finite"):
        _fit(X, n_clusters=4, init=X[:4].astype(np.float64), max_iter=2)


def test_streamed_transform_is_refused(ooc):
    X, _ = _data(n=3000, d=8, k=4)
    km = _fit(X, n_clusters=4, init=X[:4].astype(np.float64), max_iter=2)
    with pytest.raises(NotImplementedError):
        km.transform(X)
