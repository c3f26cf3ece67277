"""-m gpu: SpectralClustering (the in-tree caller of KMeans) -- the reference's own tests
(reference tests/test_spectral_clustering.py) replayed, plus the Nystrom embedding and the pairwise kernels
against the numpy restatement in the oracle."""
import numpy as np
import numpy.testing as npt
import pytest
import sklearn.cluster
import sklearn.metrics

from conftest import easy_blobs
from oracle import kmeans_oracle as orc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def X():
    from sklearn.datasets import make_blobs

    # reference: dask_ml.datasets.make_blobs(n_samples=200, chunks=100, random_state=0) -- two 100-row blocks
    proto, yp = make_blobs(n_samples=100, n_features=2, centers=3, random_state=0)
    centers = np.vstack([proto[yp == i].mean(0) for i in range(3)])
    return np.vstack([make_blobs(n_samples=100, n_features=2, centers=centers, random_state=i)[0] for i in (0, 1)])


def _sc():
    from dask_ml_b200.cluster import SpectralClustering

    return SpectralClustering


@pytest.mark.parametrize("persist_embedding", [True, False])
def test_basic(X, persist_embedding):  # reference :15-27
    sc = _sc()(n_components=25, random_state=0, persist_embedding=persist_embedding)
    sc.fit(X)
    assert len(sc.labels_) == len(X)


@pytest.mark.parametrize("assign_labels", [sklearn.cluster.KMeans(n_init=2), "sklearn-kmeans"])
def test_sklearn_kmeans(X, assign_labels):  # reference :30-41
    sc = _sc()(n_components=25, random_state=0, assign_labels=assign_labels, kmeans_params={"n_clusters": 8})
    sc.fit(X)
    assert isinstance(sc.assign_labels_, sklearn.cluster.KMeans)


def test_n_components_raises(X):  # reference :56-60
    with pytest.raises(ValueError, match="n_components"):
        _sc()(n_components=len(X)).fit(X)


def test_assign_labels_raises(X):  # reference :63-74
    with pytest.raises(ValueError, match="Unknown 'assign_labels' 'foo'"):
        _sc()(assign_labels="foo").fit(X)
    with pytest.raises(TypeError, match="Invalid type "):
        _sc()(assign_labels=dict()).fit(X)


def test_affinity_raises(X):  # reference :77-87
    with pytest.raises(ValueError, match="Unknown affinity metric name 'foo'"):
        _sc()(affinity="foo").fit(X)
    with pytest.raises(TypeError):
        _sc()(affinity=np.array([])).fit(X)


def test_spectral_clustering_easy_blobs():  # reference :90-104
    from sklearn.datasets import make_blobs

    centers = np.array([[-7, -7], [0, 0], [7, 7]])
    parts = [make_blobs(n_samples=50, n_features=2, centers=centers, cluster_std=0.1, random_state=i) for i in (0, 1)]
    Xe = np.vstack([p[0] for p in parts])
    y = np.concatenate([p[1] for p in parts])
    Xe = (Xe - Xe.mean(0)) / Xe.std(0)
    model = _sc()(random_state=0, n_clusters=3, n_components=5, gamma=None).fit(Xe)
    labels = model.labels_.compute()
    for i in range(3):
        assert len(set(labels[y == i])) == 1


@pytest.mark.parametrize("metric,kw", [("rbf", {"gamma": 0.3}), ("rbf", {}), ("linear", {}),
                                       ("polynomial", {"degree": 2, "gamma": 0.1, "coef0": 0.5}),
                                       ("sigmoid", {"gamma": 0.05, "coef0": 0.1})])
def test_pairwise_kernels_match_sklearn(metric, kw):  # reference tests/metrics/test_metrics.py:74-81
    from dask_ml_b200.metrics import pairwise_kernels

    rs = np.random.RandomState(0)
    A, B = rs.standard_normal((300, 7)), rs.standard_normal((70, 7))
    got = pairwise_kernels(A, B, metric=metric, **kw).compute()
    npt.assert_allclose(got, sklearn.metrics.pairwise.pairwise_kernels(A, B, metric=metric, **kw), rtol=1e-9, atol=1e-12)
    got32 = pairwise_kernels(A.astype("f4"), B.astype("f4"), metric=metric, **kw)
    assert got32.dtype == np.float32
    npt.assert_allclose(got32.compute(), got, rtol=2e-4, atol=2e-5)
    with pytest.raises(ValueError, match="Unknown kernel"):
        pairwise_kernels(A, B, metric="nope")


def test_nystrom_embedding_matches_oracle():
    """Same sampled rows -> same embedding (up to the sign of each singular vector) and same partition."""
    from sklearn.metrics import adjusted_rand_score

    rs = np.random.RandomState(4)
    Xb = np.vstack([rs.normal(c, 0.4, size=(150, 3)) for c in ([0, 0, 0], [4, 4, 0], [0, 4, 4], [4, 0, 4])])
    rs.shuffle(Xb)
    sc = _sc()(n_clusters=4, n_components=40, gamma=0.5, random_state=1).fit(Xb)
    U2_ref, S_ref = orc.spectral_embedding(Xb, sc._keep, 4, metric="rbf", gamma=0.5)
    U2 = sc._embedding.cpu().numpy()
    sign = np.sign((U2 * U2_ref).sum(0))
    npt.assert_allclose(U2 * sign, U2_ref, rtol=1e-6, atol=1e-7)
    npt.assert_allclose(sc.eigenvalues_.compute(), S_ref, rtol=1e-9)
    ref_labels = sklearn.cluster.KMeans(4, n_init=5, random_state=0).fit_predict(U2_ref)
    assert adjusted_rand_score(ref_labels, sc.labels_.compute()) == 1.0
