"""Pins the CPU oracle to the reference's OWN source.

``oracle/reference_loader.py`` imports the unmodified ``/root/reference/dask_ml/{cluster/k_means.py,
metrics/pairwise.py, utils.py, ...}`` on top of the eager dask stand-in in ``oracle/dask_stub`` and these tests run
the reference's functions and the oracle's restatement on the same inputs and demand the same answers: labels and
picked rows exact, floats to 1e-12 relative (they are in fact bit-equal; the margin is for other BLAS builds).

Needs the reference tree, so it runs in the build container only (CPU); on the GPU box -- where ``/root/reference``
does not exist -- the fixtures ``tests/golden/*.npz``, generated here FROM THE REFERENCE by ``make_golden.py``, carry
the pin (``tests/test_oracle_pin.py::test_golden_vectors``, ``tests/test_gpu_kmeans.py``).
"""
import numpy as np
import numpy.testing as npt
import pytest

from conftest import blobs, easy_blobs
from oracle import kmeans_oracle as orc
from oracle import reference_loader

pytestmark = pytest.mark.skipif(not reference_loader.available(), reason="reference tree not present on this box")

RTOL = 1e-12


@pytest.fixture(scope="module")
def ref():
    return reference_loader.load()


def _chunks_of(n, nblocks):
    base = n // nblocks
    sizes = [base] * nblocks
    sizes[-1] += n - base * nblocks
    return tuple(sizes)


# shapes after BASELINE.json's configs (C1: d=16 k=8 f64, C2: d=32 k=256 f32, C3: d=64 k=1024 f32, C4: d=128 k large f32)
LLOYD_CASES = [
    # name,            n,    d,   k,    dtype,      seed, max_iter, tol,  nblocks, init
    ("c1_capped",      4000, 16,  8,    np.float64, 0,    4,        0.0,  4,       "rows"),
    ("c1_converged",   3000, 16,  8,    np.float64, 1,    300,      1e-9, 3,       "near"),
    ("c2_capped",      6000, 32,  256,  np.float32, 2,    3,        0.0,  5,       "rows"),
    ("c2_converged",   5000, 32,  64,   np.float32, 3,    300,      1e-4, 2,       "near"),
    ("c3_capped",      5000, 64,  1024, np.float32, 4,    2,        0.0,  4,       "rows"),
    ("c4_capped",      3000, 128, 1500, np.float32, 5,    2,        0.0,  3,       "rows"),
    ("one_block",      1500, 5,   3,    np.float64, 6,    300,      1e-9, 1,       "rows"),
    ("many_blocks",    2100, 7,   12,   np.float64, 7,    5,        0.0,  17,      "rows"),   # two levels of dask's sum tree
    ("empty_cluster",  2000, 8,   6,    np.float32, 8,    1,        0.0,  4,       "far"),
    ("tol_break",      2500, 20,  40,   np.float32, 9,    100,      5.0,  3,       "rows"),   # shift < tol with shift > 1e-7
]


def _lloyd_inputs(n, d, k, dtype, seed, init_mode):
    X, true = blobs(n, d, k, seed, dtype)
    if init_mode == "rows":
        init = X[:k].copy()
    elif init_mode == "near":
        init = (true + 0.5).astype(dtype)
    else:   # one centre far from every row: its cluster stays empty -> origin (reference :598-601)
        init = X[:k].copy()
        init[-1] = 1e3
    return X, init


@pytest.mark.parametrize("case", LLOYD_CASES, ids=[c[0] for c in LLOYD_CASES])
def test_lloyd_reference_equals_oracle(ref, case):
    """reference dask_ml/cluster/k_means.py:549-632 (Lloyd loop + numba scatter) vs oracle.kmeans_single_lloyd."""
    _, n, d, k, dtype, seed, max_iter, tol, nblocks, init_mode = case
    X, init = _lloyd_inputs(n, d, k, dtype, seed, init_mode)
    chunks = _chunks_of(n, nblocks)
    Xd = ref.da.from_array(X, chunks=(chunks, (d,)))
    lab_r, inertia_r, C_r, it_r = ref.k_means._kmeans_single_lloyd(Xd, k, init=init, max_iter=max_iter, tol=tol)
    lab_o, inertia_o, C_o, it_o = orc.kmeans_single_lloyd(X, k, init, max_iter=max_iter, tol=tol, chunks=chunks)
    assert it_r == it_o
    lab_r = lab_r.compute()
    assert lab_r.dtype == np.int32 == lab_o.dtype
    npt.assert_array_equal(lab_r, lab_o)
    assert C_r.dtype == dtype == C_o.dtype
    npt.assert_allclose(C_r, C_o, rtol=RTOL, atol=0)
    npt.assert_allclose(float(inertia_r.compute()), inertia_o, rtol=RTOL)
    if init_mode == "far":
        npt.assert_array_equal(C_r[-1], 0.0)


def test_lloyd_inertia_branches(ref):
    """Both exits of reference :612-616: squared distances when the loop ended with shift <= 1e-7, a fresh
    NON-squared pass otherwise -- asserted on the reference itself, then oracle == reference."""
    import sklearn.metrics

    X, true = blobs(1200, 4, 5, 11, np.float64)
    Xd = ref.da.from_array(X, chunks=(400, 4))
    init = (true + 0.3)
    _, inertia_c, C_c, it_c = ref.k_means._kmeans_single_lloyd(Xd, 5, init=init, max_iter=300, tol=1e-12)
    _, d_sq = sklearn.metrics.pairwise_distances_argmin_min(X, C_c, metric_kwargs={"squared": True})
    assert it_c < 300 and np.isclose(float(inertia_c.compute()), d_sq.sum(), rtol=1e-10)
    _, inertia_n, C_n, it_n = ref.k_means._kmeans_single_lloyd(Xd, 5, init=X[:5].copy(), max_iter=1, tol=0.0)
    _, d_lin = sklearn.metrics.pairwise_distances_argmin_min(X, C_n)
    assert it_n == 1 and np.isclose(float(inertia_n.compute()), d_lin.sum(), rtol=1e-10)
    for init_, mi in ((init, 300), (X[:5].copy(), 1)):
        r = ref.k_means._kmeans_single_lloyd(Xd, 5, init=init_, max_iter=mi, tol=1e-12)
        o = orc.kmeans_single_lloyd(X, 5, init_, max_iter=mi, tol=1e-12, chunks=(400, 400, 400))
        npt.assert_allclose(float(r[1].compute()), o[1], rtol=RTOL)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("squared", [False, True])
def test_argmin_min_reference_equals_oracle(ref, dtype, squared):
    """reference dask_ml/metrics/pairwise.py:19-47."""
    X, true = blobs(3000, 24, 50, 20, dtype)
    kw = {"squared": True} if squared else None
    Xd = ref.da.from_array(X, chunks=(700, 24))
    a_r, m_r = ref.pairwise.pairwise_distances_argmin_min(Xd, true, metric="euclidean", metric_kwargs=kw)
    a_o, m_o = orc.pairwise_distances_argmin_min(X, true, chunks=Xd.chunks[0], metric_kwargs=kw)
    npt.assert_array_equal(a_r.compute(), a_o)
    npt.assert_array_equal(m_r.compute(), m_o)
    assert a_r.compute().dtype == np.int64 and m_r.compute().dtype == np.float64


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_pairwise_and_euclidean_reference_equals_oracle(ref, dtype):
    """reference dask_ml/metrics/pairwise.py:50-67 and :70-105."""
    X, true = blobs(900, 10, 7, 21, dtype)
    C = true.astype(dtype)
    Xd = ref.da.from_array(X, chunks=(400, 10))
    r = ref.pairwise.pairwise_distances(Xd, C).compute()
    o = orc.pairwise_distances(X, C, chunks=Xd.chunks[0])
    assert r.dtype == o.dtype
    npt.assert_array_equal(r, o)
    for sq in (False, True):
        r = ref.pairwise.euclidean_distances(Xd, C, squared=sq).compute()
        o = orc.euclidean_distances(X, C, squared=sq)
        assert r.dtype == o.dtype
        npt.assert_allclose(r, o, rtol=1e-6 if dtype == np.float32 else RTOL, atol=1e-4 if dtype == np.float32 else 1e-10)
    with pytest.raises(TypeError):
        ref.pairwise.pairwise_distances(Xd, ref.da.from_array(C))
    npt.assert_array_equal(ref.utils.row_norms(Xd, squared=True).compute(),
                           orc.row_norms(X, squared=True, chunks=Xd.chunks[0]))


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_evaluate_cost_and_sample_points_reference_equals_oracle(ref, dtype):
    """reference dask_ml/cluster/k_means.py:516-541 with the SAME uniform draws on both sides (a draw provider
    whose ``uniform(size=, chunks=)`` hands back a fixed vector)."""
    X, _ = blobs(5000, 12, 9, 30, dtype)
    chunks = (1300, 1300, 1300, 1100)
    Xd = ref.da.from_array(X, chunks=(chunks, (12,)))
    centers = X[[0, 17, 400]]
    cost_r = float(ref.dask.compute(ref.k_means.evaluate_cost(Xd, centers))[0])
    cost_o = orc.evaluate_cost(X, centers, chunks)
    npt.assert_allclose(cost_r, cost_o, rtol=RTOL)
    draws = np.random.RandomState(5).uniform(size=len(X)) * 5e-3   # scaled so that a few dozen rows are picked

    class Fixed:
        def uniform(self, size, chunks):
            assert size == len(draws) and tuple(chunks[0]) == chunks_expected
            return ref.da.from_array(draws, chunks=chunks)

    chunks_expected = chunks
    idx_r = ref.dask.compute(ref.k_means._sample_points(Xd, centers, 2, Fixed()))[0]
    idx_o = orc.sample_points(X, centers, 2, draws, chunks)
    assert 5 < len(idx_o) < 2000
    npt.assert_array_equal(np.sort(idx_r), idx_o)


INIT_CASES = [
    # name,              n,    d,  k,  dtype,      seed, l, init_max_iter, nblocks
    ("supplement_f64",   3000, 4,  40, np.float64, 0,    2, None,          3),    # ~2 ln(cost) candidates < k
    ("supplement_f32",   4000, 8,  64, np.float32, 1,    2, None,          2),
    ("recluster_f64",    2000, 3,  3,  np.float64, 2,    2, None,          4),    # candidates >= k -> sklearn KMeans
    ("recluster_f32",    2500, 6,  5,  np.float32, 3,    4, None,          5),
    ("capped_rounds",    2000, 5,  20, np.float64, 4,    3, 2,             2),
    ("one_block",        1000, 2,  8,  np.float64, 5,    2, None,          1),
]


@pytest.mark.parametrize("case", INIT_CASES, ids=[c[0] for c in INIT_CASES])
def test_init_scalable_reference_equals_oracle(ref, case):
    """reference dask_ml/cluster/k_means.py:435-513 driven by the stub's ``da.random.RandomState(seed)``; the oracle
    by its own restatement of that class (both follow the same reading of dask: the per-chunk seeding is the one
    ASSUMPTION of this pin)."""
    _, n, d, k, dtype, seed, l, init_max_iter, nblocks = case
    X, _ = blobs(n, d, max(k // 4, 3), 100 + seed, dtype)
    chunks = _chunks_of(n, nblocks)
    Xd = ref.da.from_array(X, chunks=(chunks, (d,)))
    C_r = ref.k_means.init_scalable(Xd, k, ref.da.random.RandomState(seed), init_max_iter, l)
    trace = []
    C_o = orc.init_scalable(X, k, orc.DaskLikeRandomState(seed), init_max_iter, l, "auto", chunks, trace=trace)
    assert C_r.shape == C_o.shape == (k, d)
    npt.assert_allclose(C_r, C_o, rtol=RTOL, atol=0)
    took_supplement = any("supplement" in t for t in trace)
    if case[0].startswith(("supplement", "capped", "recluster")):   # the branch the case is named after was taken
        assert took_supplement == (not case[0].startswith("recluster"))


def test_k_init_validation_messages(ref):
    """reference :360-401 -- the messages the product repeats."""
    X = ref.da.from_array(np.zeros((10, 3)), chunks=(5, 3))
    with pytest.raises(ValueError, match="does not match 'n_clusters'"):
        ref.k_means.k_init(X, 4, init=np.zeros((3, 3)))
    with pytest.raises(ValueError, match="do not match the number of features"):
        ref.k_means.k_init(X, 3, init=np.zeros((3, 2)))
    with pytest.raises(TypeError, match="must be an array or str"):
        ref.k_means.k_init(X, 3, init=5)
    with pytest.raises(ValueError, match="must be one of"):
        ref.k_means.k_init(X, 3, init="nope")


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_estimator_reference_equals_oracle(ref, dtype):
    """``KMeans(init='k-means||', random_state=int).fit(ndarray)`` end to end (reference :171-223): chunking by
    ``cpu_count()`` (:192-193), NaN check, k-means||, Lloyd, published attributes."""
    from multiprocessing import cpu_count

    X, _ = blobs(4000, 6, 5, 40, dtype)
    a = ref.k_means.KMeans(n_clusters=5, random_state=3, max_iter=20).fit(X)
    chunks = orc.row_chunks(len(X), ncpu=cpu_count())
    b = orc.KMeansOracle(n_clusters=5, random_state=3, max_iter=20, chunks=chunks).fit(X)
    npt.assert_array_equal(a.labels_.compute(), b.labels_)
    npt.assert_allclose(a.cluster_centers_, b.cluster_centers_, rtol=RTOL)
    npt.assert_allclose(a.inertia_, b.inertia_, rtol=RTOL)
    assert a.n_iter_ == b.n_iter_ and a.n_features_in_ == 6
    assert a.cluster_centers_.dtype == dtype
    npt.assert_array_equal(a.predict(X).compute(), b.predict(X))
    npt.assert_allclose(a.transform(X).compute(), b.transform(X), rtol=1e-5, atol=1e-4)
    Xbad = X.copy()
    Xbad[7, 2] = np.nan
    # an ndarray is rejected by scikit-learn's own check_array inside utils.check_array (reference utils.py:214) with
    # scikit-learn's message; a dask array reaches the reference's own check and message (k_means.py:195-201)
    with pytest.raises(ValueError, match="Input contains NaN"):
        ref.k_means.KMeans(n_clusters=5).fit(Xbad)
    with pytest.raises(ValueError, match="Input contains NaN, infinity or a value too large"):
        ref.k_means.KMeans(n_clusters=5).fit(ref.da.from_array(Xbad, chunks=(1000, 6)))


def test_reference_own_basic_test(ref):
    """reference tests/test_kmeans.py:64-94 (3 easy blobs against scikit-learn) run on the reference itself through
    the stub: a check of the stub, not of the oracle."""
    from sklearn.cluster import KMeans as SKKMeans

    X = easy_blobs()
    a = ref.k_means.KMeans(n_clusters=3, random_state=0).fit(ref.da.from_array(X, chunks=(50, 2)))
    b = SKKMeans(n_clusters=3, random_state=0, n_init="auto").fit(X)
    assert abs(a.inertia_ - b.inertia_) < 0.01
    ao, bo = np.argsort(a.cluster_centers_, 0)[:, 0], np.argsort(b.cluster_centers_, 0)[:, 0]
    npt.assert_allclose(a.cluster_centers_[ao], b.cluster_centers_[bo], rtol=1e-3)


def test_stub_reduction_tree_and_rng(ref):
    """The two stub behaviours results depend on, stated as tests: ``sum()`` = per-chunk float64 partials combined
    four at a time; ``RandomState`` = one MT19937 stream per chunk seeded from the parent's byte stream."""
    v = np.random.RandomState(0).standard_normal(1000).astype(np.float32) ** 2
    chunks = (130,) * 7 + (90,)
    got = ref.da.from_array(v, chunks=(chunks,)).map_blocks(lambda b: b, dtype=float).sum().compute()
    parts = [np.sum(v[i:i + 130], dtype=np.float64) for i in range(0, 1000, 130)]
    want = np.sum([np.sum(parts[:4]), np.sum(parts[4:])])
    assert got == want == orc.dask_sum(v, chunks)
    rs = ref.da.random.RandomState(42)
    u = rs.uniform(size=10, chunks=((6, 4),)).compute()
    parent = np.random.RandomState(42)
    words = np.frombuffer(parent.bytes(624 * 2 * 4), dtype="<u4").reshape(2, -1)
    want_u = np.concatenate([np.random.RandomState(words[0]).uniform(size=6), np.random.RandomState(words[1]).uniform(size=4)])
    npt.assert_array_equal(u, want_u)
    npt.assert_array_equal(orc.DaskLikeRandomState(42).uniform(10, (6, 4)), want_u)
