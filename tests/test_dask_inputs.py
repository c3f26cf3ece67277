"""dask collections as INPUT to the drop-in surface (reference dask_ml/utils.py:160-207, dask_ml/cluster/k_means.py:
171-202, tests/test_kmeans.py:200-211).  dask itself is not installable here; the eager stand-in under
``oracle/dask_stub`` (the one the unmodified reference modules run on, tests/test_reference_pin.py) provides
``dask.array.Array`` objects with real chunk structure, which is all these branches look at."""
import os
import sys

import numpy as np
import numpy.testing as npt
import pytest

STUB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "dask_stub")


@pytest.fixture(scope="module")
def da():
    if "dask" not in sys.modules:
        sys.path.insert(0, STUB)
        try:
            import dask.array  # noqa: F401
        finally:
            sys.path.remove(STUB)
    import dask.array as da_

    return da_


def test_check_array_dask_branches(da):
    from dask_ml_b200.utils import check_array

    X = np.random.RandomState(0).uniform(size=(100, 4))
    ok = da.from_array(X, chunks=(50, 4))
    assert check_array(ok) is ok                                     # reference utils.py:160-197: returned unchanged
    with pytest.raises(TypeError, match="Chunking is only allowed on the first axis"):
        check_array(da.from_array(X, chunks=(50, 2)))                # :171-178
    assert check_array(da.from_array(X, chunks=(50, 2)), accept_multiple_blocks=True) is not None
    unknown = da.where(ok[:, 0] > 0.5)[0]                            # dask gives such arrays nan chunks
    unknown._chunks = ((float("nan"),),)
    with pytest.raises(TypeError, match="unknown chunk sizes"):
        check_array(unknown, ensure_2d=False)                        # :163-170
    with pytest.raises(TypeError):
        check_array(ok, accept_dask_array=False)                     # :161-162
    import dask.dataframe as dd

    with pytest.raises(TypeError, match="does not support dask dataframes"):
        check_array(dd.DataFrame())                                  # :199-207


def test_pairwise_distances_rejects_dask_Y(da):
    from dask_ml_b200.metrics import pairwise_distances

    X = np.zeros((10, 3))
    with pytest.raises(TypeError, match="`Y` must be a numpy array"):    # reference metrics/pairwise.py:57-58
        pairwise_distances(X, da.from_array(X, chunks=(5, 3)))


@pytest.mark.gpu
def test_fit_on_dask_array_with_empty_chunk(da):
    """reference tests/test_kmeans.py:207-211 (issue 551: a zero-length chunk) and :200-204: the collection is computed
    once, block structure and all, and handed to the device; results equal the ndarray fit."""
    from dask_ml_b200.cluster import KMeans

    X = np.random.RandomState(1).uniform(size=(100, 4))
    Xd = da.from_array(X, chunks=((0, 5, 95), (4,)))
    km = KMeans(n_clusters=2, random_state=0)
    trn = km.fit_transform(Xd)
    assert tuple(trn.shape) == (100, 2)
    km2 = KMeans(n_clusters=2, random_state=0).fit(X)
    npt.assert_allclose(km.cluster_centers_, km2.cluster_centers_, rtol=1e-12)
    npt.assert_array_equal(km.labels_.compute(), km2.labels_.compute())
    npt.assert_array_equal(km.predict(Xd).compute(), km2.predict(X).compute())
