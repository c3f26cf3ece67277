import os
import sys

# see bench.py: OpenBLAS's thread tables overflow on many-core boxes; the oracle needs one BLAS thread per block
os.environ.setdefault("OPENBLAS_NUM_THREADS", "1")

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box: pytest -m gpu)")


def _cuda_ok():
    try:
        import torch

        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _cuda_ok():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this process")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def easy_blobs():
    """The reference's ``Xl_blobs_easy`` fixture rebuilt without dask
    (reference tests/conftest.py:61-70 + dask_ml/datasets.py:180-191: one sklearn make_blobs
    call per 50-row block, block i seeded with random_state=i)."""
    from sklearn.datasets import make_blobs

    centers = np.array([[-7, -7], [0, 0], [7, 7]])
    return np.vstack([
        make_blobs(n_samples=50, n_features=2, centers=centers, cluster_std=0.1, shuffle=True,
                   center_box=(-10, 10), random_state=i)[0]
        for i in (0, 1)
    ])


@pytest.fixture
def Xl_blobs_easy():
    return easy_blobs()


@pytest.fixture
def X_blobs():
    """Stand-in for the reference's ``X_blobs`` (tests/conftest.py:49-58, 73-78): 1000 x 4 in two
    500-row blocks.  The reference draws it from da.random (not reproducible without dask); its
    tests only compare against sklearn on the same data, so the identity of the data is free."""
    return np.random.RandomState(1).standard_normal((1000, 4))


def blobs(n, d, k, seed, dtype=np.float64, std=1.0, box=10.0):
    rs = np.random.RandomState(seed)
    centers = rs.uniform(-box, box, size=(k, d))
    lab = rs.randint(0, k, size=n)
    X = centers[lab] + std * rs.standard_normal((n, d))
    return X.astype(dtype), centers
