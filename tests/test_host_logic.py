"""CPU-only tests of the host-side mirror of the reference interface (no kernels involved)."""
import importlib

import numpy as np
import pytest
import torch

kmod = importlib.import_module("dask_ml_b200.cluster.k_means")
from dask_ml_b200.cluster import KMeans  # noqa: E402
from dask_ml_b200.utils import DeviceArray, _timed, _timer, check_array, row_norms  # noqa: E402


def test_constructor_matches_reference_signature():
    # reference dask_ml/cluster/k_means.py:143-157
    km = KMeans()
    assert km.get_params() == dict(
        n_clusters=8, init="k-means||", oversampling_factor=2, max_iter=300, tol=0.0001,
        precompute_distances="auto", random_state=None, copy_x=True, n_jobs=1, algorithm="full",
        init_max_iter=None, n_init="auto")
    from sklearn.base import clone

    assert clone(KMeans(n_clusters=3, tol=0.5)).get_params()["tol"] == 0.5
    assert not hasattr(km, "fit_predict") and not hasattr(km, "score")  # API gaps kept
    import inspect

    sig = inspect.signature(kmod.k_means)
    assert sig.parameters["n_jobs"].default == -1 and sig.parameters["init"].default == "k-means||"


def test_k_init_validation_messages():
    # reference dask_ml/cluster/k_means.py:360-401
    X = torch.zeros(10, 4)
    with pytest.raises(ValueError, match=r"Number of centers in provided 'init' \(2\) does not match 'n_clusters' \(3\)"):
        kmod.k_init(X, 3, init=np.zeros((2, 4)))
    with pytest.raises(ValueError, match="Number of features in the provided 'init'"):
        kmod.k_init(X, 3, init=np.zeros((3, 2)))
    with pytest.raises(TypeError, match="'init' must be an array or str"):
        kmod.k_init(X, 3, init=5)
    init = np.ones((3, 4))
    assert kmod.k_init(X, 3, init=init) is init


def test_philox_random_state_is_deterministic_and_o_of_need():
    a, b = kmod.PhiloxRandomState(5), kmod.PhiloxRandomState(5)
    np.testing.assert_array_equal(a.choice(10**12, 7), b.choice(10**12, 7))   # no arange(n) materialised
    assert len(set(a.choice(100, 100).tolist())) == 100                       # without replacement
    assert a.randint(0, 10) == b.randint(0, 10)
    assert kmod.PhiloxRandomState(None).seed != kmod.PhiloxRandomState(None).seed


def test_check_array_and_row_norms_host_paths():
    import sklearn.utils.extmath as skm

    X = np.random.RandomState(0).rand(20, 3)
    np.testing.assert_array_equal(row_norms(X, squared=True), skm.row_norms(X, squared=True))
    assert check_array(X) is not None
    with pytest.raises(ValueError):
        check_array(np.zeros((0, 3)))
    t = torch.zeros(5, 3)
    assert check_array(t) is t
    with pytest.raises(ValueError):
        check_array(torch.zeros(5, 0))


def test_device_array_surface():
    a = DeviceArray(torch.arange(6, dtype=torch.int64).reshape(3, 2))
    assert a.shape == (3, 2) and a.dtype == np.int64 and len(a) == 3 and a.ndim == 2
    np.testing.assert_array_equal(a.compute(), np.arange(6).reshape(3, 2))
    np.testing.assert_array_equal(np.asarray(a), a.compute())
    assert a.astype(np.int32).dtype == np.int32
    assert a[1].compute().tolist() == [2, 3]


def test_timer_logs(caplog):
    import logging

    lg = logging.getLogger("t")
    with caplog.at_level(logging.INFO, logger="t"):
        with _timer("thing", _logger=lg):
            pass

        @_timed(_logger=lg)
        def f(x):
            return x + 1

        assert f(1) == 2
    msgs = [r.getMessage() for r in caplog.records]
    assert any(m.startswith("Starting thing") for m in msgs) and any(m.startswith("Finished f in") for m in msgs)


def test_error_code_mapping():
    from dask_ml_b200 import _lib

    with pytest.raises(ValueError, match="invalid argument"):
        _lib.check(-1, "op")
    with pytest.raises(_lib.B200KernelError):
        _lib.check(700, "op")  # cudaErrorIllegalAddress
    _lib.check(0)
