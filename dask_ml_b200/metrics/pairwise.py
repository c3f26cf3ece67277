"""Device versions of the three distance operators on the k-means path.

Signatures, argument meaning and error behaviour follow reference
dask_ml/metrics/pairwise.py:19-105.  Where the reference returns lazy dask arrays these return
``DeviceArray`` (``.compute()`` gives the ndarray).
"""
from __future__ import annotations

import warnings
from typing import Any, Dict, Optional

import numpy as np
import torch

from .._engine import get_engine
from ..utils import DeviceArray, _adopt_device_buffer

__all__ = ["pairwise_distances_argmin_min", "pairwise_distances", "euclidean_distances", "pairwise_kernels",
           "rbf_kernel", "linear_kernel", "polynomial_kernel", "sigmoid_kernel", "PAIRWISE_KERNEL_FUNCTIONS"]


def _dev(X, eng=None, dtype=None) -> torch.Tensor:
    """numpy / pandas / DeviceArray / torch -> 2-D float tensor on the engine's device."""
    eng = eng or get_engine()
    if isinstance(X, DeviceArray):
        X = X.tensor
    if hasattr(X, "compute") and not isinstance(X, torch.Tensor):  # a dask collection
        X = X.compute()
    if not isinstance(X, (torch.Tensor, np.ndarray)):
        X = _adopt_device_buffer(X)   # CuPy / Numba / JAX device arrays: zero-copy through DLPack
    if not isinstance(X, torch.Tensor):
        X = np.asarray(X)
        if X.dtype not in (np.float32, np.float64):
            X = X.astype(np.float32 if X.dtype in (np.int32, np.float16) else np.float64)
    elif X.dtype not in (torch.float32, torch.float64):
        X = X.to(torch.float32 if X.dtype in (torch.int32, torch.float16, torch.bfloat16) else torch.float64)
    t = eng.to_device(X, dtype)
    if t.dim() != 2:
        raise ValueError(f"Expected 2D array, got {t.dim()}D array instead")
    return t


def _check_same_features(Xd: torch.Tensor, Yd: torch.Tensor) -> None:
    """The reference fails inside da.dot / sklearn with a ValueError when the feature counts differ."""
    if Xd.shape[1] != Yd.shape[1]:
        raise ValueError(
            "Incompatible dimension for X and Y matrices: X.shape[1] == %d while Y.shape[1] == %d"
            % (Xd.shape[1], Yd.shape[1])
        )


def pairwise_distances_argmin_min(X, Y, axis: int = 1, metric="euclidean", batch_size: Optional[int] = None,
                                  metric_kwargs: Optional[Dict[str, Any]] = None):
    """reference dask_ml/metrics/pairwise.py:19-47.

    Returns ``(argmins int64, mins float64)`` like the reference (:36-44).  ``metric_kwargs=
    {"squared": True}`` gives squared distances (what the Lloyd loop asks for, reference
    dask_ml/cluster/k_means.py:576-578).  Computed by K1 (+ the fp64 re-decision for float32 X).
    """
    if batch_size is not None:
        msg = "'batch_size' is deprecated. Use sklearn.config_context instead.'"
        warnings.warn(msg, FutureWarning)
    if axis != 1:
        raise ValueError("only axis=1 (argmin over the rows of Y) is on the k-means path")
    squared = bool((metric_kwargs or {}).get("squared", False))
    if metric == "sqeuclidean":
        squared = True
    elif metric != "euclidean":
        raise ValueError(f"metric={metric!r}: only 'euclidean' is on the k-means path")
    eng = get_engine()
    Xd = _dev(X, eng)
    Yd = _dev(Y, eng, torch.float64)
    _check_same_features(Xd, Yd)
    pack = eng.pack(Yd)
    labels, mind2 = eng.assign(Xd, pack, want_mind2=True)
    mins = mind2.to(torch.float64)
    if not squared:
        mins = mins.sqrt_()
    return DeviceArray(labels.to(torch.int64)), DeviceArray(mins)


def pairwise_distances(X, Y, metric="euclidean", n_jobs: Optional[int] = None, **kwargs):
    """reference dask_ml/metrics/pairwise.py:50-67 (``Y`` must be an in-memory array, :57-58)."""
    if type(Y).__module__.startswith("dask."):
        raise TypeError("`Y` must be a numpy array")
    if metric not in ("euclidean", "sqeuclidean"):
        raise ValueError(f"metric={metric!r}: only 'euclidean' is on the k-means path")
    eng = get_engine()
    Xd = _dev(X, eng)
    Yd = _dev(Y, eng)
    _check_same_features(Xd, Yd)
    if Yd.dtype != Xd.dtype:
        Xd, Yd = Xd.to(torch.float64), Yd.to(torch.float64)
    squared = metric == "sqeuclidean" or bool(kwargs.get("squared", False))
    return DeviceArray(eng.cdist(Xd, Yd, squared=squared))


def euclidean_distances(X, Y=None, Y_norm_squared=None, squared: bool = False, X_norm_squared=None):
    """reference dask_ml/metrics/pairwise.py:70-105: ``sqrt(max(-2 X.Y^T + XX + YY, 0))`` in the
    input precision, optional precomputed norms with the reference's shape checks (:80-96)."""
    eng = get_engine()
    Xd = _dev(X, eng)
    same = Y is None or Y is X
    Yd = Xd if same else _dev(Y, eng)
    _check_same_features(Xd, Yd)
    if Yd.dtype != Xd.dtype:  # numpy promotion of -2 * dot(X, Y.T) + XX + YY (:101)
        Xd, Yd = Xd.to(torch.float64), Yd.to(torch.float64)
    xn = yn = None
    if X_norm_squared is not None:
        XX = np.asarray(X_norm_squared.compute() if hasattr(X_norm_squared, "compute") else X_norm_squared)
        if XX.shape == (1, Xd.shape[0]):
            XX = XX.T
        elif XX.shape != (Xd.shape[0], 1):
            raise ValueError("Incompatible dimensions for X and X_norm_squared")
        xn = eng.to_device(XX.reshape(-1), Xd.dtype)
    if same:
        yn = xn
    elif Y_norm_squared is not None:
        YY = np.asarray(Y_norm_squared.compute() if hasattr(Y_norm_squared, "compute") else Y_norm_squared)
        if YY.ndim < 2:
            YY = YY[:, np.newaxis]
        if YY.shape != (1, Yd.shape[0]):
            raise ValueError("Incompatible dimensions for Y and Y_norm_squared")
        yn = eng.to_device(YY.reshape(-1), Xd.dtype)
    return DeviceArray(eng.cdist(Xd, Yd, squared=squared, xnorm2=xn, ynorm2=yn))


# ----------------
# Kernel functions (reference dask_ml/metrics/pairwise.py:135-226) -- the affinity blocks of
# SpectralClustering's Nystrom embedding, the only in-tree caller of KMeans
# ----------------
def _kernel(X, Y, kind, gamma=None, coef0=1.0, degree=3):
    eng = get_engine()
    Xd = _dev(X, eng)
    Yd = Xd if Y is None else _dev(Y, eng)
    _check_same_features(Xd, Yd)  # check_pairwise_arrays (:108-128)
    if Yd.dtype != Xd.dtype:
        Xd, Yd = Xd.to(torch.float64), Yd.to(torch.float64)
    if gamma is None:
        gamma = 1.0 / Xd.shape[1]
    return DeviceArray(eng.pairwise_kernel(Xd, Yd, kind, gamma=gamma, coef0=coef0, degree=degree))


def linear_kernel(X, Y=None):
    """reference :135-138 -- ``dot(X, Y.T)``."""
    return _kernel(X, Y, "linear")


def rbf_kernel(X, Y=None, gamma=None):
    """reference :141-151 -- ``exp(-gamma * euclidean_distances(X, Y, squared=True))``, gamma default 1/d."""
    return _kernel(X, Y, "rbf", gamma=gamma)


def polynomial_kernel(X, Y=None, degree=3, gamma=None, coef0=1):
    """reference :154-167 -- ``(gamma * dot(X, Y.T) + coef0) ** degree``."""
    return _kernel(X, Y, "polynomial", gamma=gamma, coef0=coef0, degree=degree)


def sigmoid_kernel(X, Y=None, gamma=None, coef0=1):
    """reference :170-185 -- ``tanh(gamma * dot(X, Y.T) + coef0)``."""
    return _kernel(X, Y, "sigmoid", gamma=gamma, coef0=coef0)


PAIRWISE_KERNEL_FUNCTIONS = {
    "rbf": rbf_kernel,
    "linear": linear_kernel,
    "polynomial": polynomial_kernel,
    "sigmoid": sigmoid_kernel,
}


def pairwise_kernels(X, Y=None, metric="linear", filter_params=False, n_jobs=1, **kwds):
    """reference :199-226 (same dispatch, same errors)."""
    from sklearn.gaussian_process.kernels import Kernel as GPKernel
    from sklearn.metrics.pairwise import KERNEL_PARAMS

    if metric == "precomputed":
        Xd = _dev(X)
        Yn = Xd.shape[0] if Y is None else _dev(Y).shape[0]
        if Xd.shape[1] != Yn:
            raise ValueError(
                "Precomputed metric requires shape (n_queries, n_indexed). Got (%d, %d) for %d indexed."
                % (Xd.shape[0], Xd.shape[1], Yn)
            )
        return DeviceArray(Xd)
    elif isinstance(metric, GPKernel):
        raise NotImplementedError()
    elif isinstance(metric, str) and metric in PAIRWISE_KERNEL_FUNCTIONS:
        if filter_params:
            kwds = dict((k, kwds[k]) for k in kwds if k in KERNEL_PARAMS[metric])
        func = PAIRWISE_KERNEL_FUNCTIONS[metric]
    elif callable(metric):
        raise NotImplementedError()
    else:
        raise ValueError("Unknown kernel %r" % metric)
    return func(X, Y, **kwds)
