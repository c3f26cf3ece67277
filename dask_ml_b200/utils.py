"""Helpers on the k-means path: ``row_norms``, ``check_array``, ``_timer`` / ``_timed`` and the
lazy-array stand-in ``DeviceArray``.

Mirrors the parts of reference dask_ml/utils.py that the path touches (row_norms :83-88,
check_array :123-214, _timer/_timed :515-559); nothing else from that module is in scope.
"""
from __future__ import annotations

import contextlib
import os
import functools
import logging
from timeit import default_timer as tic

import numpy as np
import torch
from sklearn.utils import validation as sk_validation

logger = logging.getLogger(__name__)

__all__ = ["DeviceArray", "row_norms", "check_array", "_timer", "_timed", "as_numpy"]


class DeviceArray:
    """What the reference returns as a lazy dask array (``labels_``, ``predict``, ``transform``,
    ``pairwise_distances*``) is here a device-resident tensor with the small dask-like surface
    callers use: ``.compute()``, ``.dtype``, ``.shape``, ``.astype``, ``np.asarray(...)``."""

    __array_priority__ = 100

    def __init__(self, tensor: torch.Tensor):
        self._t = tensor

    @property
    def tensor(self) -> torch.Tensor:
        return self._t

    def compute(self, **_ignored) -> np.ndarray:
        return self._t.detach().cpu().numpy()

    def persist(self):
        return self

    def __array__(self, dtype=None, copy=None):
        a = self.compute()
        return a.astype(dtype) if dtype is not None else a

    @property
    def dtype(self):
        return np.dtype(str(self._t.dtype).replace("torch.", ""))

    @property
    def shape(self):
        return tuple(self._t.shape)

    @property
    def ndim(self):
        return self._t.dim()

    @property
    def chunks(self):
        return tuple((s,) for s in self.shape)

    def __len__(self):
        return self._t.shape[0]

    def __getitem__(self, item):
        out = self._t[item]
        return DeviceArray(out) if isinstance(out, torch.Tensor) and out.dim() > 0 else out.item()

    def astype(self, dtype):
        return DeviceArray(self._t.to(getattr(torch, np.dtype(dtype).name)))

    def sum(self):
        from ._engine import get_engine

        if self._t.dtype in (torch.float32, torch.float64):
            return DeviceArray(get_engine(self._t.device).reduce_sum(self._t.reshape(-1)).reshape(()))
        return DeviceArray(self._t.sum())

    def item(self):
        return self._t.item()

    def __repr__(self):
        return f"DeviceArray(shape={self.shape}, dtype={self.dtype}, device={self._t.device})"


def as_numpy(a) -> np.ndarray:
    if isinstance(a, DeviceArray):
        return a.compute()
    if isinstance(a, torch.Tensor):
        return a.detach().cpu().numpy()
    return np.asarray(a)


def row_norms(X, squared=False):
    """reference dask_ml/utils.py:83-88.  ndarray in -> sklearn on the host exactly as the
    reference; device input -> b200km_row_norms (DeviceArray out)."""
    if isinstance(X, np.ndarray):
        from sklearn.utils import extmath as skm

        return skm.row_norms(X, squared=squared)
    from ._engine import get_engine

    t = X.tensor if isinstance(X, DeviceArray) else X
    return DeviceArray(get_engine(t.device).row_norms(t, squared=squared))


def _adopt_device_buffer(array):
    """Zero-copy view of a foreign DEVICE array (CuPy, Numba, JAX, another framework's tensor) as a torch tensor:
    ``__dlpack__`` when the producer says the buffer lives on a CUDA device, else ``__cuda_array_interface__``.
    Host-side DLPack producers and everything else are returned unchanged (sklearn's validation handles them).
    SURVEY section 8(f).3: the reference takes such blocks only inside dask arrays."""
    dev = getattr(array, "__dlpack_device__", None)
    if dev is not None and hasattr(array, "__dlpack__"):
        try:
            kind = int(dev()[0])
        except Exception:  # pragma: no cover - a producer that cannot report its device
            kind = -1
        if kind in (2, 13):  # kDLCUDA, kDLCUDAManaged
            return torch.from_dlpack(array)
        return array
    if hasattr(array, "__cuda_array_interface__"):
        return torch.as_tensor(array, device="cuda")
    return array


def check_array(array, *args, accept_dask_array=True, accept_dask_dataframe=False, accept_unknown_chunks=False,
                accept_multiple_blocks=False, preserve_pandas_dataframe=False, remove_zero_chunks=True, **kwargs):
    """reference dask_ml/utils.py:123-214.

    Device tensors / DeviceArrays play the role of dask arrays: they are validated through a
    small all-ones sample of the same dtype and trailing shape (as the reference does for
    ``da.Array``, :187-197) and returned unchanged; everything else goes to sklearn's
    ``check_array``.  dask collections are accepted when dask is importable.
    """
    if not isinstance(array, (torch.Tensor, DeviceArray, np.ndarray)):
        array = _adopt_device_buffer(array)
    if isinstance(array, (torch.Tensor, DeviceArray)):
        t = array.tensor if isinstance(array, DeviceArray) else array
        shape = tuple(t.shape)
        if len(shape) == 2:
            shape = (min(10, shape[0]), shape[1])
        elif len(shape) == 1:
            shape = (min(10, shape[0]),)
        sample = np.ones(shape=shape, dtype=str(t.dtype).replace("torch.", ""))
        sk_validation.check_array(sample, *args, **kwargs)
        return array
    mod = type(array).__module__
    if mod.startswith("dask."):  # optional: only when the caller already has dask
        import dask.array as da  # noqa: F401

        if type(array).__name__ == "DataFrame":
            if not accept_dask_dataframe:
                raise TypeError(
                    "This estimator does not support dask dataframes. This might be resolved with one of\n\n"
                    "    1. ddf.to_dask_array(lengths=True)\n"
                    "    2. ddf.to_dask_array()  # may cause other issues because of unknown chunk sizes"
                )
            return array
        if not accept_dask_array:
            raise TypeError
        if not accept_unknown_chunks and np.isnan(array.shape[0]):
            raise TypeError(
                "Cannot operate on Dask array with unknown chunk sizes. Use the following the compute chunk sizes:\n\n"
                "   >>> X.compute_chunk_sizes()  # if Dask.Array\n"
                "   >>> ddf.to_dask_array(lengths=True)  # if Dask.Dataframe"
            )
        if not accept_multiple_blocks and array.ndim > 1 and len(array.chunks[1]) > 1:
            raise TypeError(
                "Chunking is only allowed on the first axis. Use 'array.rechunk({1: array.shape[1]})' to "
                "rechunk to a single block along the second axis."
            )
        return array
    try:
        import pandas as pd

        if isinstance(array, pd.DataFrame) and preserve_pandas_dataframe:
            return array
    except ImportError:  # pragma: no cover
        pass
    return sk_validation.check_array(array, *args, **kwargs)


def _format(seconds: float) -> str:
    return "%.2f s" % seconds if seconds >= 1 else "%.2f ms" % (seconds * 1e3)


_NVTX = os.environ.get("B200KM_NVTX", "") not in ("", "0")


@contextlib.contextmanager
def _timer(name, _logger=None, level="info"):
    """reference dask_ml/utils.py:515-536 -- log start / finish + wall time.  With ``B200KM_NVTX=1`` every timed
    section (a Lloyd iteration, a k-means|| round, an initialiser) is also an NVTX range, so the kernels of one
    iteration group together on a profiler's timeline."""
    start = tic()
    _logger = _logger or logger
    _logger.info("Starting %s", name)
    pushed = False
    if _NVTX:
        try:
            torch.cuda.nvtx.range_push(str(name))
            pushed = True
        except Exception:  # pragma: no cover - NVTX is optional
            pushed = False
    try:
        yield
    finally:
        if pushed:
            torch.cuda.nvtx.range_pop()
    getattr(_logger, level)("Finished %s in %s", name, _format(tic() - start))


def _timed(_logger=None, level="info"):
    """reference dask_ml/utils.py:539-559 -- decorator form of ``_timer``."""

    def wrap(func):
        @functools.wraps(func)
        def inner(*args, **kwargs):
            with _timer(func.__name__, _logger=_logger, level=level):
                return func(*args, **kwargs)

        return inner

    return wrap
