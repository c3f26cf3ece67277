"""Imports the UNMODIFIED reference modules of the k-means path and runs them on the eager dask stand-in.

TEST INFRASTRUCTURE ONLY: used by ``tests/test_reference_pin.py`` and ``tests/golden/make_golden.py`` in the build
container.  The reference tree (``/root/reference``) does not travel to the GPU box; there the committed fixtures
under ``tests/golden/`` (generated HERE from the reference through this loader) carry the pin.

How: ``oracle/dask_stub`` provides ``dask`` / ``dask.array`` / ``dask.dataframe`` / ``distributed`` (see its README);
a synthetic ``dask_ml`` package (and synthetic ``dask_ml.metrics`` / ``dask_ml.cluster`` sub-packages) is registered
with ``__path__`` pointing INTO the reference tree, so that the reference's own files

    dask_ml/_typing.py  _utils.py  _compat.py  utils.py  metrics/pairwise.py  cluster/_compat.py  cluster/k_means.py

are executed from where they lie, byte for byte, while the package ``__init__`` files -- which pull in the whole of
dask-ml (model selection, GLMs, ...) -- are skipped.  Nothing of the reference is copied into this repository.
"""
from __future__ import annotations

import importlib
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("B200KM_REFERENCE_ROOT", "/root/reference")
_STUB = os.path.join(os.path.dirname(os.path.abspath(__file__)), "dask_stub")
_loaded = None


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "dask_ml", "cluster", "k_means.py"))


def _synthetic_package(name: str, path: str) -> types.ModuleType:
    mod = types.ModuleType(name)
    mod.__path__ = [path]
    mod.__package__ = name
    sys.modules[name] = mod
    return mod


def load():
    """-> namespace with ``k_means`` (reference dask_ml/cluster/k_means.py), ``pairwise`` (metrics/pairwise.py),
    ``utils`` (utils.py), ``da`` (the stub's dask.array) and ``dask``."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not available():
        raise RuntimeError("reference tree not found at %s" % REFERENCE_ROOT)
    for name in ("dask", "distributed", "dask_ml"):
        if name in sys.modules and not getattr(sys.modules[name], "__file__", "").startswith(_STUB) \
                and name != "dask_ml":
            raise RuntimeError("a real %s is already imported; the stub must not shadow it" % name)
    sys.path.insert(0, _STUB)
    try:
        import dask  # noqa: F401  (the stub)
        import dask.array as da
        import distributed  # noqa: F401

        base = os.path.join(REFERENCE_ROOT, "dask_ml")
        _synthetic_package("dask_ml", base)
        metrics_pkg = _synthetic_package("dask_ml.metrics", os.path.join(base, "metrics"))
        _synthetic_package("dask_ml.cluster", os.path.join(base, "cluster"))
        utils = importlib.import_module("dask_ml.utils")
        pairwise = importlib.import_module("dask_ml.metrics.pairwise")
        # what ``dask_ml/metrics/__init__.py`` re-exports and k_means.py imports from ``..metrics``
        for fn in ("euclidean_distances", "pairwise_distances", "pairwise_distances_argmin_min"):
            setattr(metrics_pkg, fn, getattr(pairwise, fn))
        k_means = importlib.import_module("dask_ml.cluster.k_means")
    finally:
        sys.path.remove(_STUB)
    for mod in (utils, pairwise, k_means):
        assert os.path.realpath(mod.__file__).startswith(os.path.realpath(REFERENCE_ROOT)), mod.__file__
    _loaded = types.SimpleNamespace(k_means=k_means, pairwise=pairwise, utils=utils, da=da, dask=sys.modules["dask"])
    return _loaded
