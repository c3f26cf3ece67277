"""Eager numpy-backed stand-in for the part of ``dask`` that dask-ml's k-means path imports.

TEST INFRASTRUCTURE ONLY (see ../README.md).  ``delayed`` runs the function at call time; ``compute`` unwraps.
"""
import contextlib

__version__ = "2025.5.0"   # what the reference pins as its floor (pyproject.toml: dask >= 2025.1.0)

from ._delayed import Delayed, delayed  # noqa: E402,F401


def is_dask_collection(x) -> bool:
    from .array import Array

    return isinstance(x, (Array, Delayed))


def _unwrap(x):
    from .array import Array

    if isinstance(x, Array):
        return x.compute()
    if isinstance(x, Delayed):
        return x.compute()
    if isinstance(x, tuple):
        return tuple(_unwrap(v) for v in x)
    if isinstance(x, list):
        return [_unwrap(v) for v in x]
    if isinstance(x, dict):
        return {k: _unwrap(v) for k, v in x.items()}
    return x


def compute(*args, **kwargs):
    """``dask.compute(a, b)`` -> ``(a_value, b_value)``; python containers are traversed."""
    return tuple(_unwrap(a) for a in args)


def persist(*args, **kwargs):
    return args


@contextlib.contextmanager
def annotate(**kwargs):
    yield


from . import utils  # noqa: E402,F401
from . import array  # noqa: E402,F401
from . import dataframe  # noqa: E402,F401
