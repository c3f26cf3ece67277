"""``dask.utils``: the two helpers the path needs."""
import numpy as np


def derived_from(original_klass, version=None, ua_args=None, skipblocks=0, inconsistencies=None):
    """Decorator that, in dask, copies the docstring of ``original_klass.<name>``."""

    def wrapper(method):
        try:
            method.__doc__ = getattr(original_klass, method.__name__).__doc__
        except AttributeError:
            pass
        return method

    return wrapper


def random_state_data(n, random_state=None):
    """``n`` seeds of 624 uint32 words each, drawn from ``random_state``'s byte stream -- how dask gives every chunk
    of a random array its own MT19937 stream (ASSUMPTION: restated from dask/utils.py, not verifiable here)."""
    if not all(hasattr(random_state, attr) for attr in ["normal", "beta", "bytes", "uniform"]):
        random_state = np.random.RandomState(random_state)
    random_data = random_state.bytes(624 * n * 4)
    out = list(np.frombuffer(random_data, dtype="<u4").reshape((n, -1)))
    assert len(out) == n
    return out
