"""``dask.array.utils.assert_eq``: value comparison after computing both sides."""
import numpy as np


def assert_eq(a, b, check_shape=True, check_dtype=True, **kwargs):
    a = a.compute() if hasattr(a, "compute") else a
    b = b.compute() if hasattr(b, "compute") else b
    a, b = np.asarray(a), np.asarray(b)
    if check_shape:
        assert a.shape == b.shape, (a.shape, b.shape)
    if check_dtype:
        assert a.dtype == b.dtype, (a.dtype, b.dtype)
    np.testing.assert_allclose(a, b, **{k: v for k, v in kwargs.items() if k in ("rtol", "atol", "equal_nan")})
    return True
