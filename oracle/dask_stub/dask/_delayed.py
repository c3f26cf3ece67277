"""``dask.delayed`` evaluated eagerly.  ``sum(list_of_delayed)`` is Python's left fold starting at 0, exactly the
chain the reference builds with ``sum(r.to_delayed().flatten())`` (dask_ml/cluster/k_means.py:595-597)."""


class Delayed:
    def __init__(self, value, nout=None):
        self._value = value
        self._nout = nout

    def compute(self, **kwargs):
        return self._value

    def __iter__(self):
        if self._nout is None:
            raise TypeError("Delayed objects of unspecified length are not iterable")
        for i in range(self._nout):
            yield Delayed(self._value[i])

    def __len__(self):
        if self._nout is None:
            raise TypeError("Delayed objects of unspecified length have no len()")
        return self._nout

    def __add__(self, other):
        return Delayed(self._value + _val(other))

    def __radd__(self, other):
        return Delayed(_val(other) + self._value)

    def __getitem__(self, i):
        return Delayed(self._value[i])


def _val(x):
    from .array import Array

    if isinstance(x, Delayed):
        return x._value
    if isinstance(x, Array):
        return x.compute()
    if isinstance(x, (list, tuple)):
        return type(x)(_val(v) for v in x)
    return x


class _DelayedFunc:
    def __init__(self, func, nout):
        self._func = func
        self._nout = nout

    def __call__(self, *args, **kwargs):
        args = [_val(a) for a in args]
        kwargs = {k: _val(v) for k, v in kwargs.items()}
        return Delayed(self._func(*args, **kwargs), self._nout)


def delayed(obj, name=None, pure=None, nout=None, traverse=True):
    if callable(obj):
        return _DelayedFunc(obj, nout)
    return Delayed(_val(obj), nout)
