"""``dask.array`` evaluated eagerly on numpy arrays that remember their chunking.

TEST INFRASTRUCTURE ONLY (see ../../README.md).  An ``Array`` holds the whole ndarray plus a ``chunks`` tuple;
anything whose result can depend on the block structure (``map_blocks``, ``blockwise``, ``to_delayed``, ``dot``,
``sum``) is executed block by block in block order, everything elementwise on the whole array (IEEE elementwise
results do not depend on how the array is cut).  Only the surface that dask-ml's k-means path touches is here:

    Array: shape dtype ndim chunks numblocks blocks T astype rechunk compute persist map_blocks to_delayed
           sum min max any all __getitem__ __len__ arithmetic / comparison operators, numpy ufunc protocol
    from_array from_delayed asarray concatenate blockwise map_blocks dot bincount where maximum minimum sqrt exp
    tanh isnull isnan isinf compute random.RandomState utils.assert_eq
"""
from __future__ import annotations

import itertools
import math
import operator

import numpy as np

from .._delayed import Delayed, _val

__all__ = ["Array", "from_array", "from_delayed", "asarray", "concatenate", "blockwise", "map_blocks", "dot",
           "bincount", "where", "maximum", "minimum", "sqrt", "exp", "tanh", "isnull", "isnan", "isinf", "compute",
           "normalize_chunks"]

_AUTO_LIMIT = 128 * 2**20   # dask's default array.chunk-size


def normalize_chunks(chunks, shape, dtype=np.float64):
    """int / tuple of ints (block sizes) / tuple of tuples / dict / 'auto'  ->  tuple of tuples of block lengths."""
    shape = tuple(int(s) for s in shape)
    if chunks == "auto" or chunks is None:
        if int(np.prod(shape, dtype=np.int64)) * np.dtype(dtype).itemsize > _AUTO_LIMIT:
            raise NotImplementedError("stub: chunks='auto' beyond one default-size chunk")
        return tuple((s,) for s in shape)
    if isinstance(chunks, (int, np.integer)):
        chunks = (int(chunks),) * len(shape)
    if isinstance(chunks, dict):
        raise NotImplementedError("stub: dict chunks are handled by Array.rechunk")
    chunks = tuple(chunks)
    if len(shape) == 1 and len(chunks) > 1 and all(isinstance(c, (int, np.integer)) for c in chunks):
        chunks = (chunks,)      # block lengths of a 1-d array given flat
    if len(chunks) != len(shape):
        raise ValueError("Chunks and shape must be of the same length/dimension. Got chunks=%r, shape=%r"
                         % (chunks, shape))
    out = []
    for c, s in zip(chunks, shape):
        if isinstance(c, (tuple, list)):
            c = tuple(int(v) if not (isinstance(v, float) and math.isnan(v)) else v for v in c)
            if not any(isinstance(v, float) for v in c) and sum(c) != s:
                raise ValueError("Chunks do not add up to shape. Got chunks=%r, shape=%r" % (chunks, shape))
            out.append(c)
        elif c == -1 or c is None:
            out.append((s,))
        else:
            c = int(c)
            if s == 0:
                out.append((0,))
            else:
                full, rem = divmod(s, c)
                out.append((c,) * full + ((rem,) if rem else ()))
    return tuple(out)


def _bounds(chunk_lens):
    b = [0]
    for c in chunk_lens:
        b.append(b[-1] + c)
    return b


class _Blocks:
    def __init__(self, arr):
        self._arr = arr

    def __getitem__(self, idx):
        if not isinstance(idx, tuple):
            idx = (idx,)
        a = self._arr
        sl, ch = [], []
        for dim in range(a.ndim):
            b = _bounds(a.chunks[dim])
            if dim < len(idx) and isinstance(idx[dim], (int, np.integer)):
                i = int(idx[dim])
                sl.append(slice(b[i], b[i + 1]))
                ch.append((a.chunks[dim][i],))
            else:
                sl.append(slice(None))
                ch.append(a.chunks[dim])
        return Array(a._a[tuple(sl)], tuple(ch), a._meta)

    def ravel(self):
        return [self[i] for i in itertools.product(*[range(n) for n in self._arr.numblocks])]


class Array:
    __array_priority__ = 11

    def __init__(self, a, chunks=None, meta=None):
        a = np.asarray(a)
        self._a = a
        # dask reports the DECLARED dtype (``map_blocks(dtype=float)``, ``from_delayed(..., dtype)``) and never casts
        # the blocks to it; reductions take their accumulator dtype from the declared one
        self._meta = np.dtype(meta) if meta is not None else a.dtype
        if chunks is None:
            chunks = tuple((s,) for s in a.shape)
        self._chunks = tuple(tuple(c) for c in chunks)
        assert len(self._chunks) == a.ndim, (self._chunks, a.shape)

    # -- metadata ----------------------------------------------------------------------------------------
    @property
    def shape(self):
        if any(isinstance(v, float) for c in self._chunks for v in c):   # unknown chunk sizes
            return tuple(float("nan") if any(isinstance(v, float) for v in c) else sum(c) for c in self._chunks)
        return self._a.shape

    dtype = property(lambda self: self._meta)
    ndim = property(lambda self: self._a.ndim)
    chunks = property(lambda self: self._chunks)
    numblocks = property(lambda self: tuple(len(c) for c in self._chunks))
    size = property(lambda self: self._a.size)
    nbytes = property(lambda self: self._a.nbytes)
    blocks = property(lambda self: _Blocks(self))

    def __len__(self):
        if self.ndim == 0:
            raise TypeError("len() of unsized object")
        n = self.shape[0]
        if isinstance(n, float):
            raise ValueError("Cannot call len() on object with unknown chunk size.")
        return n

    def __repr__(self):
        return "stub.dask.array<shape=%r, dtype=%s, chunks=%r>" % (self.shape, self.dtype, self._chunks)

    # -- evaluation ----------------------------------------------------------------------------------------
    def compute(self, **kwargs):
        return self._a[()] if self._a.ndim == 0 else self._a

    def persist(self, **kwargs):
        return self

    def __array__(self, dtype=None, copy=None):
        return self._a if dtype is None else self._a.astype(dtype)

    def __bool__(self):
        return bool(self._a)

    def __float__(self):
        return float(self._a)

    def __int__(self):
        return int(self._a)

    def item(self):
        return self._a.item()

    # -- block access ----------------------------------------------------------------------------------------
    def _block_arrays(self):
        """The blocks, in C order of their block index, as ndarrays."""
        bounds = [_bounds(c) for c in self._chunks]
        out = []
        for idx in itertools.product(*[range(n) for n in self.numblocks]):
            out.append(self._a[tuple(slice(bounds[d][i], bounds[d][i + 1]) for d, i in enumerate(idx))])
        return out

    def to_delayed(self, optimize_graph=True):
        blocks = self._block_arrays()
        out = np.empty(len(blocks), dtype=object)
        for i, b in enumerate(blocks):
            out[i] = Delayed(b)
        return out.reshape(self.numblocks)

    def map_blocks(self, func, *args, **kwargs):
        return map_blocks(func, self, *args, **kwargs)

    def rechunk(self, chunks="auto", **kwargs):
        if isinstance(chunks, dict):
            new = list(self._chunks)
            for dim, c in chunks.items():
                new[dim] = normalize_chunks((c,), (self._a.shape[dim],))[0]
            return Array(self._a, tuple(new), self._meta)
        return Array(self._a, normalize_chunks(chunks, self._a.shape, self.dtype), self._meta)

    def astype(self, dtype, **kwargs):
        return Array(self._a.astype(dtype), self._chunks)

    @property
    def T(self):
        return Array(self._a.T, self._chunks[::-1], self._meta)

    def transpose(self, *axes):
        if axes:
            raise NotImplementedError
        return self.T

    def reshape(self, *shape):
        return Array(self._a.reshape(*shape), None, self._meta)

    # -- indexing ----------------------------------------------------------------------------------------
    def __getitem__(self, key):
        res = self._a[_val(key) if not isinstance(key, tuple) else tuple(_val(k) for k in key)]
        res = np.asarray(res)
        keys = key if isinstance(key, tuple) else (key,)
        # keep the chunking where a dimension is taken whole; anything else becomes one chunk
        ch, dim = [], 0
        for k in keys:
            if k is None:
                ch.append((1,))
            elif isinstance(k, slice):
                if k == slice(None):
                    ch.append(self._chunks[dim])
                else:
                    ch.append((len(range(*k.indices(self._a.shape[dim]))),))
                dim += 1
            elif isinstance(k, (int, np.integer)):
                dim += 1
            else:   # list / ndarray index
                kk = np.asarray(_val(k))
                ch.append((int(kk.sum()) if kk.dtype == bool else kk.size,))
                dim += 1
        ch.extend(self._chunks[dim:])
        if len(ch) != res.ndim:
            ch = [(s,) for s in res.shape]
        return Array(res, tuple(ch), self._meta)

    # -- reductions ----------------------------------------------------------------------------------------
    def _tree_reduce(self, chunk_fn, agg_fn, axis, split_every=4, dtype=None):
        """dask.array.reductions.reduction / _tree_reduce: ``chunk_fn(block, keepdims=True)`` on every block, then
        groups of ``n`` neighbouring partials per reduced axis are concatenated and reduced with ``agg_fn`` until one
        block is left (``n = max(int(split_every ** (1 / len(axes))), 2)``; config default split_every = 4).
        ``dtype`` (sum only) is passed to both stages, as dask does with the dtype derived from the DECLARED one."""
        kw = {} if dtype is None else {"dtype": dtype}
        meta = self._meta if dtype is None else np.dtype(dtype)
        if chunk_fn in (np.any, np.all):
            meta = np.dtype(bool)
        if axis is None:
            axes = tuple(range(self.ndim))
        elif isinstance(axis, (int, np.integer)):
            axes = (int(axis) % max(self.ndim, 1),)
        else:
            axes = tuple(int(a) % self.ndim for a in axis)
        if self.ndim == 0:
            return Array(chunk_fn(self._a, axis=None, keepdims=False, **kw), None, meta)
        nb = self.numblocks
        parts = np.empty(nb, dtype=object)
        for idx, blk in zip(itertools.product(*[range(n) for n in nb]), self._block_arrays()):
            parts[idx] = chunk_fn(blk, axis=axes, keepdims=True, **kw)
        n = max(int(split_every ** (1.0 / (len(axes) or 1))), 2)
        while any(parts.shape[a] > 1 for a in axes):
            new_shape = tuple(-(-s // n) if d in axes else s for d, s in enumerate(parts.shape))
            new = np.empty(new_shape, dtype=object)
            for idx in itertools.product(*[range(s) for s in new_shape]):
                sl = tuple(slice(i * n, (i + 1) * n) if d in axes else slice(i, i + 1) for d, i in enumerate(idx))
                group = parts[sl]
                cat = np.block(group.tolist()) if group.size > 1 else group.ravel()[0]
                new[idx] = agg_fn(cat, axis=axes, keepdims=True, **kw)
            parts = new
        full = np.block(parts.tolist()) if parts.size > 1 else parts.ravel()[0]
        out = full.reshape([s for d, s in enumerate(full.shape) if d not in axes])
        ch = tuple(c for d, c in enumerate(self._chunks) if d not in axes)
        return Array(out, ch, meta)

    def sum(self, axis=None, dtype=None, keepdims=False, split_every=None):
        if keepdims:
            raise NotImplementedError
        if dtype is None:   # dask: getattr(np.zeros(1, dtype=a.dtype).sum(), "dtype", object), a.dtype the declared one
            dtype = np.zeros(1, dtype=self._meta).sum().dtype
        return self._tree_reduce(np.sum, np.sum, axis, split_every or 4, dtype=dtype)

    def min(self, axis=None, keepdims=False, split_every=None):
        return self._tree_reduce(np.min, np.min, axis, split_every or 4)

    def max(self, axis=None, keepdims=False, split_every=None):
        return self._tree_reduce(np.max, np.max, axis, split_every or 4)

    def any(self, axis=None, keepdims=False, split_every=None):
        return self._tree_reduce(np.any, np.any, axis, split_every or 4)

    def all(self, axis=None, keepdims=False, split_every=None):
        return self._tree_reduce(np.all, np.all, axis, split_every or 4)

    def argmin(self, axis=None):
        return Array(np.argmin(self._a, axis=axis))

    def mean(self, axis=None):
        raise NotImplementedError("stub: mean is not on the k-means path")

    # -- elementwise ----------------------------------------------------------------------------------------
    def __array_ufunc__(self, ufunc, method, *inputs, **kwargs):
        if method != "__call__" or kwargs.get("out") is not None:
            return NotImplemented
        return _elemwise(ufunc, *inputs, **kwargs)

    def __neg__(self):
        return _elemwise(operator.neg, self)

    def __abs__(self):
        return _elemwise(abs, self)


def _result_chunks(res, operands):
    """Chunks of a broadcast elementwise result: per dimension, those of an operand that spans it."""
    ch = [None] * res.ndim
    for op in operands:
        if not isinstance(op, Array):
            continue
        off = res.ndim - op.ndim
        for d in range(op.ndim):
            if op._a.shape[d] == res.shape[off + d] and (ch[off + d] is None or len(op._chunks[d]) > len(ch[off + d])):
                ch[off + d] = op._chunks[d]
    return tuple(c if c is not None else (s,) for c, s in zip(ch, res.shape))


def _elemwise(fn, *operands, **kwargs):
    raw = [o._a if isinstance(o, Array) else _val(o) for o in operands]
    with np.errstate(all="ignore"):
        res = np.asarray(fn(*raw, **kwargs))
        # declared dtype of the result: the function applied to one-element stand-ins of the declared dtypes
        try:
            meta = np.asarray(fn(*[np.ones((1,) * o.ndim, dtype=o._meta) if isinstance(o, Array) else r
                                   for o, r in zip(operands, raw)], **kwargs)).dtype
        except Exception:
            meta = res.dtype
    return Array(res, _result_chunks(res, operands), meta)


def _binop(name, fn):
    def fwd(self, other):
        if isinstance(other, Delayed):
            return NotImplemented
        return _elemwise(fn, self, other)

    def rev(self, other):
        return _elemwise(fn, other, self)

    setattr(Array, "__%s__" % name, fwd)
    setattr(Array, "__r%s__" % name, rev)


for _n, _f in [("add", operator.add), ("sub", operator.sub), ("mul", operator.mul), ("truediv", operator.truediv),
               ("floordiv", operator.floordiv), ("pow", operator.pow), ("mod", operator.mod),
               ("and", operator.and_), ("or", operator.or_), ("xor", operator.xor)]:
    _binop(_n, _f)
for _n, _f in [("lt", operator.lt), ("le", operator.le), ("gt", operator.gt), ("ge", operator.ge),
               ("eq", operator.eq), ("ne", operator.ne)]:
    setattr(Array, "__%s__" % _n, (lambda f: lambda self, other: _elemwise(f, self, other))(_f))
Array.__hash__ = None


def sqrt(x):
    return _elemwise(np.sqrt, x)


def exp(x):
    return _elemwise(np.exp, x)


def tanh(x):
    return _elemwise(np.tanh, x)


def maximum(a, b):
    return _elemwise(np.maximum, a, b)


def minimum(a, b):
    return _elemwise(np.minimum, a, b)


def isnan(x):
    return _elemwise(np.isnan, x)


def isnull(x):
    import pandas as pd

    return _elemwise(pd.isnull, x)


def isinf(x):
    return _elemwise(np.isinf, x)


# -- construction ------------------------------------------------------------------------------------------
def from_array(x, chunks="auto", **kwargs):
    if isinstance(x, Array):
        return x.rechunk(chunks)
    x = np.asarray(x)
    return Array(x, normalize_chunks(chunks, x.shape, x.dtype))


def asarray(x, **kwargs):
    return x if isinstance(x, Array) else from_array(x)


def from_delayed(value, shape, dtype=None, meta=None, name=None):
    """One block holding the delayed value AS IT IS: dask does not cast to the declared dtype (the reference declares
    X.dtype for float64 block sums, k_means.py:594-596)."""
    v = np.asarray(_val(value))
    assert tuple(v.shape) == tuple(shape), (v.shape, shape)
    return Array(v, None, dtype)


def concatenate(seq, axis=0, allow_unknown_chunksizes=False):
    seq = [asarray(s) for s in seq]
    a = np.concatenate([s._a for s in seq], axis=axis)
    ch = list(seq[0]._chunks)
    ch[axis] = tuple(itertools.chain.from_iterable(s._chunks[axis] for s in seq))
    return Array(a, tuple(ch), np.result_type(*[s._meta for s in seq]))


def compute(*args, **kwargs):
    from .. import compute as _compute

    return _compute(*args, **kwargs)


# -- block-structured operations -----------------------------------------------------------------------------
def _row_blocks(x: Array):
    if any(n != 1 for n in x.numblocks[1:]):
        raise NotImplementedError("stub: block-wise functions need a single block along every axis but the first")
    b = _bounds(x._chunks[0])
    return [x._a[b[i]:b[i + 1]] for i in range(len(b) - 1)]


def map_blocks(func, *args, dtype=None, chunks=None, drop_axis=None, new_axis=None, name=None, token=None,
               meta=None, **kwargs):
    arrs = [a for a in args if isinstance(a, Array)]
    nblk = arrs[0].numblocks[0]
    per = [(_row_blocks(a) if isinstance(a, Array) else None) for a in args]
    for p in per:
        if p is not None and len(p) != nblk:
            raise ValueError("stub: map_blocks operands must have matching row blocks")
    outs = []
    for i in range(nblk):
        outs.append(np.asarray(func(*[(p[i] if p is not None else a) for p, a in zip(per, args)], **kwargs)))
    res = np.concatenate(outs, axis=0) if outs[0].ndim else np.asarray(outs)
    ch = (tuple(o.shape[0] for o in outs),) + tuple((s,) for s in res.shape[1:])
    return Array(res, ch, dtype)


def blockwise(func, out_ind, *args, name=None, token=None, dtype=None, adjust_chunks=None, new_axes=None,
              align_arrays=True, concatenate=None, meta=None, **kwargs):
    """Only the pattern the reference uses (k_means.py:582-593): every array operand is cut along the first output
    index only; ``func`` runs once per block row and the per-block results are stacked along that index
    (``adjust_chunks`` says each has its own length there).  (array, index) pairs come from ``partition(2, args)``,
    which silently DROPS a trailing unpaired argument -- the reference passes one."""
    pairs = [(args[i], args[i + 1]) for i in range(0, len(args) - 1, 2)]
    arrays = [(a, ind) for a, ind in pairs if ind is not None]
    first = out_ind[0]
    for a, ind in arrays:
        if not isinstance(a, Array) or ind[0] != first:
            raise NotImplementedError("stub: blockwise operands must be arrays indexed by the first output index")
    nblk = arrays[0][0].numblocks[0]
    per = [(_row_blocks(a) if ind is not None else None) for a, ind in pairs]
    outs = []
    for i in range(nblk):
        outs.append(np.asarray(func(*[(p[i] if p is not None else a) for p, (a, _) in zip(per, pairs)], **kwargs)))
    res = np.concatenate(outs, axis=0)
    ch = (tuple(o.shape[0] for o in outs),) + tuple((s,) for s in res.shape[1:])
    return Array(res, ch, dtype)


def dot(a, b):
    """tensordot over the shared axis, one ``np.dot`` per (row block of a, column block of b)."""
    a, b = asarray(a), asarray(b)
    if a.ndim != 2 or b.ndim != 2 or a.numblocks[1] != 1 or b.numblocks[0] != 1:
        raise NotImplementedError("stub: dot needs an unchunked contraction axis")
    rb, cb = _bounds(a._chunks[0]), _bounds(b._chunks[1])
    out = np.empty((a._a.shape[0], b._a.shape[1]), dtype=np.result_type(a.dtype, b.dtype))
    for i in range(len(rb) - 1):
        for j in range(len(cb) - 1):
            out[rb[i]:rb[i + 1], cb[j]:cb[j + 1]] = np.dot(a._a[rb[i]:rb[i + 1]], b._a[:, cb[j]:cb[j + 1]])
    return Array(out, (a._chunks[0], b._chunks[1]))


def bincount(x, weights=None, minlength=0, split_every=None):
    if weights is not None:
        raise NotImplementedError
    x = asarray(x)
    total = None
    for blk in _row_blocks(x):      # per-chunk counts, then summed: integers, so exact in any order
        c = np.bincount(blk, minlength=minlength)
        total = c if total is None else total + c
    return Array(total)


def where(condition, x=None, y=None):
    if (x is None) != (y is None):
        raise ValueError("either both or neither of x and y should be given")
    if x is None:
        condition = asarray(condition)
        idx = np.nonzero(condition._a)
        # dask returns arrays of unknown length (nan chunks); only compute() is meaningful on them
        return tuple(Array(i) for i in idx)
    return _elemwise(np.where, condition, x, y)


from . import random  # noqa: E402,F401
from . import utils  # noqa: E402,F401
