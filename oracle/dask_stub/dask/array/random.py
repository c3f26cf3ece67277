"""``dask.array.random.RandomState``: a parent MT19937 stream that hands every chunk of every call its own seed.

ASSUMPTION (restated from dask/array/random.py and dask/utils.py as of dask 2025.x; the source is not on this box):
``RandomState(seed)`` keeps ``np.random.RandomState(seed)``; each ``uniform`` / ``randint`` / ``choice`` call draws
``random_state_data(nchunks, parent)`` -- 624 uint32 words per chunk from ``parent.bytes`` -- and chunk i is
``np.random.RandomState(words_i).<method>(*args, size=chunk_shape)``.  So the draws depend on the chunk layout, and
every call advances the parent stream.  ``choice(replace=False)`` refuses more than one chunk.
"""
import itertools

import numpy as np

from ..utils import random_state_data
from . import Array, normalize_chunks


class RandomState:
    def __init__(self, seed=None, RandomState=None):
        self._numpy_state = np.random.RandomState(seed)
        self._RandomState = RandomState or np.random.RandomState

    def seed(self, seed=None):
        self._numpy_state.seed(seed)

    def _wrap(self, funcname, *args, size=None, chunks="auto", **kwargs):
        if size is not None and not isinstance(size, (tuple, list)):
            size = (size,)
        size = tuple(size) if size is not None else ()
        chunks = normalize_chunks(chunks, size, dtype=kwargs.get("dtype", np.float64))
        sizes = list(itertools.product(*chunks))
        seeds = random_state_data(len(sizes), self._numpy_state)
        blocks = np.empty(tuple(len(c) for c in chunks), dtype=object)
        for idx, seed, sz in zip(itertools.product(*[range(len(c)) for c in chunks]), seeds, sizes):
            state = self._RandomState(seed)
            blocks[idx] = np.asarray(getattr(state, funcname)(*args, size=sz, **kwargs))
        full = np.block(blocks.tolist()) if blocks.ndim else blocks[()]
        return Array(full, chunks)

    def uniform(self, low=0.0, high=1.0, size=None, chunks="auto", **kwargs):
        return self._wrap("uniform", low, high, size=size, chunks=chunks, **kwargs)

    def random_sample(self, size=None, chunks="auto", **kwargs):
        return self._wrap("random_sample", size=size, chunks=chunks, **kwargs)

    def normal(self, loc=0.0, scale=1.0, size=None, chunks="auto", **kwargs):
        return self._wrap("normal", loc, scale, size=size, chunks=chunks, **kwargs)

    def randint(self, low, high=None, size=None, chunks="auto", **kwargs):
        return self._wrap("randint", low, high, size=size, chunks=chunks, **kwargs)

    def choice(self, a, size=None, replace=True, p=None, chunks="auto"):
        if isinstance(a, Array):
            a = a.compute()
        if not isinstance(a, (int, np.integer)):
            a = np.asarray(a)
        if size is None:
            size = ()
        elif not isinstance(size, (tuple, list)):
            size = (size,)
        chunks = normalize_chunks(chunks, size, dtype=np.float64)
        if not replace and len(chunks[0]) > 1:
            raise NotImplementedError("replace=False is not currently supported for dask.array.choice with multi-chunk "
                                      "output arrays")
        sizes = list(itertools.product(*chunks))
        seeds = random_state_data(len(sizes), self._numpy_state)
        outs = [np.random.RandomState(seed).choice(a, size=sz, replace=replace, p=p) for seed, sz in zip(seeds, sizes)]
        return Array(np.concatenate([np.atleast_1d(o) for o in outs]) if size else outs[0], chunks)
