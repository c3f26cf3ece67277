"""Synthetic blobs generated ON THE DEVICE.

Mirrors the signature and the distribution of reference dask_ml/datasets.py:78-205 ``make_blobs``: isotropic
Gaussian blobs, equal expected points per centre, rows in random class order.  It is NOT the reference's bit
stream (that one comes from one ``sklearn.datasets.make_blobs`` call per row block, seeded with the block index,
:180-191); here every row is a pure function of (random_state, global row index), so any sharding over GPUs
generates the same array, and N = 1e9 rows never touch the host.
"""
from __future__ import annotations

import numbers

import numpy as np
import torch

from ._engine import get_engine
from .utils import DeviceArray

__all__ = ["make_blobs"]


def make_blobs(n_samples=100, n_features=2, centers=None, cluster_std=1.0, center_box=(-10.0, 10.0), shuffle=True,
               random_state=None, chunks=None, dtype="f8", row_offset=0):
    """Generate isotropic Gaussian blobs for clustering (reference dask_ml/datasets.py:78-205).

    Parameters as in the reference; ``chunks`` is accepted and ignored (there is one device-resident block),
    ``shuffle`` is always effectively True (the class of a row is a hash of its index).  Extra: ``dtype``
    ('f8' like the reference, or 'f4'), ``row_offset`` (global index of the first row, for a rank that
    generates only its shard).

    Returns
    -------
    X : DeviceArray (n_samples, n_features)
    y : DeviceArray (n_samples,) int64 -- cluster membership of each sample
    """
    eng = get_engine()
    if centers is None:
        centers = 3  # reference :159-161
    if isinstance(random_state, np.random.RandomState):
        rs = random_state
    else:
        rs = np.random.RandomState(random_state)
    if isinstance(centers, numbers.Integral):
        # sklearn.datasets.make_blobs draws the centres first, uniformly in center_box (the reference then
        # re-estimates them as block-0 class means, :171-178 -- an artefact not reproduced)
        centers = rs.uniform(center_box[0], center_box[1], size=(int(centers), n_features))
    centers = np.asarray(centers, dtype=np.float64)
    if centers.ndim != 2 or centers.shape[1] != n_features:
        n_features = centers.shape[1]
    seed = int(rs.randint(0, 2**31 - 1))
    tdt = torch.float64 if np.dtype(dtype) == np.float64 else torch.float32
    y = torch.empty(int(n_samples), dtype=torch.int32, device=eng.device)
    X = eng.make_blobs(int(n_samples), int(n_features), torch.from_numpy(centers), float(cluster_std), seed=seed,
                       row_offset=int(row_offset), dtype=tdt, y=y)
    return DeviceArray(X), DeviceArray(y.to(torch.int64))
