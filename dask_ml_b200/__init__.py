"""dask_ml_b200 -- the k-means hot path of dask-ml (Lloyd + k-means||) on B200.

Python host code over hand-written sm_100a CUDA kernels behind a flat C ABI
(``include/b200km.h`` / ``libb200km.so``).  Drop-in for ``dask_ml.cluster.KMeans``,
``dask_ml.metrics.pairwise_distances_argmin_min`` / ``pairwise_distances`` /
``euclidean_distances`` and ``dask_ml.utils.row_norms`` on that path; nothing else of dask-ml.

There is no CPU fallback: without the library or without a CUDA device every call raises.
"""
__version__ = "0.1.0"

from . import _lib  # noqa: F401  (does not load the .so until first use)

__all__ = ["cluster", "metrics", "utils", "datasets"]
