"""Unsupervised learning: the k-means hot path and its only in-tree caller, the Nystrom
SpectralClustering (reference dask_ml/cluster/__init__.py exports exactly these two)."""
from .k_means import KMeans, k_means  # noqa: F401
from .spectral import SpectralClustering  # noqa: F401

__all__ = ["KMeans", "k_means", "SpectralClustering"]
