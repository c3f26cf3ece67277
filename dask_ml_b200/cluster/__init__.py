"""Unsupervised learning: the k-means hot path (reference dask_ml/cluster/__init__.py exports
KMeans and SpectralClustering; only KMeans is in scope)."""
from .k_means import KMeans, k_means  # noqa: F401

__all__ = ["KMeans", "k_means"]
