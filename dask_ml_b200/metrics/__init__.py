from .pairwise import euclidean_distances, pairwise_distances, pairwise_distances_argmin_min  # noqa: F401

__all__ = ["euclidean_distances", "pairwise_distances", "pairwise_distances_argmin_min"]
