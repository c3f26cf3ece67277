from .pairwise import (  # noqa: F401
    euclidean_distances,
    pairwise_distances,
    pairwise_distances_argmin_min,
    pairwise_kernels,
)

__all__ = ["euclidean_distances", "pairwise_distances", "pairwise_distances_argmin_min", "pairwise_kernels"]
