"""Nystrom SpectralClustering on the device -- the one in-tree CALLER of the k-means hot path
(reference dask_ml/cluster/spectral.py:22-375).

Same constructor, attributes, validation and error messages as the reference.  The two affinity blocks
(A: n_components x n_components, B: n_components x rest) come from ``b200km_pairwise_kernel_*``; the final
clustering of the embedding is the B200 ``KMeans`` with the embedding passed device-to-device.  The small
dense algebra in between (pinv / svd of an n_components-sized matrix, one (n x n_components) x (n_components
x k) product, row scalings) is plain library work and uses torch's cuSOLVER / cuBLAS calls, as the reference
uses numpy / scipy there.
"""
from __future__ import annotations

import logging

import numpy as np
import sklearn.cluster
import torch
from sklearn.base import BaseEstimator, ClusterMixin
from sklearn.utils import check_random_state

from .._engine import get_engine
from ..metrics.pairwise import PAIRWISE_KERNEL_FUNCTIONS, _dev, pairwise_kernels
from ..utils import DeviceArray, check_array
from .k_means import KMeans

logger = logging.getLogger(__name__)

__all__ = ["SpectralClustering", "embed"]


class SpectralClustering(BaseEstimator, ClusterMixin):
    """Apply parallel Spectral Clustering (Nystrom approximation; reference spectral.py:22-179).

    Parameters are the reference's: ``n_clusters, eigen_solver, random_state, n_init, gamma, affinity,
    n_neighbors, eigen_tol, assign_labels, degree, coef0, kernel_params, n_jobs, n_components,
    persist_embedding, kmeans_params``.

    Attributes
    ----------
    assign_labels_ : the fitted clustering estimator
    labels_ : DeviceArray (or ndarray for a scikit-learn ``assign_labels``)
    eigenvalues_ : DeviceArray -- the leading singular values of the normalised affinity block
    """

    def __init__(self, n_clusters=8, eigen_solver=None, random_state=None, n_init="auto", gamma=1.0, affinity="rbf",
                 n_neighbors=10, eigen_tol=0.0, assign_labels="kmeans", degree=3, coef0=1, kernel_params=None,
                 n_jobs=1, n_components=100, persist_embedding=False, kmeans_params=None):
        self.n_clusters = n_clusters
        self.eigen_solver = eigen_solver
        self.random_state = random_state
        self.n_init = n_init
        self.gamma = gamma
        self.affinity = affinity
        self.n_neighbors = n_neighbors
        self.eigen_tol = eigen_tol
        self.assign_labels = assign_labels
        self.degree = degree
        self.coef0 = coef0
        self.kernel_params = kernel_params
        self.n_jobs = n_jobs
        self.n_components = n_components
        self.persist_embedding = persist_embedding
        self.kmeans_params = kmeans_params

    def _check_array(self, X):
        """reference :155-159 -- validated and cast to float64."""
        logger.info("Starting check array")
        X = check_array(X, accept_dask_dataframe=False)
        result = _dev(X, get_engine()).to(torch.float64)
        logger.info("Finished check array")
        return result

    def fit(self, X, y=None):
        X = self._check_array(X)
        n_components = self.n_components
        metric = self.affinity
        rng = check_random_state(self.random_state)
        n_clusters = self.n_clusters

        # kmeans for final clustering (reference :187-207)
        if isinstance(self.assign_labels, str):
            if self.assign_labels == "kmeans":
                km = KMeans(n_clusters=n_clusters, random_state=int(rng.randint(0, np.iinfo("i4").max)),
                            n_init=self.n_init)
            elif self.assign_labels == "sklearn-kmeans":
                km = sklearn.cluster.KMeans(n_clusters=n_clusters, random_state=rng, n_init=self.n_init)
            else:
                msg = "Unknown 'assign_labels' {!r}".format(self.assign_labels)
                raise ValueError(msg)
        elif isinstance(self.assign_labels, BaseEstimator):
            km = self.assign_labels
        else:
            raise TypeError("Invalid type {} for 'assign_labels'".format(type(self.assign_labels)))

        if self.kmeans_params:
            km.set_params(**self.kmeans_params)

        n = len(X)
        if n <= n_components:
            msg = ("'n_components' must be smaller than the number of samples."
                   " Got {} components and {} samples".format(n_components, n))
            raise ValueError(msg)

        params = self.kernel_params or {}
        params["gamma"] = self.gamma
        params["degree"] = self.degree
        params["coef0"] = self.coef0

        # indices for our exact / approximate blocks (:226-231)
        inds = np.arange(n)
        keep = rng.choice(inds, n_components, replace=False)
        keep.sort()
        rest = ~np.isin(inds, keep)
        dev = X.device
        keep_t = torch.from_numpy(keep).to(dev)
        rest_t = torch.from_numpy(inds[rest]).to(dev)
        X_keep = X[keep_t]
        X_rest = X[rest_t]

        A, B = embed(X_keep, X_rest, n_components, metric, params)  # :242
        A, B = A.tensor, B.tensor

        # now the approximation of C (:246-263)
        a = A.sum(0)
        b1 = B.sum(1)
        b2 = B.sum(0)
        A_inv = torch.linalg.pinv(A)
        inner = A_inv @ b1
        d1_si = 1 / torch.sqrt(a + b1)
        d2_si = 1 / torch.sqrt(b2 + B.T @ inner)
        A2 = d1_si.reshape(-1, 1) * A * d1_si.reshape(1, -1)
        B2 = d1_si.reshape(-1, 1) * B * d2_si.reshape(1, -1)

        U_A, S_A, _ = torch.linalg.svd(A2)  # :265-269

        # Eq 16 (:272-275)
        V2 = np.sqrt(float(n_components) / n) * torch.vstack([A2, B2.T]) @ U_A[:, :n_clusters] @ torch.diag(
            1.0 / torch.sqrt(S_A[:n_clusters]))
        # normalize (Eq. 4, :283)
        U2 = (V2.T / torch.sqrt((V2 ** 2).sum(1))).T
        # Recover original indices (:288): rows of U2 are ordered [keep, rest]
        order = torch.argsort(torch.cat([keep_t, rest_t]))
        U2 = U2[order].contiguous()

        self._embedding = U2   # (n, n_clusters), original row order; kept for tests / inspection
        self._keep = keep
        logger.info("k-means for assign_labels[starting]")
        if isinstance(km, KMeans):
            km.fit(U2)  # device-to-device: the embedding never visits the host
        else:
            km.fit(U2.cpu().numpy())
        logger.info("k-means for assign_labels[finished]")

        self.assign_labels_ = km
        self.labels_ = km.labels_
        self.eigenvalues_ = DeviceArray(S_A[:n_clusters])
        return self


def embed(X_keep, X_rest, n_components, metric, kernel_params):
    """reference :312-336 -- the exact block A and the approximate block B of the affinity matrix."""
    if isinstance(metric, str):
        if metric not in PAIRWISE_KERNEL_FUNCTIONS:
            msg = "Unknown affinity metric name '{}'. Expected one " "of '{}'".format(
                metric, PAIRWISE_KERNEL_FUNCTIONS.keys())
            raise ValueError(msg)
        A = pairwise_kernels(X_keep, metric=metric, filter_params=True, **kernel_params)
        B = pairwise_kernels(X_keep, X_rest, metric=metric, filter_params=True, **kernel_params)
    elif callable(metric):
        A = metric(X_keep, **kernel_params)
        B = metric(X_keep, X_rest, **kernel_params)
        A = A if isinstance(A, DeviceArray) else DeviceArray(_dev(A))
        B = B if isinstance(B, DeviceArray) else DeviceArray(_dev(B))
    else:
        msg = "Unexpected type for 'affinity' '{}'. Must be string kernel name, array, or callable"
        raise TypeError(msg)
    return A, B
