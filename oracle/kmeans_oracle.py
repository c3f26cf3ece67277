"""CPU oracle for the k-means hot path of dask-ml.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this module.  The product package
(``dask_ml_b200``) never imports it and has no CPU fallback.

What this is
------------
A restatement of the reference algorithm with the dask layer removed.  dask and
distributed are not installable in the build container (no network, no wheel), so
a plain ``import dask_ml`` is impossible; the arithmetic of the path, however, lives in
third-party packages that ARE importable (scikit-learn's Cython ``ArgKmin`` /
``euclidean_distances`` and a numba scatter loop) and this module calls those same
functions per row-block, replacing only dask's chunking / graph execution:

* blocks      = explicit row chunk sizes (``row_chunks``), mirroring
                ``da.from_array(X, chunks=(max(1, len(X)//cpu_count()), d))``
                (reference dask_ml/cluster/k_means.py:192-193)
* chain-sum   = left fold over block partials in block order (k_means.py:595-597)
* bincount    = ``np.bincount(labels, minlength=k)`` (k_means.py:598)
* scheduler   = optional ThreadPoolExecutor over blocks (threaded scheduler stand-in)

Parity pin status: PINNED to the reference's own source
-------------------------------------------------------
* ``oracle/reference_loader.py`` imports the UNMODIFIED reference modules (dask_ml/cluster/k_means.py,
  metrics/pairwise.py, utils.py, ...) from /root/reference on top of an eager numpy-backed stand-in for the small
  dask surface they touch (``oracle/dask_stub``), and ``tests/test_reference_pin.py`` demands that every function
  here returns what the reference returns on the same inputs: Lloyd at C1-C4 shapes (k up to 1500), both inertia
  branches, an empty cluster, the ``shift < tol`` exit, k-means|| through the supplement and the re-cluster branch,
  ``evaluate_cost`` / ``_sample_points`` with shared draws, the three pairwise functions, the estimator end to end --
  labels / picked rows exact, floats to 1e-12 (bit-equal on this box).  ``tests/golden/*.npz`` are generated from the
  reference (``tests/golden/make_golden.py``), not from this module.
* The reference's own tests for the path (tests/metrics/test_metrics.py:15-71, tests/test_kmeans.py:36-94, 184-197;
  expectations computed live from scikit-learn) are replayed in tests/test_oracle_pin.py.
* What remains an ASSUMPTION: the stand-in's reading of dask itself, which is not on this box -- per-chunk MT19937
  seeding of ``da.random.RandomState`` (``DaskLikeRandomState`` below follows the same reading) and the fan-in-4
  tree of ``Array.sum()`` (``dask_sum``).  Neither touches labels, centres or picked rows beyond the last ulp of a
  sum, except that the concrete rows k-means|| draws for a given seed depend on the seeding scheme.

Every function cites the reference lines it follows (R/ = /root/reference/).
"""
from __future__ import annotations

import math
from concurrent.futures import ThreadPoolExecutor
from typing import Callable, Iterable, List, Optional, Sequence, Tuple

import numba
import numpy as np
import sklearn.cluster
from sklearn import metrics as skmetrics
from sklearn.utils import extmath as skextmath

__all__ = [
    "row_chunks",
    "split_blocks",
    "dask_sum",
    "row_norms",
    "pairwise_distances_argmin_min",
    "pairwise_distances",
    "euclidean_distances",
    "centers_dense",
    "lloyd_step",
    "kmeans_single_lloyd",
    "evaluate_cost",
    "sample_points",
    "init_scalable",
    "KMeansOracle",
    "DaskLikeRandomState",
    "PhiloxDraws",
    "philox4x32_10",
    "philox_uniform_f64",
]


# --------------------------------------------------------------------------------------
# Block structure (the only thing dask contributes besides scheduling)
# --------------------------------------------------------------------------------------
def row_chunks(n: int, chunk: Optional[int] = None, ncpu: Optional[int] = None) -> Tuple[int, ...]:
    """Row chunk sizes as ``KMeans._check_array`` makes them for an ndarray.

    R/dask_ml/cluster/k_means.py:192-193: ``da.from_array(X, chunks=(max(1, len(X) //
    cpu_count()), X.shape[-1]))`` -> equal chunks of that size plus one remainder chunk.
    """
    if chunk is None:
        import multiprocessing

        ncpu = ncpu or multiprocessing.cpu_count()
        chunk = max(1, n // ncpu)
    sizes = [chunk] * (n // chunk)
    if n % chunk:
        sizes.append(n % chunk)
    return tuple(sizes)


def split_blocks(X: np.ndarray, chunks: Sequence[int]) -> List[np.ndarray]:
    """Views of X per row block; zero-length chunks are dropped (R/dask_ml/utils.py:180-185)."""
    out, start = [], 0
    for c in chunks:
        if c:
            out.append(X[start : start + c])
        start += c
    assert start == len(X), "chunks must cover X"
    return out


def dask_sum(v: np.ndarray, chunks: Sequence[int], split_every: int = 4):
    """``dask.array.Array.sum()`` of a 1-d array cut into ``chunks``: one ``np.sum(block, dtype=acc)`` per block, then
    groups of ``split_every`` neighbouring partials are summed until one is left (dask/array/reductions.py
    ``reduction`` / ``_tree_reduce``; config default ``split_every`` = 4).  The accumulator dtype comes from the
    array's DECLARED dtype; everything the path sums is declared float64 (``map_blocks(dtype=float)``,
    R/dask_ml/metrics/pairwise.py:59-66; ``from_delayed(..., "f8")``, :41-43) even where scikit-learn hands back
    float32 blocks, so float32 values are accumulated in float64."""
    parts = [np.sum(b, dtype=np.float64, keepdims=True) for b in split_blocks(v, chunks)]
    if not parts:
        return np.float64(0.0)
    while len(parts) > 1:
        parts = [np.sum(np.concatenate(parts[i : i + split_every]), dtype=np.float64, keepdims=True)
                 for i in range(0, len(parts), split_every)]
    return parts[0][0]


def _map_blocks(fn: Callable, blocks: Iterable, pool: Optional[ThreadPoolExecutor]):
    if pool is None:
        return [fn(b) for b in blocks]
    return list(pool.map(fn, blocks))


# --------------------------------------------------------------------------------------
# metrics/pairwise.py + utils.row_norms
# --------------------------------------------------------------------------------------
def row_norms(X: np.ndarray, squared: bool = False, chunks: Optional[Sequence[int]] = None) -> np.ndarray:
    """R/dask_ml/utils.py:83-88 -- sklearn ``row_norms`` per block."""
    if chunks is None:
        return skextmath.row_norms(X, squared=squared)
    return np.concatenate([skextmath.row_norms(b, squared=squared) for b in split_blocks(X, chunks)])


def pairwise_distances_argmin_min(
    X: np.ndarray,
    Y: np.ndarray,
    chunks: Optional[Sequence[int]] = None,
    metric: str = "euclidean",
    metric_kwargs: Optional[dict] = None,
    pool: Optional[ThreadPoolExecutor] = None,
) -> Tuple[np.ndarray, np.ndarray]:
    """R/dask_ml/metrics/pairwise.py:19-47.

    One ``sklearn.metrics.pairwise_distances_argmin_min(x, Y, metric=, metric_kwargs=)``
    call per row block (:31-33); argmins are int64, mins float64 (:36-44); blocks are
    concatenated in order (:45-46).
    """
    chunks = chunks if chunks is not None else (len(X),)

    def one(b):
        return skmetrics.pairwise_distances_argmin_min(b, Y, metric=metric, metric_kwargs=metric_kwargs)

    res = _map_blocks(one, split_blocks(X, chunks), pool)
    if not res:
        return np.empty(0, np.int64), np.empty(0, np.float64)
    arg = np.concatenate([r[0] for r in res]).astype(np.int64, copy=False)
    mn = np.concatenate([r[1] for r in res]).astype(np.float64, copy=False)
    return arg, mn


def pairwise_distances(
    X: np.ndarray,
    Y: np.ndarray,
    chunks: Optional[Sequence[int]] = None,
    metric: str = "euclidean",
    pool: Optional[ThreadPoolExecutor] = None,
    **kwargs,
) -> np.ndarray:
    """R/dask_ml/metrics/pairwise.py:50-67 -- ``sklearn.metrics.pairwise_distances`` per block.

    Note the result dtype is whatever sklearn returns per block (float32 when both inputs
    are float32: SK/metrics/pairwise.py:581,636), not the ``dtype=float`` dask declares.
    """
    chunks = chunks if chunks is not None else (len(X),)
    res = _map_blocks(lambda b: skmetrics.pairwise_distances(b, Y, metric=metric, **kwargs), split_blocks(X, chunks), pool)
    return np.concatenate(res, axis=0) if res else np.empty((0, len(Y)))


def euclidean_distances(
    X: np.ndarray,
    Y: Optional[np.ndarray] = None,
    Y_norm_squared: Optional[np.ndarray] = None,
    squared: bool = False,
    X_norm_squared: Optional[np.ndarray] = None,
) -> np.ndarray:
    """R/dask_ml/metrics/pairwise.py:70-105 -- dask-ml's OWN formula (no float64 up-cast):
    ``sqrt(max(-2 X.Y^T + XX + YY, 0))`` in the promoted input dtype, with the same
    shape checks on precomputed norms (:80-96)."""
    if Y is None:
        Y = X
    if X_norm_squared is not None:
        XX = X_norm_squared
        if XX.shape == (1, X.shape[0]):
            XX = XX.T
        elif XX.shape != (X.shape[0], 1):
            raise ValueError("Incompatible dimensions for X and X_norm_squared")
    else:
        XX = skextmath.row_norms(X, squared=True)[:, np.newaxis]
    if X is Y:
        YY = XX.T
    elif Y_norm_squared is not None:
        YY = Y_norm_squared[:, np.newaxis] if Y_norm_squared.ndim < 2 else Y_norm_squared
        if YY.shape != (1, Y.shape[0]):
            raise ValueError("Incompatible dimensions for Y and Y_norm_squared")
    else:
        YY = skextmath.row_norms(Y, squared=True)[np.newaxis, :]
    dist = -2 * np.dot(X, Y.T) + XX + YY
    dist = np.maximum(dist, 0)
    return dist if squared else np.sqrt(dist)


# --------------------------------------------------------------------------------------
# Lloyd
# --------------------------------------------------------------------------------------
@numba.njit(nogil=True, fastmath=True, cache=True)
def centers_dense(X, labels, n_clusters):
    """R/dask_ml/cluster/k_means.py:622-632 -- per-block scatter-add of rows by label into a
    float64 (k, d) table, serial over rows (numba, nogil, fastmath)."""
    acc = np.zeros((n_clusters, X.shape[1]), dtype=np.float64)
    for row in range(X.shape[0]):
        dst = labels[row]
        for col in range(X.shape[1]):
            acc[dst, col] += X[row, col]
    return acc


def lloyd_step(
    X: np.ndarray,
    centers: np.ndarray,
    n_clusters: int,
    chunks: Sequence[int],
    pool: Optional[ThreadPoolExecutor] = None,
):
    """One pass of the loop body, R/dask_ml/cluster/k_means.py:576-602.

    returns (labels int32, squared min distances f64, new_centers f64 (k,d), counts int64 (k,))
    """
    labels, distances = pairwise_distances_argmin_min(
        X, centers, chunks, metric="euclidean", metric_kwargs={"squared": True}, pool=pool
    )  # :576-578
    labels = labels.astype(np.int32)  # :579
    xb = split_blocks(X, chunks)
    lb = split_blocks(labels, chunks)
    parts = _map_blocks(lambda t: centers_dense(t[0], t[1], n_clusters), list(zip(xb, lb)), pool)  # :582-594
    total = parts[0]
    for p in parts[1:]:  # python sum(): left fold in block order (:595-597)
        total = total + p
    counts = np.bincount(labels, minlength=n_clusters)  # :598
    counts = np.maximum(counts, 1)  # :600
    new_centers = total / counts[:, None]  # :601
    return labels, distances, new_centers, counts


def kmeans_single_lloyd(
    X: np.ndarray,
    n_clusters: int,
    centers: np.ndarray,
    max_iter: int = 300,
    tol: float = 1e-4,
    chunks: Optional[Sequence[int]] = None,
    pool: Optional[ThreadPoolExecutor] = None,
):
    """R/dask_ml/cluster/k_means.py:549-619 with ``centers`` = result of ``k_init``.

    Quirks kept on purpose: raw ``tol`` against the squared Frobenius shift (:605-609);
    on convergence the OLD centres are returned (break at :608-609 precedes :610);
    empty clusters go to the origin (:598-601); after the loop, if ``shift > 1e-7`` labels
    and distances are recomputed with the NON-squared metric and ``inertia`` is their sum
    (:612-616); centres are cast back to X.dtype (:617); ``n_iter = i + 1`` (:619).

    returns (labels int32, inertia float, centers (k,d) X.dtype, n_iter)
    """
    chunks = chunks if chunks is not None else row_chunks(len(X))
    dt = X.dtype
    shift = None
    labels = distances = None
    i = -1
    for i in range(max_iter):
        labels, distances, new_centers, _ = lloyd_step(X, centers, n_clusters, chunks, pool)
        shift = skextmath.squared_norm(centers - new_centers)  # :605
        if shift < tol:  # :608
            break
        centers = new_centers  # :610
    if shift > 1e-7:  # :612 (max_iter == 0 raises in the reference too: shift unbound)
        labels, distances = pairwise_distances_argmin_min(X, centers, chunks, pool=pool)  # :613
        labels = labels.astype(np.int32)
    inertia = float(dask_sum(distances, chunks))  # :616
    centers = centers.astype(dt)  # :617
    return labels, inertia, centers, i + 1


# --------------------------------------------------------------------------------------
# k-means|| (init_scalable)
# --------------------------------------------------------------------------------------
def _min_dist_sq(X, centers, chunks, pool=None) -> np.ndarray:
    """``pairwise_distances(X, centers).min(1) ** 2`` per block -- sqrt then re-square, in the
    dtype sklearn returns (R/dask_ml/cluster/k_means.py:519, :533)."""
    res = _map_blocks(
        lambda b: skmetrics.pairwise_distances(b, centers, metric="euclidean").min(1) ** 2,
        split_blocks(X, chunks),
        pool,
    )
    return np.concatenate(res)


def evaluate_cost(X: np.ndarray, centers: np.ndarray, chunks: Optional[Sequence[int]] = None, pool=None) -> float:
    """R/dask_ml/cluster/k_means.py:516-519: ``(pairwise_distances(X, centers).min(1) ** 2).sum()`` -- float32
    blocks for float32 X (scikit-learn), summed in float64 by dask's reduction tree (``dask_sum``)."""
    chunks = chunks if chunks is not None else (len(X),)
    d2 = _min_dist_sq(X, centers, chunks, pool)
    return float(dask_sum(d2, chunks))


def sample_points(
    X: np.ndarray,
    centers: np.ndarray,
    oversampling_factor: float,
    draws: np.ndarray,
    chunks: Optional[Sequence[int]] = None,
    pool=None,
    return_p: bool = False,
):
    """R/dask_ml/cluster/k_means.py:522-541 with the uniform ``draws`` vector supplied by the
    caller (one U[0,1) per row).  ``p = l * d2 / d2.sum()`` (:533-535); picked = p > draws
    (:538); returns the row indices (:540-541).  For float32 X: ``l * d2`` stays float32, the float64 ``denom``
    (``dask_sum``) promotes the quotient to float64 (numpy >= 2 promotion, which the reference's pins imply)."""
    chunks = chunks if chunks is not None else (len(X),)
    distances = _min_dist_sq(X, centers, chunks, pool)
    denom = dask_sum(distances, chunks)
    p = oversampling_factor * distances / denom
    (idx,) = np.where(p > draws)
    if return_p:
        return idx, p
    return idx


class DaskLikeRandomState:
    """Restatement of ``dask.array.random.RandomState`` for the three calls the path makes
    (k_means.py:386, :496-498, :505-508, :537).  UNVERIFIED: written from memory of
    dask/array/random.py + dask/utils.py::random_state_data (dask is not on this box).

    Parent ``np.random.RandomState(seed)``; every call draws ``624*4*nchunks`` bytes from the
    parent and chunk i gets ``np.random.RandomState(<its 624 uint32 words>)``.
    """

    def __init__(self, seed=None):
        self._numpy_state = np.random.RandomState(seed)

    def _state_data(self, n):
        raw = self._numpy_state.bytes(624 * n * 4)
        return list(np.frombuffer(raw, dtype="<u4").reshape((n, -1)))

    def uniform(self, size: int, chunks: Sequence[int]) -> np.ndarray:
        chunks = [c for c in chunks]
        seeds = self._state_data(len(chunks))
        return np.concatenate([np.random.RandomState(s).uniform(size=c) for s, c in zip(seeds, chunks)]) if chunks else np.empty(0)

    def choice(self, n: int, size: int) -> np.ndarray:
        (s,) = self._state_data(1)
        return np.random.RandomState(s).choice(np.arange(0, n), size=size, replace=False)

    def randint(self, low, high) -> int:
        (s,) = self._state_data(1)
        return int(np.random.RandomState(s).randint(low, high))


# ---- Philox4x32-10: the counter-based generator the CUDA path uses ---------------------
_PHILOX_M0 = np.uint64(0xD2511F53)
_PHILOX_M1 = np.uint64(0xCD9E8D57)
_PHILOX_W0 = np.uint32(0x9E3779B9)
_PHILOX_W1 = np.uint32(0xBB67AE85)


def philox4x32_10(counter: np.ndarray, key: Tuple[int, int]) -> np.ndarray:
    """Philox4x32 with 10 rounds (Salmon et al., SC'11 -- published algorithm; no reference
    counterpart).  ``counter``: (n, 4) uint32; ``key``: two uint32.  Returns (n, 4) uint32."""
    c = np.array(counter, dtype=np.uint32, copy=True)
    k0 = np.uint32(key[0] & 0xFFFFFFFF)
    k1 = np.uint32(key[1] & 0xFFFFFFFF)
    mask = np.uint64(0xFFFFFFFF)
    with np.errstate(over="ignore"):
        for _ in range(10):
            p0 = _PHILOX_M0 * c[:, 0].astype(np.uint64)
            p1 = _PHILOX_M1 * c[:, 2].astype(np.uint64)
            hi0 = (p0 >> np.uint64(32)).astype(np.uint32)
            lo0 = (p0 & mask).astype(np.uint32)
            hi1 = (p1 >> np.uint64(32)).astype(np.uint32)
            lo1 = (p1 & mask).astype(np.uint32)
            n0 = hi1 ^ c[:, 1] ^ k0
            n2 = hi0 ^ c[:, 3] ^ k1
            c = np.stack([n0, lo1, n2, lo0], axis=1)
            k0 = np.uint32((int(k0) + int(_PHILOX_W0)) & 0xFFFFFFFF)
            k1 = np.uint32((int(k1) + int(_PHILOX_W1)) & 0xFFFFFFFF)
    return c


def philox_uniform_f64(seed: int, round_: int, row_start: int, n: int) -> np.ndarray:
    """U[0,1) doubles exactly as the CUDA kernel ``b200km_bernoulli_pick`` draws them:
    key = (seed lo32, seed hi32); counter = (row lo32, row hi32, round, 0x6b6d7c7c);
    u = ((w0 >> 5) * 2^26 + (w1 >> 6)) / 2^53 (the 53-bit construction numpy's
    ``random_sample`` uses)."""
    rows = np.arange(row_start, row_start + n, dtype=np.uint64)
    ctr = np.empty((n, 4), dtype=np.uint32)
    ctr[:, 0] = (rows & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    ctr[:, 1] = (rows >> np.uint64(32)).astype(np.uint32)
    ctr[:, 2] = np.uint32(round_ & 0xFFFFFFFF)
    ctr[:, 3] = np.uint32(0x6B6D7C7C)
    w = philox4x32_10(ctr, (seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF))
    a = (w[:, 0] >> np.uint32(5)).astype(np.float64)
    b = (w[:, 1] >> np.uint32(6)).astype(np.float64)
    return (a * 67108864.0 + b) / 9007199254740992.0


class PhiloxDraws:
    """Draw provider matching the CUDA path's counter-based RNG (round counter advances per
    call, like the product's init_scalable)."""

    def __init__(self, seed: int):
        self.seed = int(seed)
        self.round = 0

    def uniform(self, size: int, chunks=None) -> np.ndarray:
        u = philox_uniform_f64(self.seed, self.round, 0, size)
        self.round += 1
        return u

    # host-side draws of the two tail branches, as the product's PhiloxRandomState makes them
    def choice(self, n: int, size: int) -> np.ndarray:
        return np.random.Generator(np.random.Philox(key=[self.seed, 0x63686F69])).choice(n, size=size, replace=False)

    def randint(self, low: int, high: int) -> int:
        return int(np.random.Generator(np.random.Philox(key=[self.seed, 0x72616E64])).integers(low, high))


def init_scalable(
    X: np.ndarray,
    n_clusters: int,
    random_state,
    max_iter: Optional[int] = None,
    oversampling_factor: float = 2,
    n_init="auto",
    chunks: Optional[Sequence[int]] = None,
    pool=None,
    supplement: Optional[Callable[[int, int], np.ndarray]] = None,
    trace: Optional[list] = None,
):
    """R/dask_ml/cluster/k_means.py:435-513.

    ``random_state`` needs ``uniform(size, chunks)`` and, for the two tail branches,
    ``choice(n, size)`` / ``randint(low, high)`` (DaskLikeRandomState) -- or pass
    ``supplement(n, need) -> indices`` to inject the random-supplement rows.
    ``trace`` (a list) receives per-round dicts for parity tests.
    """
    chunks = chunks if chunks is not None else row_chunks(len(X))
    idx = 0  # :451  first candidate is always row 0
    centers = X[idx, np.newaxis]  # :452
    c_idx = {idx}  # :453
    cost = evaluate_cost(X, centers, chunks, pool)  # :456
    n_iter = 0 if cost == 0 else int(np.round(np.log(cost)))  # :458-461
    if max_iter is not None:
        n_iter = min(max_iter, n_iter)  # :463-464
    if trace is not None:
        trace.append({"cost0": cost, "n_iter": n_iter})
    for _ in range(n_iter):  # :467
        draws = random_state.uniform(size=len(X), chunks=chunks)  # :537
        new_idxs = sample_points(X, centers, oversampling_factor, draws, chunks, pool)  # :472
        c_idx |= set(int(v) for v in new_idxs)  # :473-474
        centers = X[sorted(c_idx)]  # :479
        if trace is not None:
            trace.append({"picked": sorted(int(v) for v in new_idxs), "n_centers": len(c_idx)})
    if len(centers) < n_clusters:  # :489
        need = n_clusters - len(centers)  # :494
        if supplement is not None:
            locs = sorted(int(v) for v in supplement(len(X), need))
        else:
            locs = sorted(int(v) for v in random_state.choice(len(X), need))  # :495-499
        extra = X[locs]  # :500
        if trace is not None:
            trace.append({"supplement": locs})
        return np.vstack([centers, extra])  # :501
    # Steps 7, 8 without weights (:503-513)
    rng2 = random_state.randint(0, np.iinfo("i4").max - 1)  # :505-509
    km = sklearn.cluster.KMeans(n_clusters, random_state=rng2, n_init=n_init)  # :510
    km.fit(centers)  # :511
    return km.cluster_centers_  # :513


# --------------------------------------------------------------------------------------
# Estimator shell
# --------------------------------------------------------------------------------------
class KMeansOracle:
    """R/dask_ml/cluster/k_means.py:31-251 restated over numpy blocks (fit / predict / transform
    and the published attributes).  ``init`` may be 'k-means||' or a (k, d) ndarray."""

    def __init__(self, n_clusters=8, init="k-means||", oversampling_factor=2, max_iter=300, tol=1e-4,
                 random_state=None, init_max_iter=None, n_init="auto", chunks=None, n_threads=None,
                 draws=None):
        self.n_clusters = n_clusters
        self.init = init
        self.oversampling_factor = oversampling_factor
        self.max_iter = max_iter
        self.tol = tol
        self.random_state = random_state
        self.init_max_iter = init_max_iter
        self.n_init = n_init
        self.chunks = chunks
        self.n_threads = n_threads
        self.draws = draws

    def _check_array(self, X):
        X = np.asarray(X)
        if X.dtype == np.int32:  # :187-190
            X = X.astype(np.float32)
        elif X.dtype == np.int64:
            X = X.astype(np.float64)
        from sklearn.utils.validation import check_array

        X = check_array(X, accept_sparse=False, ensure_all_finite=False)
        if np.isnan(X).any() or np.isinf(X).any():  # :195-201
            raise ValueError("Input contains NaN, infinity or a value too large for dtype('float64').")
        return X

    def _k_init(self, X, chunks, pool):
        init = self.init
        if isinstance(init, np.ndarray):  # :360-377
            K, P = init.shape
            if K != self.n_clusters:
                raise ValueError(
                    "Number of centers in provided 'init' ({}) does not match 'n_clusters' ({})".format(K, self.n_clusters)
                )
            if P != X.shape[1]:
                raise ValueError(
                    "Number of features in the provided 'init' ({}) do not match the number of features in 'X'".format(P)
                )
            return init
        if not isinstance(init, str):
            raise TypeError("'init' must be an array or str, got {}".format(type(init)))
        if init != "k-means||":
            raise ValueError("oracle restates only init='k-means||' and ndarray init")
        rs = self.draws if self.draws is not None else DaskLikeRandomState(self.random_state)
        return init_scalable(X, self.n_clusters, rs, self.init_max_iter, self.oversampling_factor, self.n_init, chunks, pool)

    def fit(self, X, y=None):
        X = self._check_array(X)
        chunks = self.chunks if self.chunks is not None else row_chunks(len(X))
        pool = ThreadPoolExecutor(self.n_threads) if self.n_threads else None
        try:
            centers = self._k_init(X, chunks, pool)
            labels, inertia, centers, n_iter = kmeans_single_lloyd(
                X, self.n_clusters, centers, self.max_iter, self.tol, chunks, pool
            )
        finally:
            if pool is not None:
                pool.shutdown()
        self.cluster_centers_ = centers
        self.labels_ = labels
        self.inertia_ = inertia
        self.n_iter_ = n_iter
        self.n_features_in_ = X.shape[1]
        return self

    def predict(self, X):  # :230-251
        X = self._check_array(X)
        return pairwise_distances_argmin_min(X, self.cluster_centers_, self.chunks)[0].astype(np.int32)

    def transform(self, X, y=None):  # :225-228
        X = self._check_array(X)
        return euclidean_distances(X, self.cluster_centers_)

    def fit_transform(self, X, y=None):
        return self.fit(X).transform(X)


# --------------------------------------------------------------------------------------
# SpectralClustering's Nystrom embedding (the in-tree caller of KMeans)
# --------------------------------------------------------------------------------------
def _pairwise_kernel(X, Y, metric, gamma=None, degree=3, coef0=1):
    """R/dask_ml/metrics/pairwise.py:135-191 in numpy (gamma default 1/d)."""
    if gamma is None:
        gamma = 1.0 / X.shape[1]
    if metric == "rbf":
        return np.exp(-gamma * euclidean_distances(X, Y, squared=True))
    if metric == "linear":
        return X @ Y.T
    if metric == "polynomial":
        return (gamma * (X @ Y.T) + coef0) ** degree
    if metric == "sigmoid":
        return np.tanh(gamma * (X @ Y.T) + coef0)
    raise ValueError("Unknown kernel %r" % metric)


def spectral_embedding(X, keep, n_clusters, metric="rbf", gamma=1.0, degree=3, coef0=1):
    """R/dask_ml/cluster/spectral.py:226-288 with ``keep`` (the sorted sampled row indices) supplied.
    Returns (U2 (n, n_clusters) in the ORIGINAL row order, leading singular values)."""
    from scipy.linalg import pinv, svd

    X = np.asarray(X, dtype=float)  # :157
    n = len(X)
    n_components = len(keep)
    inds = np.arange(n)
    rest = ~np.isin(inds, keep)  # :231
    X_keep, X_rest = X[keep], X[rest]
    A = _pairwise_kernel(X_keep, X_keep, metric, gamma, degree, coef0)  # :318-319
    B = _pairwise_kernel(X_keep, X_rest, metric, gamma, degree, coef0)  # :320-322
    a = A.sum(0)  # :246-248
    b1 = B.sum(1)
    b2 = B.sum(0)
    A_inv = pinv(A)  # :251
    inner = A_inv.dot(b1)  # :253
    d1_si = 1 / np.sqrt(a + b1)  # :254
    d2_si = 1 / np.sqrt(b2 + B.T.dot(inner))  # :256
    A2 = d1_si.reshape(-1, 1) * A * d1_si.reshape(1, -1)  # :260
    B2 = d1_si.reshape(-1, 1) * B * d2_si.reshape(1, -1)  # :264
    U_A, S_A, V_A = svd(A2)  # :267
    V2 = np.sqrt(float(n_components) / n) * np.vstack([A2, B2.T]).dot(U_A[:, :n_clusters]).dot(
        np.diag(1.0 / np.sqrt(S_A[:n_clusters])))  # :272-275
    U2 = (V2.T / np.sqrt((V2 ** 2).sum(1))).T  # :283
    idx = np.argsort(np.concatenate([keep, inds[rest]]))  # :288 (_slice_mostly_sorted)
    return U2[idx], S_A[:n_clusters]
