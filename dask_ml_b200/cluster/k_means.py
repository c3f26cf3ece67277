"""KMeans on B200: the Lloyd iteration and the k-means|| initialisation.

Drop-in for reference dask_ml/cluster/k_means.py: same class, constructor, attributes, module
functions, argument meaning, quirks and error messages; underneath, every pass over the data
is a hand-written sm_100a kernel called through ``libb200km.so`` and the only cross-GPU
traffic is one small allreduce per Lloyd iteration / one small allgather per k-means|| round.

Quirks of the reference kept on purpose:
  * distances are float64-accurate even for float32 X (K1 + fp64 re-decision of near ties);
  * ``inertia_`` is the sum of NON-squared distances unless the loop converged to
    ``shift <= 1e-7`` (reference :612-616);
  * on convergence the OLD centres are returned (:608-610); ``n_iter_ = i + 1`` (:619);
  * empty clusters go to the origin (:598-601); ``tol`` is compared raw to the squared
    Frobenius shift (:605-609);
  * k-means|| samples ~``oversampling_factor`` rows per round (:535), starts from row 0 (:451),
    runs ``round(ln cost)`` rounds (:461) and supplements with random rows when fewer than k
    candidates were found (:489-501).

Multi-GPU: run one process per GPU with ``torch.distributed`` initialised (NCCL); each rank
passes ITS contiguous row shard to ``fit``.  Attributes are then identical on every rank except
``labels_``, which covers the rank's own rows.
"""
from __future__ import annotations

import logging
from numbers import Integral

import numpy as np
import sklearn.cluster
import torch
from sklearn.base import BaseEstimator, TransformerMixin
from sklearn.utils.validation import check_is_fitted

from .. import _comm
from .._engine import any_rank_flag, get_engine
from ..metrics.pairwise import _dev, euclidean_distances, pairwise_distances_argmin_min
from .._hostmatrix import HostMatrix, row_chunks
from ..utils import DeviceArray, _timed, _timer, check_array

logger = logging.getLogger(__name__)

__all__ = ["KMeans", "k_means", "k_init", "init_scalable", "init_random", "init_pp", "evaluate_cost",
           "_sample_points", "_kmeans_single_lloyd", "PhiloxRandomState", "candidate_weights"]


# ---------------------------------------------------------------------------------------------
# random streams
# ---------------------------------------------------------------------------------------------
class PhiloxRandomState:
    """What ``k_init`` builds for ``init='k-means||'`` when ``random_state`` is an int / None --
    the stand-in for ``da.random.RandomState(seed)`` (reference :384-386).

    The Bernoulli draws are generated ON THE DEVICE by a counter-based Philox4x32-10 keyed by
    ``seed`` with counter (global row, round): independent of the sharding, so 1 GPU and 8 GPUs
    pick the same rows.  dask's per-chunk MT19937 streams cannot be reproduced by construction
    (they depend on the chunk layout); pass an object with ``uniform(size=, chunks=)`` /
    ``choice(n, size)`` / ``randint(low, high)`` as ``random_state`` to inject host draws.
    """

    def __init__(self, seed=None):
        if seed is None:
            seed = int(np.random.SeedSequence().generate_state(1, dtype=np.uint64)[0])
        self.seed = int(seed) & (2**64 - 1)
        self.round = 0
        self.picks = []   # sorted global row indices picked in each k-means|| round so far (a record, never read back)

    def _host(self, tag: int) -> np.random.Generator:
        return np.random.Generator(np.random.Philox(key=[self.seed, tag]))

    def choice(self, n: int, size: int) -> np.ndarray:
        # O(size), not O(n): the reference permutes arange(n) (8 GB at n = 1e9), :495-499
        return self._host(0x63686F69).choice(n, size=size, replace=False)

    def randint(self, low: int, high: int) -> int:
        return int(self._host(0x72616E64).integers(low, high))


class _NumpyRandomState:
    """Adapter giving ``np.random.RandomState`` the three-call protocol above."""

    def __init__(self, rs: np.random.RandomState):
        self.rs = rs

    def uniform(self, size, chunks=None):
        return self.rs.uniform(size=size)

    def choice(self, n, size):
        return self.rs.choice(n, size=size, replace=False)

    def randint(self, low, high):
        return int(self.rs.randint(low, high))


# ---------------------------------------------------------------------------------------------
# estimator
# ---------------------------------------------------------------------------------------------
class KMeans(TransformerMixin, BaseEstimator):
    """Scalable KMeans for clustering -- same constructor and attributes as reference
    dask_ml/cluster/k_means.py:31-251.

    Parameters
    ----------
    n_clusters : int, default 8
    init : {'k-means||', 'k-means++', 'random'} or ndarray (n_clusters, n_features)
    oversampling_factor : int, default 2
    max_iter : int
    tol : float -- compared raw with the squared Frobenius norm of the centre shift
    random_state : int, RandomState instance or None
    init_max_iter : int or None
    n_init : 'auto' or int -- used only by the candidate re-cluster of k-means||
    precompute_distances, copy_x, n_jobs, algorithm : accepted and ignored, as in the reference

    Attributes
    ----------
    cluster_centers_ : np.ndarray (n_clusters, n_features), X's dtype
    labels_ : DeviceArray int32 (``.compute()`` -> ndarray)
    inertia_ : float
    n_iter_ : int
    n_features_in_ : int
    """

    def __init__(self, n_clusters=8, init="k-means||", oversampling_factor=2, max_iter=300, tol=0.0001,
                 precompute_distances="auto", random_state=None, copy_x=True, n_jobs=1, algorithm="full",
                 init_max_iter=None, n_init="auto"):
        self.n_clusters = n_clusters
        self.init = init
        self.oversampling_factor = oversampling_factor
        self.random_state = random_state
        self.max_iter = max_iter
        self.init_max_iter = init_max_iter
        self.algorithm = algorithm
        self.tol = tol
        self.precompute_distances = precompute_distances
        self.n_jobs = n_jobs
        self.copy_x = copy_x
        self.n_init = n_init

    _NONFINITE_MSG = "Input contains NaN, infinity or a value too large for dtype('float64')."

    @_timed(_logger=logger)
    def _check_array(self, X, defer=False):
        """reference :171-202 -> a (n, d) float32/float64 tensor on the device.

        ``defer=True`` (used by ``fit``): a large HOST input is uploaded in row chunks on a side stream and the
        NaN/inf scan runs chunk by chunk behind the copies; the tensor comes back with the outstanding chunks
        attached so the first Lloyd pass can start on the first rows while the rest is still crossing PCIe.
        The scan's verdict is read -- and the reference's ValueError raised -- at the first host
        synchronisation of the fit (end of Lloyd iteration 0)."""
        try:
            import pandas as pd

            if isinstance(X, pd.DataFrame):
                X = X.values
        except ImportError:  # pragma: no cover
            pass
        if type(X).__module__.startswith("dask.") and type(X).__name__ == "DataFrame":
            X = X.to_dask_array(lengths=True)
        X = check_array(X, accept_dask_dataframe=False, accept_unknown_chunks=False, accept_sparse=False,
                        remove_zero_chunks=True, ensure_all_finite=False)
        eng = get_engine()
        host = isinstance(X, np.ndarray) or (isinstance(X, torch.Tensor) and not X.is_cuda)
        if host and getattr(X, "ndim", 0) == 2 and str(X.dtype).replace("torch.", "") in ("float32", "float64") \
                and _out_of_core(eng, X):
            # larger than the device can hold: X stays in (page-locked) host memory and every pass streams it;
            # the NaN/inf scan (:195-201) rides on the first pass
            return HostMatrix(X, eng, nonfinite_msg=self._NONFINITE_MSG, share_verdict=defer)
        if (defer and host and hasattr(eng, "to_device_streamed") and X.ndim == 2
                and str(X.dtype).replace("torch.", "") in ("float32", "float64")
                and X.shape[0] * X.shape[1] * (4 if "32" in str(X.dtype) else 8) >= eng.STREAM_MIN_BYTES):
            Xd = eng.to_device_streamed(X)
            # the scan itself (:195-201) is interleaved with Lloyd iteration 0, chunk by chunk
            Xd._b200_bad = torch.zeros(1, dtype=torch.int32, device=eng.device)
            return Xd
        Xd = _dev(X, eng)  # int32 -> float32, int64 -> float64 (:187-190)
        flag = eng.nonfinite_flag(Xd)  # :195-201
        if defer and _comm.default_comm().world > 1:
            # multi-rank fit: the verdict is shared at the fit's first exchange step, so that every rank raises
            # together (a rank raising alone would leave the others waiting in the next collective)
            Xd._b200_bad = flag
        elif int(flag.item()):
            raise ValueError(self._NONFINITE_MSG)
        return Xd

    def fit(self, X, y=None):
        X = self._check_array(X, defer=True)
        labels, centroids, inertia, n_iter = k_means(
            X, self.n_clusters, oversampling_factor=self.oversampling_factor, random_state=self.random_state,
            init=self.init, return_n_iter=True, max_iter=self.max_iter, init_max_iter=self.init_max_iter,
            tol=self.tol, n_init=self.n_init,
        )
        self.cluster_centers_ = centroids
        self.labels_ = labels
        self.inertia_ = inertia.compute().item()
        self.n_iter_ = n_iter
        self.n_features_in_ = X.shape[1]
        return self

    def transform(self, X, y=None):
        check_is_fitted(self, "cluster_centers_")
        X = self._check_array(X)
        if isinstance(X, HostMatrix):
            raise NotImplementedError("transform() of a matrix that does not fit on the device: the (n, n_clusters) "
                                      "result would not fit either; transform row blocks instead")
        return euclidean_distances(X, self.cluster_centers_)

    def predict(self, X):
        """Index of the closest centre for every row (int32), reference :230-251."""
        check_is_fitted(self, "cluster_centers_")
        X = self._check_array(X)
        if isinstance(X, HostMatrix):   # streamed: the same K1 calls block by block
            eng = get_engine()
            pack = eng.pack(eng.to_device(np.ascontiguousarray(self.cluster_centers_), torch.float64))
            out = torch.empty(X.shape[0], dtype=torch.int32, device=eng.device)
            for r0, r1, Xs in row_chunks(X):
                eng.assign(Xs, pack, out[r0:r1], None, want_mind2=True)
            return DeviceArray(out)
        labels = pairwise_distances_argmin_min(X, self.cluster_centers_)[0].astype(np.int32)
        return labels


def k_means(X, n_clusters, init="k-means||", precompute_distances="auto", n_init="auto", max_iter=300,
            verbose=False, tol=1e-4, random_state=None, copy_x=True, n_jobs=-1, algorithm="full",
            return_n_iter=False, oversampling_factor=2, init_max_iter=None):
    """K-means algorithm for clustering (reference :254-294; same defaults, incl. n_jobs=-1)."""
    labels, inertia, centers, n_iter = _kmeans_single_lloyd(
        X, n_clusters, max_iter=max_iter, init=init, verbose=verbose, tol=tol, random_state=random_state,
        oversampling_factor=oversampling_factor, init_max_iter=init_max_iter, n_init=n_init,
    )
    if return_n_iter:
        return labels, centers, inertia, n_iter
    return labels, centers, inertia


def _out_of_core(eng, X) -> bool:
    """Does a host matrix have to stay on the host?  ``B200KM_OUT_OF_CORE=1`` / ``0`` forces the answer; otherwise
    X plus its per-row outputs (labels, distances, flag list: 12 bytes a row) must fit in 80 % of the free device
    memory."""
    import os

    env = os.environ.get("B200KM_OUT_OF_CORE")
    if env is not None and env != "":
        return env not in ("0", "false", "no")
    if not hasattr(eng, "device") or not torch.cuda.is_available():
        return False
    itemsize = 4 if "32" in str(X.dtype) else 8
    need = X.shape[0] * (X.shape[1] * itemsize + 12)
    free, _total = torch.cuda.mem_get_info(torch.device(eng.device))
    return need > 0.8 * free


# ---------------------------------------------------------------------------------------------
# initialisation
# ---------------------------------------------------------------------------------------------
def _as_device_matrix(X) -> torch.Tensor:
    if isinstance(X, HostMatrix):   # stays on the host; every pass streams it (row_chunks)
        return X
    if isinstance(X, torch.Tensor) and X.dim() == 2 and X.dtype in (torch.float32, torch.float64) and \
            X.device == torch.device(get_engine().device):
        return X
    return _dev(X)


def k_init(X, n_clusters, init="k-means||", random_state=None, max_iter=None, oversampling_factor=2, n_init="auto"):
    """Choose the initial centres (reference :310-401; same validation and messages)."""
    if isinstance(init, torch.Tensor):
        init = init.detach().cpu().numpy()
    if isinstance(init, np.ndarray):
        K, P = init.shape
        if K != n_clusters:
            msg = "Number of centers in provided 'init' ({}) does not match 'n_clusters' ({})"
            raise ValueError(msg.format(K, n_clusters))
        if P != X.shape[1]:
            msg = "Number of features in the provided 'init' ({}) do not match the number of features in 'X'"
            raise ValueError(msg.format(P, X.shape[1]))
        return init
    elif not isinstance(init, str):
        raise TypeError("'init' must be an array or str, got {}".format(type(init)))

    valid = {"k-means||", "k-means++", "random"}

    if isinstance(random_state, Integral) or random_state is None:
        if init == "k-means||":
            random_state = PhiloxRandomState(_agree_seed(random_state))
        else:
            random_state = np.random.RandomState(_agree_seed(random_state) % (2**32))
    if isinstance(random_state, np.random.RandomState) and init == "k-means||":
        random_state = _NumpyRandomState(random_state)

    if init == "k-means||":
        return init_scalable(X, n_clusters, random_state, max_iter, oversampling_factor, n_init)
    elif init == "k-means++":
        return init_pp(X, n_clusters, random_state)
    elif init == "random":
        return init_random(X, n_clusters, random_state)
    else:
        raise ValueError("'init' must be one of {}, got {}".format(valid, init))


def _agree_seed(seed):
    """None -> a fresh seed, the SAME on every rank (rank 0 draws, others receive it)."""
    comm = _comm.default_comm()
    if seed is not None:
        return int(seed)
    s = int(np.random.SeedSequence().generate_state(1, dtype=np.uint64)[0] >> 1)
    if comm.world > 1:
        eng = get_engine()
        t = torch.tensor([s], dtype=torch.int64, device=eng.device)
        comm.broadcast_(t, 0)
        s = int(t.item())
    return s


def _gather_global_rows(X: torch.Tensor, gidx, comm, row0: int) -> torch.Tensor:
    """Rows of the (sharded) X at sorted global indices ``gidx`` -> (m, d) on every rank.
    reference: ``X[sorted(c_idx)].compute()`` (:479), ``X[locs].compute()`` (:500)."""
    eng = get_engine(X.device)
    g = np.asarray(gidx, dtype=np.int64)
    n_local = X.shape[0]
    mine = (g >= row0) & (g < row0 + n_local)
    loc = torch.from_numpy(g[mine] - row0)
    rows = X.rows(loc) if isinstance(X, HostMatrix) else eng.gather(X, loc)
    if comm.world > 1:
        # every rank knows which positions of the (sorted) index list it owns: each fills in its rows, zeros elsewhere,
        # and ONE allreduce puts the (m, d) matrix on every rank -- exact (x + 0 + ... + 0), no exchange of sizes
        full = torch.zeros(len(g), rows.shape[1], dtype=rows.dtype, device=rows.device)
        if mine.any():
            full[torch.from_numpy(np.flatnonzero(mine)).to(rows.device)] = rows
        comm.allreduce_sum_(full)
        rows = full
    return rows


def _search_cumsum(comm, closest: torch.Tensor, vals: np.ndarray, n_total: int) -> np.ndarray:
    """``np.searchsorted(stable_cumsum(closest_dist_sq), vals)`` over the row-sharded vector: the global index of
    the first row whose running total reaches ``vals[j]`` = the number of rows, over all ranks, whose running total
    is still below it (clipped to the last row as scikit-learn does)."""
    cs = torch.cumsum(closest.to(torch.float64), 0)
    total = cs[-1:].clone() if cs.numel() else torch.zeros(1, dtype=torch.float64, device=closest.device)
    offset = 0.0
    if comm.world > 1:
        totals = comm.allgather_varlen(total)
        offset = float(totals[: comm.rank].sum().item())
    v = torch.as_tensor(np.asarray(vals, dtype=np.float64) - offset, device=closest.device)
    counts = torch.searchsorted(cs, v) if cs.numel() else torch.zeros_like(v, dtype=torch.int64)
    counts = counts.to(torch.float64)          # exact below 2^53; the comm shim reduces floating tensors
    comm.allreduce_sum_(counts)
    return np.minimum(counts.cpu().numpy().astype(np.int64), n_total - 1)


def init_pp(X, n_clusters, random_state):
    """K-means initialization using k-means++ (reference :404-423 hands the fully materialised X to scikit-learn's
    ``_kmeans_plusplus``).  Same algorithm and the same draws from ``random_state`` (one uniform for the first
    centre, ``2 + int(log(k))`` uniforms per further centre, candidates by searching the running total of the
    closest squared distances, best candidate by potential), but X stays where it is: the running minimum lives on
    the device, each step is one pass for the trial potentials (``b200km_mindist_pots``) and one for the update
    (K4), row-sharded and streamed like every other pass.  SURVEY section 8(f).4."""
    logger.info("Initializing with k-means++")
    X = _as_device_matrix(X)
    eng = get_engine(X.device)
    comm = _comm.default_comm()
    n_local, d = X.shape
    row0, n_total = comm.row_range(n_local)
    n_local_trials = 2 + int(np.log(n_clusters))
    with _timer("initialization of %2d centers" % n_clusters, _logger=logger):
        center_id = min(int(random_state.random_sample() * n_total), n_total - 1)   # choice(n, p=uniform)
        picked = [_gather_global_rows(X, [center_id], comm, row0)]
        closest = eng._buf("mind2_init", (n_local,), X.dtype)
        pot_t = _mindist_pass(eng, X, picked[0], closest, True)
        comm.allreduce_sum_(pot_t)
        pot = float(pot_t.item())
        for _c in range(1, n_clusters):
            rand_vals = random_state.uniform(size=n_local_trials) * pot
            cand_ids = _search_cumsum(comm, closest, rand_vals, n_total)
            uniq = sorted(set(int(g) for g in cand_ids))
            rows = _gather_global_rows(X, uniq, comm, row0)
            where = {g: i for i, g in enumerate(uniq)}
            cand_rows = rows[[where[int(g)] for g in cand_ids]]
            pots = torch.zeros(len(cand_ids), dtype=torch.float64, device=eng.device)
            for r0, r1, Xs in row_chunks(X):
                if r1 > r0:
                    pots += eng.mindist_pots(Xs, cand_rows, closest[r0:r1])
            comm.allreduce_sum_(pots)
            pots_h = pots.cpu().numpy()
            best = int(np.argmin(pots_h))
            pot = float(pots_h[best])
            _mindist_pass(eng, X, cand_rows[best : best + 1], closest, False)
            picked.append(cand_rows[best : best + 1])
    return torch.cat(picked).cpu().numpy()


@_timed(_logger=logger)
def init_random(X, n_clusters, random_state):
    """K-means initialization using randomly chosen points (reference :426-432)."""
    logger.info("Initializing randomly")
    X = _as_device_matrix(X)
    comm = _comm.default_comm()
    row0, n_total = comm.row_range(X.shape[0])
    idx = sorted(random_state.randint(0, n_total, size=n_clusters))
    return _gather_global_rows(X, idx, comm, row0).cpu().numpy()


def _mindist_pass(eng, X, C, mind2, init: bool) -> torch.Tensor:
    """K4 over all rows of X (resident: one launch sequence; HostMatrix: block by block) -> cost (1-element tensor)."""
    cost = None
    for r0, r1, Xs in row_chunks(X):
        c = eng.mindist_update(Xs, C, mind2[r0:r1], init=init)
        cost = c.clone() if cost is None else cost.add_(c)
    if cost is None:
        cost = torch.zeros(1, dtype=torch.float64, device=eng.device)
    return cost


def evaluate_cost(X, centers):
    """phi_X(C) = sum_i min_j ||x_i - c_j||^2  (reference :516-519).  Returns a Python float
    (the reference returns a lazy scalar that its only caller computes at once, :456)."""
    X = _as_device_matrix(X)
    eng = get_engine(X.device)
    comm = _comm.default_comm()
    C = eng.to_device(np.asarray(centers) if not isinstance(centers, torch.Tensor) else centers, X.dtype)
    mind2 = eng._buf("mind2_init", (X.shape[0],), X.dtype)
    cost = _mindist_pass(eng, X, C, mind2, True)
    comm.allreduce_sum_(cost)
    return float(cost.item())


def _pick(eng, comm, mind2, cost, oversampling_factor, random_state, row0, n_total) -> np.ndarray:
    """Second half of _sample_points (reference :535-541): p = l * d2 / cost ; picked = p > draws."""
    scale = float(oversampling_factor) / cost
    draws_t = None
    if isinstance(random_state, PhiloxRandomState):
        seed, rnd = random_state.seed, random_state.round
        random_state.round += 1
    else:
        # injected host draws (np RandomState, or an object restating dask's RandomState)
        try:
            u = random_state.uniform(size=n_total, chunks=(n_total,))  # dask-style signature
        except TypeError:
            u = random_state.uniform(size=n_total)
        u = np.asarray(u, dtype=np.float64)[row0 : row0 + mind2.numel()]
        seed, rnd, draws_t = 0, 0, eng.to_device(u)
    if hasattr(eng, "bernoulli_buf"):
        # one collective and ONE device->host read per round: every rank hands over [count, picks...] in a
        # fixed-size buffer (no exchange of sizes first); more picks than room -> once more with room (same draws)
        cap = 1024
        while True:
            buf = eng.bernoulli_buf(mind2, scale, seed, rnd, row0, draws=draws_t, capacity=cap)
            h = comm.allgather_fixed(buf).cpu().numpy()
            if int(h[:, 0].max()) <= cap:
                break
            cap = int(h[:, 0].max())
        picked_h = np.concatenate([h[r, 1 : 1 + int(h[r, 0])] for r in range(h.shape[0])])
        out = np.sort(picked_h.astype(np.int64))
        if isinstance(getattr(random_state, "picks", None), list):
            random_state.picks.append(out)
        return out
    picked = eng.bernoulli(mind2, scale, seed, rnd, row0, draws=draws_t)
    if comm.world > 1:
        picked = comm.allgather_varlen(picked)
    out = np.sort(picked.cpu().numpy())
    if isinstance(getattr(random_state, "picks", None), list):
        random_state.picks.append(out)
    return out


def _sample_points(X, centers, oversampling_factor, random_state):
    r"""Sample points independently with probability
    :math:`p_x = \ell \cdot d^2(x, C) / \phi_X(C)` (reference :522-541).
    Returns the sorted GLOBAL row indices (ndarray int64)."""
    X = _as_device_matrix(X)
    eng = get_engine(X.device)
    comm = _comm.default_comm()
    row0, n_total = comm.row_range(X.shape[0])
    C = eng.to_device(np.asarray(centers) if not isinstance(centers, torch.Tensor) else centers, X.dtype)
    mind2 = eng._buf("mind2_init", (X.shape[0],), X.dtype)
    cost = _mindist_pass(eng, X, C, mind2, True)
    comm.allreduce_sum_(cost)
    if isinstance(random_state, np.random.RandomState):
        random_state = _NumpyRandomState(random_state)
    return _pick(eng, comm, mind2, float(cost.item()), oversampling_factor, random_state, row0, n_total)


def candidate_weights(X, candidates):
    """Step 7 of k-means|| (Bahmani et al., algorithm 2): ``w_c`` = number of rows of X whose closest candidate is
    ``c`` -- one K1 pass against the candidate set (float64-exact labels, lowest index on ties, as
    ``pairwise_distances_argmin_min``) and a histogram of the labels (``b200km_label_counts``), summed over the ranks.
    The reference leaves this step out (``k_means.py:481-487``: "scikit-learn doesn't have weighted k-means");
    scikit-learn has had ``sample_weight`` since 0.20, so ``init_scalable(..., weighted=True)`` can use it.

    Returns an ndarray (len(candidates),) float64 of integer counts; their sum is the total number of rows."""
    X = _as_device_matrix(X)
    eng = get_engine(X.device)
    comm = _comm.default_comm()
    C = np.ascontiguousarray(candidates.detach().cpu().numpy() if isinstance(candidates, torch.Tensor) else candidates,
                             dtype=np.float64)
    m = C.shape[0]
    counts = torch.zeros(m, dtype=torch.float64, device=eng.device)
    pack = eng.pack(eng.to_device(C, torch.float64))
    labels = eng._buf("cand_labels", (max(1, min(X.shape[0], getattr(X, "chunk_rows", X.shape[0]))),), torch.int32)
    for r0, r1, Xs in row_chunks(X):
        if r1 > r0:
            ls = labels[: r1 - r0]
            eng.assign(Xs, pack, ls, None, want_mind2=True)
            eng.label_counts(ls, m, counts)
    comm.allreduce_sum_(counts)
    return counts.cpu().numpy()


@_timed(_logger=logger)
def init_scalable(X, n_clusters, random_state=None, max_iter=None, oversampling_factor=2, n_init="auto", weighted=False):
    """K-Means initialization using k-means|| (reference :435-513; algorithm 2 of Bahmani et al.
    as the reference modifies it).

    A running per-row minimum is kept on the device, so each round costs one pass of K4 over the
    rows for the NEW candidates only plus one pass of K5 over the minima; the reference
    recomputes the full n x |C| block matrices every round (:533).

    ``weighted`` (an extension, default off so that the reference's behaviour is untouched): when at least
    ``n_clusters`` candidates were found, re-cluster them WEIGHTED by the number of rows closest to each
    (``candidate_weights``; the paper's steps 7-8) instead of unweighted (:503-513).
    """
    logger.info("Initializing with k-means||")
    X = _as_device_matrix(X)
    eng = get_engine(X.device)
    comm = _comm.default_comm()
    n_local, d = X.shape
    row0, n_total = comm.row_range(n_local)
    if isinstance(random_state, Integral) or random_state is None:
        random_state = PhiloxRandomState(_agree_seed(random_state))
    elif isinstance(random_state, np.random.RandomState):
        random_state = _NumpyRandomState(random_state)

    # Step 1: Initialize Centers (:451-453) -- always global row 0
    idx = 0
    c_idx = {idx}
    new_rows = _gather_global_rows(X, [idx], comm, row0)
    centers_dev = new_rows

    # Step 2: Initialize cost (:456)
    mind2 = eng._buf("mind2_init", (n_local,), X.dtype)
    cost_t = _mindist_pass(eng, X, new_rows, mind2, True)
    comm.allreduce_sum_(cost_t)
    cost = float(cost_t.item())

    if cost == 0:
        n_iter = 0
    else:
        n_iter = int(np.round(np.log(cost)))  # :458-461
    if max_iter is not None:
        n_iter = min(max_iter, n_iter)  # :463-464

    # Steps 3 - 6: update candidate Centers (:467-479)
    for i in range(n_iter):
        with _timer("init iteration %2d/%2d , %2d centers" % (i + 1, n_iter, len(c_idx)), _logger=logger):
            picked = _pick(eng, comm, mind2, cost, oversampling_factor, random_state, row0, n_total)
            fresh = sorted(set(int(v) for v in picked) - c_idx)
            c_idx |= set(fresh)
            if fresh:
                new_rows = _gather_global_rows(X, fresh, comm, row0)
                cost_t = _mindist_pass(eng, X, new_rows, mind2, False)
                comm.allreduce_sum_(cost_t)
                cost = float(cost_t.item())
    centers = _gather_global_rows(X, sorted(c_idx), comm, row0).cpu().numpy()  # :479

    if len(centers) < n_clusters:
        logger.warning("Found fewer than %d clusters in init (found %d).", n_clusters, len(centers))
        # supplement with random (:493-501)
        need = n_clusters - len(centers)
        locs = sorted(int(v) for v in random_state.choice(n_total, need))
        extra = _gather_global_rows(X, locs, comm, row0).cpu().numpy()
        return np.vstack([centers, extra])
    else:
        # Step 7, 8 without weights (:503-513): scikit-learn on the (tiny) candidate set, on the
        # host, exactly as the reference does
        rng2 = random_state.randint(0, np.iinfo("i4").max - 1)
        km = sklearn.cluster.KMeans(n_clusters, random_state=rng2, n_init=n_init)
        if weighted:
            km.fit(centers, sample_weight=candidate_weights(X, centers))
        else:
            km.fit(centers)
    return km.cluster_centers_


# ---------------------------------------------------------------------------------------------
# EM steps
# ---------------------------------------------------------------------------------------------
def _kmeans_single_lloyd(X, n_clusters, max_iter=300, init="k-means||", verbose=False, x_squared_norms=None,
                         random_state=None, tol=1e-4, precompute_distances=True, oversampling_factor=2,
                         init_max_iter=None, n_init="auto"):
    """Lloyd loop (reference :549-619).  Per iteration: pack centres -> K1 assign (+ refine) ->
    K2 sums/counts/inertia -> ONE allreduce -> K3 finalize; the 8-byte shift is the only
    device->host read (the reference's one ``compute()`` per iteration, :602)."""
    X = _as_device_matrix(X)
    eng = get_engine(X.device)
    comm = _comm.default_comm()
    streamed = getattr(X, "_b200_host", None) is not None   # deferred upload (KMeans._check_array(defer=True))
    if not isinstance(init, (np.ndarray, torch.Tensor)):
        if streamed:
            eng.wait_pending(X)                    # the initialisers read all of X: no overlap, scan it now
            streamed = False
            X._b200_bad = eng.nonfinite_flag(X)
        bad = getattr(X, "_b200_bad", None)        # set by fit(): deferred scan, or a multi-rank fit
        if bad is not None:
            X._b200_bad = None
            if any_rank_flag(bad):
                raise ValueError(KMeans._NONFINITE_MSG)
    centers = k_init(X, n_clusters, init=init, oversampling_factor=oversampling_factor, random_state=random_state,
                     max_iter=init_max_iter, n_init=n_init)
    dt = X.dtype
    n_local, P = X.shape
    k = n_clusters
    centers64 = eng.to_device(np.ascontiguousarray(centers), torch.float64)
    new64 = torch.empty_like(centers64)
    # one extra slot behind [sums | counts | inertia]: the NaN/inf flag of a deferred scan rides through iteration 0's
    # allreduce, so every rank learns of a bad value on ANY rank without a collective of its own
    acc_n = eng.acc_len(k, P)
    acc = torch.zeros(acc_n + 1, dtype=torch.float64, device=eng.device)
    shift_t = torch.zeros(1, dtype=torch.float64, device=eng.device)
    labels = torch.empty(n_local, dtype=torch.int32, device=eng.device)
    mind2 = torch.empty(n_local, dtype=dt, device=eng.device)
    pack = None
    shift = None
    i = -1
    pending = None
    if streamed:
        # every small host->device copy of the set-up is queued by now: start the big one behind them
        eng.start_upload(X)
        pending = X._b200_pending
    # k <= 16, d <= 32 on one GPU with X resident: the WHOLE iteration is one launch (b200km_lloyd_small_*: the CTA
    # that finishes last divides, measures the shift and hands acc back zeroed)
    one_launch = (comm.world == 1 and hasattr(eng, "lloyd_small") and not isinstance(X, HostMatrix)
                  and getattr(eng, "can_fuse", lambda *_: False)(k, P) and 0 < n_local < (1 << 30))
    if one_launch:
        ticket = torch.zeros(1, dtype=torch.int32, device=eng.device)
        inertia_one = 0.0
    acc_dirty = False
    last_one = False
    for i in range(max_iter):
        with _timer("Lloyd loop %2d." % i, _logger=logger):
            if one_launch and not pending:
                if acc_dirty:
                    acc.zero_()
                    acc_dirty = False
                # :576-605, every remaining iteration: iteration j + 1 is queued behind iteration j BEFORE j's shift is
                # read (the kernels hand centres, acc and the ticket to one another on the device), so the host's round
                # trip (:602) and the launch latency sit under the next kernel.  A break at j leaves j's centres,
                # labels and distances untouched: three centre buffers and two label / distance buffers rotate; the
                # one queued iteration that is not needed then only costs its time.
                cbuf = [centers64, new64, torch.empty_like(centers64)]
                lbuf = [labels, torch.empty_like(labels)]
                mbuf = [mind2, torch.empty_like(mind2)]

                def _launch(j, i0=i):
                    q = j - i0
                    return eng.lloyd_small_launch(X, cbuf[q % 3], lbuf[q & 1], mbuf[q & 1], acc, ticket,
                                                  cbuf[(q + 1) % 3], slot=q & 1)

                i0, tag = i, _launch(i)
                for i in range(i0, max_iter):
                    nxt = _launch(i + 1) if i + 1 < max_iter else None
                    shift, inertia_one = eng.lloyd_small_wait(tag, slot=(i - i0) & 1)
                    logger.info("Shift: %0.4f", shift)
                    tag = nxt
                    if shift < tol:
                        break
                q = i - i0
                labels, mind2 = lbuf[q & 1], mbuf[q & 1]
                # :606-610 -- a break keeps the centres the last pass was made with, a full run ends on the update
                centers64 = cbuf[q % 3] if shift < tol else cbuf[(q + 1) % 3]
                new64 = cbuf[(q + 2) % 3]
                last_one = True
                break
            last_one = False
            acc_dirty = True
            acc.zero_()
            fused = getattr(eng, "can_fuse", lambda *_: False)(k, P)
            split = not fused and getattr(eng, "lloyd_split", lambda *_: False)(X, k, P)
            if not fused:
                pack = eng.pack(centers64, pack)
            # iteration 0 of a streamed upload runs chunk by chunk behind the copy engine; a HostMatrix is streamed
            # block by block on EVERY pass; a resident X is one span
            if pending:
                spans = ((r0, r1, X[r0:r1], ev) for r0, r1, ev in pending)
            else:
                spans = ((r0, r1, Xs, None) for r0, r1, Xs in row_chunks(X))
            for r0, r1, Xs, ev in spans:
                if r1 <= r0:
                    continue
                if ev is not None:
                    torch.cuda.current_stream(eng.device).wait_event(ev)
                ls, ms = labels[r0:r1], mind2[r0:r1]
                if ev is not None:
                    eng.nonfinite_flag(Xs, X._b200_bad)  # the deferred NaN/inf scan (:195-201)
                if fused:
                    # HBM-bound regime (tiny k, d): labels + sums + counts + inertia in ONE pass over X
                    eng.assign_update(Xs, centers64, ls, ms, acc)  # :576-598, :616
                elif split:
                    # small k, d: K1 decides the labels only; K2 sums ||x - c||^2 (float64) from the rows it reads anyway
                    eng.assign(Xs, pack, ls, ms, lloyd=True)  # :576-579
                    eng.update_inertia(Xs, ls, centers64, k, acc)  # :582-598, :616
                else:
                    eng.assign(Xs, pack, ls, ms)  # :576-579
                    eng.update(Xs, ls, ms, k, acc)  # :582-598, :616
            if pending:
                pending = None
                X._b200_pending = None
            bad = getattr(X, "_b200_bad", None) if i == 0 else None
            if bad is not None:
                X._b200_bad = None
                acc[acc_n:].copy_((bad != 0).to(torch.float64))
            comm.allreduce_sum_(acc)  # the chain-sum over blocks / workers (:595-597)
            eng.finalize(acc, k, P, centers64, new64, shift_t)  # :600-605
            shift = float(shift_t.item())  # :602 -- the one round trip per iteration
            if i == 0 and (bad is not None or comm.world > 1) and float(acc[acc_n].item()) > 0:
                raise ValueError(KMeans._NONFINITE_MSG)  # NaN/inf verdict (:195-201), the same on every rank
            logger.info("Shift: %0.4f", shift)
            if shift < tol:
                break
            centers64, new64 = new64, centers64  # :610

    if shift > 1e-7:  # :612-614 -- labels and NON-squared distances w.r.t. the current centres
        pack = eng.pack(centers64, pack)
        for r0, r1, Xs in row_chunks(X):
            if r1 > r0:
                eng.assign(Xs, pack, labels[r0:r1], mind2[r0:r1])
        inertia_t = eng.reduce_sum(mind2, sqrt=True)
        comm.allreduce_sum_(inertia_t)
    elif last_one:
        inertia_t = torch.tensor([inertia_one], dtype=torch.float64, device=eng.device)
    else:
        inertia_t = acc[acc_n - 1 : acc_n].clone()
    inertia = DeviceArray(inertia_t.reshape(()))  # :616
    centers_out = centers64.to(dt).cpu().numpy()  # :617
    return DeviceArray(labels), inertia, centers_out, i + 1
