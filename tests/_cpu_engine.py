"""TEST-ONLY engine: the CudaEngine interface implemented with the CPU oracle on torch CPU tensors.

Lets ``tests/test_dist_gloo.py`` drive the product's host logic (Lloyd loop, k-means|| driver, row
sharding, the packed allreduce, the variable-length allgather) with world_size = 2 over gloo, on a
machine without a GPU.  It lives under tests/ because it imports the oracle; the product never
imports it (``dask_ml_b200._engine.set_engine_override`` is the only hook)."""
import numpy as np
import torch

from oracle import kmeans_oracle as orc


class _Pack:
    def __init__(self, c):
        self.centers64 = c
        self.k, self.d = c.shape


class NumpyEngine:
    MINDIST_MAX_M = 8

    def __init__(self):
        self.device = torch.device("cpu")
        self.launches = 0
        self.k1_events = None
        self._scratch = {}

    def _buf(self, key, shape, dtype):
        n = int(np.prod(shape))
        t = self._scratch.get(key)
        if t is None or t.dtype != dtype or t.numel() < n:
            t = torch.empty(max(n, 1), dtype=dtype)
            self._scratch[key] = t
        return t[:n].view(*shape)

    def to_device(self, a, dtype=None):
        t = a if isinstance(a, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(a))
        if dtype is not None and t.dtype != dtype:
            t = t.to(dtype)
        return t.contiguous()

    def pack(self, centers64, out=None):
        return _Pack(centers64)

    def assign(self, X, pack, labels=None, mind2=None, want_mind2=True, refine=True):
        # (a CUDA kernel does not raise on a NaN, it computes garbage; scikit-learn would raise here -- on this rank only)
        lab, d2 = orc.pairwise_distances_argmin_min(np.nan_to_num(X.numpy()), pack.centers64.numpy(),
                                                    metric_kwargs={"squared": True})
        lt = torch.from_numpy(lab.astype(np.int32))
        dt = torch.from_numpy(d2).to(X.dtype)
        if labels is not None:
            labels.copy_(lt)
            lt = labels
        if mind2 is not None:
            mind2.copy_(dt)
            dt = mind2
        return lt, (dt if want_mind2 else None)

    def flagged_count(self):
        return 0

    def label_counts(self, labels, k, counts=None):
        if counts is None:
            counts = torch.zeros(k, dtype=torch.float64)
        counts += torch.from_numpy(np.bincount(labels.numpy(), minlength=k).astype(np.float64))
        return counts

    def acc_len(self, k, d):
        return k * d + k + 1

    def update(self, X, labels, mind2, k, acc=None, sqrt=False):
        d = X.shape[1]
        if acc is None:
            acc = torch.zeros(self.acc_len(k, d), dtype=torch.float64)
        sums = orc.centers_dense(X.numpy(), labels.numpy(), k)
        acc[: k * d] += torch.from_numpy(sums.reshape(-1))
        acc[k * d : k * d + k] += torch.from_numpy(np.bincount(labels.numpy(), minlength=k).astype(np.float64))
        if mind2 is not None:
            v = mind2.double()
            acc[k * d + k] += (v.sqrt() if sqrt else v).sum()
        return acc

    def finalize(self, acc, k, d, centers_old, centers_new=None, shift=None):
        sums = acc[: k * d].view(k, d)
        counts = torch.clamp(acc[k * d : k * d + k], min=1.0)
        new = sums / counts[:, None]
        if centers_new is None:
            centers_new = torch.empty(k, d, dtype=torch.float64)
        centers_new.copy_(new)
        if shift is None:
            shift = torch.empty(1, dtype=torch.float64)
        shift[0] = ((centers_old - centers_new) ** 2).sum()
        return centers_new, shift

    def reduce_sum(self, v, sqrt=False, out=None):
        if out is None:
            out = torch.zeros(1, dtype=torch.float64)
        x = v.double()
        out[0] += (x.sqrt() if sqrt else x).sum()
        return out

    def mindist_update(self, X, cand, mind2, init):
        D = ((X.double()[:, None, :] - cand.double()[None]) ** 2).sum(-1).min(1).values.to(X.dtype)
        if init:
            mind2.copy_(D)
        else:
            torch.minimum(mind2, D, out=mind2)
        return mind2.double().sum().reshape(1)

    def bernoulli(self, mind2, scale, seed, round_, row_offset, draws=None, capacity=4096):
        n = mind2.numel()
        u = draws.numpy() if draws is not None else orc.philox_uniform_f64(seed, round_, row_offset, n)
        p = scale * mind2.double().numpy()
        return torch.from_numpy(row_offset + np.where(p > u)[0]).to(torch.int64)

    def gather(self, X, idx_local):
        return X[idx_local.long()].clone()

    def cdist(self, X, Y, squared=False, xnorm2=None, ynorm2=None):
        return torch.from_numpy(orc.euclidean_distances(X.numpy(), Y.numpy(), squared=squared))

    def row_norms(self, X, squared=False):
        return torch.from_numpy(orc.row_norms(X.numpy(), squared=squared))

    def nonfinite_flag(self, X, bad=None):
        if bad is None:
            bad = torch.zeros(1, dtype=torch.int32)
        if not bool(torch.isfinite(X).all()):
            bad += 1
        return bad

    def has_nonfinite(self, X):
        from dask_ml_b200._engine import any_rank_flag

        return any_rank_flag(self.nonfinite_flag(X))
