"""Row-sharded multi-GPU plumbing: one process per GPU, ``torch.distributed`` (NCCL over
NVLink/NVSwitch on the GPU box, gloo in the CPU tests).

The reference has no collectives: its cross-worker reductions are dask graph nodes --
``sum(r.to_delayed().flatten())`` (dask_ml/cluster/k_means.py:595-597), ``da.bincount`` (:598),
``.sum()`` (:519, :534, :616), ``da.where`` (:540).  Here every rank owns a contiguous row range
and the path has exactly these exchange steps:

* Lloyd: ONE allreduce(sum) per iteration of the packed float64 buffer
  ``[k*d sums | k counts | inertia]``; every rank then finalizes redundantly (bit-identical
  centres everywhere, so no broadcast).
* k-means||: one scalar allreduce (cost) and one variable-length allgather (picked global
  indices, then their rows) per round.
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch

__all__ = ["Comm", "NullComm", "TorchDistComm", "default_comm", "local_only"]


class Comm:
    rank = 0
    world = 1

    def allreduce_sum_(self, t: torch.Tensor) -> torch.Tensor:
        raise NotImplementedError

    def allgather_varlen(self, t: torch.Tensor) -> torch.Tensor:
        """Concatenate along axis 0 the (possibly empty, different-length) tensors of all ranks,
        in rank order; every rank gets the full result."""
        raise NotImplementedError

    def broadcast_(self, t: torch.Tensor, src: int = 0) -> torch.Tensor:
        raise NotImplementedError

    def allgather_fixed(self, t: torch.Tensor) -> torch.Tensor:
        """(world, *t.shape): the same-shaped tensors of all ranks, in rank order -- ONE collective, no size exchange."""
        raise NotImplementedError

    def row_range(self, n_local: int) -> Tuple[int, int]:
        """(global offset of this rank's first row, total rows)."""
        raise NotImplementedError


class NullComm(Comm):
    def allreduce_sum_(self, t):
        return t

    def allgather_varlen(self, t):
        return t

    def broadcast_(self, t, src=0):
        return t

    def allgather_fixed(self, t):
        return t[None]

    def row_range(self, n_local):
        return 0, int(n_local)


class TorchDistComm(Comm):
    def __init__(self, group=None):
        import torch.distributed as dist

        self.dist = dist
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)

    def allreduce_sum_(self, t):
        self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM, group=self.group)
        return t

    def broadcast_(self, t, src=0):
        self.dist.broadcast(t, src=src, group=self.group)
        return t

    def allgather_fixed(self, t):
        bufs = [torch.empty_like(t) for _ in range(self.world)]
        self.dist.all_gather(bufs, t.contiguous(), group=self.group)
        return torch.stack(bufs)

    def _counts(self, n: int, device) -> List[int]:
        mine = torch.tensor([n], dtype=torch.int64, device=device)
        allc = [torch.zeros(1, dtype=torch.int64, device=device) for _ in range(self.world)]
        self.dist.all_gather(allc, mine, group=self.group)
        return [int(c.item()) for c in allc]

    def allgather_varlen(self, t):
        counts = self._counts(t.shape[0], t.device)
        mx = max(counts)
        if mx == 0:
            return t
        pad = torch.zeros((mx,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        pad[: t.shape[0]] = t
        bufs = [torch.empty_like(pad) for _ in range(self.world)]
        self.dist.all_gather(bufs, pad, group=self.group)
        return torch.cat([b[:c] for b, c in zip(bufs, counts)], dim=0)

    def row_range(self, n_local):
        dev = torch.device("cuda", torch.cuda.current_device()) if self.dist.get_backend(self.group) == "nccl" else "cpu"
        counts = self._counts(int(n_local), dev)
        return sum(counts[: self.rank]), sum(counts)


_override: Optional[Comm] = None


class local_only:
    """``with local_only(): KMeans(...).fit(X)`` -- inside the block this process fits on its own even though a
    process group is initialised (no collectives are issued; X is the WHOLE data, not a shard).  Used to check a
    sharded fit against a one-rank fit of the same rows in the same job."""

    def __enter__(self):
        global _override
        self._prev, _override = _override, NullComm()
        return self

    def __exit__(self, *exc):
        global _override
        _override = self._prev
        return False


def default_comm() -> Comm:
    """TorchDistComm when a process group is initialised with more than one rank (each rank then
    passes ITS row shard to ``fit``), NullComm otherwise (or inside ``local_only()``)."""
    if _override is not None:
        return _override
    try:
        import torch.distributed as dist

        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            return TorchDistComm()
    except Exception:  # pragma: no cover
        pass
    return NullComm()
