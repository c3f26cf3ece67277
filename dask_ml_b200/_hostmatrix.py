"""A row-major matrix that STAYS in host memory and is streamed through the device on every pass.

SURVEY section 8(f).3 ("pinned-host streaming for N larger than aggregate HBM").  The reference's advice for data
that is read several times is ``X.persist()`` (reference dask_ml/cluster/k_means.py:138-140) -- it assumes the
cluster's memory holds X.  Here X may exceed the device's memory: ``HostMatrix`` keeps the caller's host buffer
(page-locked in place, no second copy), and ``chunks()`` yields device views of consecutive row blocks, double
buffered so that the copy of block c+1 runs on the copy engine while the kernels work on block c.  Every pass of
the k-means path (k-means|| rounds, Lloyd iterations, the final labelling) is written as a loop over such
blocks, so the same kernels serve both the resident and the streamed case; a streamed pass is PCIe-bound.

The NaN/inf scan of the reference (``k_means.py:195-201``) rides on the first complete pass.
"""
from __future__ import annotations

import os
from typing import Iterator, Optional, Tuple

import numpy as np
import torch

__all__ = ["HostMatrix", "row_chunks"]


class HostMatrix:
    CHUNK_BYTES = 256 << 20          # per device buffer (two of them); tests shrink it

    def __init__(self, host, eng, nonfinite_msg: Optional[str] = None, share_verdict: bool = False):
        t = host if isinstance(host, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(host))
        if t.is_cuda or t.dim() != 2 or t.dtype not in (torch.float32, torch.float64):
            raise ValueError("HostMatrix wraps a 2-D float32/float64 HOST array")
        self.host = t.contiguous()
        self.eng = eng
        self.shape = tuple(self.host.shape)
        self.dtype = self.host.dtype
        self.device = torch.device(eng.device)
        self.ndim = 2
        self.nonfinite_msg = nonfinite_msg
        self.share_verdict = share_verdict   # fit(): all ranks learn of a NaN / inf on any rank (a collective)
        self._scanned = nonfinite_msg is None
        self._bufs = None
        self._registered = False
        self._pin()
        row_bytes = max(1, self.shape[1] * self.host.element_size())
        env = os.environ.get("B200KM_CHUNK_BYTES")
        budget = int(env) if env else self.CHUNK_BYTES
        self.chunk_rows = max(1, min(max(1, self.shape[0]), budget // row_bytes))

    # ---- host memory ------------------------------------------------------------------------
    def _pin(self) -> None:
        """Page-lock the caller's buffer IN PLACE (cudaHostRegister) so the copies are asynchronous and run at
        full PCIe rate; a refusal (locked-memory limit) only costs speed."""
        if self.host.is_pinned() or self.host.numel() == 0:
            return
        try:
            rc = torch.cuda.cudart().cudaHostRegister(self.host.data_ptr(), self.host.numel() * self.host.element_size(), 0)
            self._registered = int(rc) == 0
        except Exception:  # pragma: no cover - registration is an optimisation
            self._registered = False

    def close(self) -> None:
        if self._registered:
            try:
                torch.cuda.synchronize(self.device)
                torch.cuda.cudart().cudaHostUnregister(self.host.data_ptr())
            except Exception:  # pragma: no cover
                pass
            self._registered = False
        self._bufs = None

    def __del__(self):  # pragma: no cover - best effort
        try:
            self.close()
        except Exception:
            pass

    # ---- tensor-like surface the k-means path uses ---------------------------------------------
    def dim(self) -> int:
        return 2

    def rows(self, idx) -> torch.Tensor:
        """Rows at LOCAL indices ``idx`` (CPU int64 tensor / array) -> (m, d) device tensor."""
        idx = torch.as_tensor(np.asarray(idx, dtype=np.int64))
        return self.host.index_select(0, idx).to(self.device)

    def chunks(self) -> Iterator[Tuple[int, int, torch.Tensor]]:
        """(r0, r1, device view of rows r0..r1) for consecutive row blocks.  The view is valid until the next
        item is requested; work on it must be issued on the current stream."""
        eng, dev = self.eng, self.device
        n, d = self.shape
        rows = self.chunk_rows
        if self._bufs is None or self._bufs[0].shape[0] != rows:
            self._bufs = [torch.empty((rows, d), dtype=self.dtype, device=dev) for _ in range(2)]
        if not hasattr(eng, "_copy_stream"):
            eng._copy_stream = torch.cuda.Stream(device=dev)
        cs = eng._copy_stream
        cur = torch.cuda.current_stream(dev)
        spans = [(r0, min(n, r0 + rows)) for r0 in range(0, n, rows)]
        ready = [None, None]
        free = [None, None]
        scanning = not self._scanned
        bad = torch.zeros(1, dtype=torch.int32, device=dev) if scanning else None

        def issue(c: int) -> None:
            b = c & 1
            r0, r1 = spans[c]
            with torch.cuda.stream(cs):
                if free[b] is not None:
                    cs.wait_event(free[b])
                self._bufs[b][: r1 - r0].copy_(self.host[r0:r1], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(cs)
                ready[b] = ev

        if spans:
            cs.wait_stream(cur)          # earlier work on the compute stream may still read the buffers
            issue(0)
        for c, (r0, r1) in enumerate(spans):
            if c + 1 < len(spans):
                issue(c + 1)             # crosses PCIe while block c is being worked on
            b = c & 1
            cur.wait_event(ready[b])
            view = self._bufs[b][: r1 - r0]
            if scanning:
                eng.nonfinite_flag(view, bad)
            yield r0, r1, view
            ev = torch.cuda.Event()
            ev.record(cur)
            free[b] = ev
        if scanning:
            self._scanned = True
            from ._engine import any_rank_flag

            # fit(): every rank finishes its first pass here and the verdict is shared; predict / transform are not
            # collective calls, there the rank decides alone
            if any_rank_flag(bad) if self.share_verdict else bool(int(bad.item())):
                raise ValueError(self.nonfinite_msg)


def row_chunks(X) -> Iterator[Tuple[int, int, torch.Tensor]]:
    """Row blocks of a resident tensor (one block) or of a HostMatrix (streamed blocks)."""
    if isinstance(X, HostMatrix):
        yield from X.chunks()
    else:
        yield 0, X.shape[0], X
