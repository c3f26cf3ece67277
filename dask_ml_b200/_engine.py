"""Device engine: thin Python wrappers that hand torch CUDA buffers to the C ABI.

torch is used only for device memory, streams and (elsewhere) NCCL; every computation on the
path is one of the ``b200km_*`` entry points.  There is no fallback: constructing the engine
without a CUDA device or without ``libb200km.so`` raises.

The engine interface (``pack / assign / update / finalize / mindist_update / bernoulli / gather /
cdist / row_norms / has_nonfinite / reduce_sum``) is what the host drivers in
``cluster/k_means.py`` and ``metrics/pairwise.py`` program against.
"""
from __future__ import annotations

from typing import Optional, Tuple

import numpy as np
import torch

from . import _lib

__all__ = ["CudaEngine", "get_engine", "CenterPack"]

_MAX_ROWS_PER_CALL = 1 << 30  # the kernels index rows with int32


def _suffix(t: torch.Tensor) -> str:
    if t.dtype == torch.float32:
        return "f32"
    if t.dtype == torch.float64:
        return "f64"
    raise TypeError(f"dask_ml_b200 kernels take float32 or float64, got {t.dtype}")


class CenterPack:
    """Device-side expansion of the (k, d) float64 centres (see b200km_pack_centers)."""

    __slots__ = ("buf", "k", "d", "centers64")

    def __init__(self, buf, k, d, centers64):
        self.buf, self.k, self.d, self.centers64 = buf, k, d, centers64


class CudaEngine:
    def __init__(self, device=None):
        self.lib = _lib.load()
        if not torch.cuda.is_available():
            raise _lib.B200LibraryError(
                "dask_ml_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback."
            )
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        if self.device.type != "cuda":
            raise _lib.B200LibraryError(f"dask_ml_b200 runs on CUDA devices only, got {self.device}")
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        self.launches = 0  # kernels launched through this engine (bench.py reports it)
        self.k1_events = None  # bench.py: set to a list to collect (start, end) CUDA events around K1
        self.pruned_refine = True   # tests flip this to compare with the full float64 scan
        self.f64_on_tensor_cores = True   # float64 X, k > 16: centre + round, tcgen05 decision, float64 re-decision
        self.cdist_on_tensor_cores = True   # float32 X, d <= 128: distance / rbf matrices on the tcgen05 kernel

        self.trace = None      # tools/: set to a list to collect (entry point, event) per launch
        self._scratch = {}

    # ---- plumbing -------------------------------------------------------------------------
    def _stream(self) -> int:
        return torch.cuda.current_stream(self.device).cuda_stream

    def _call(self, name: str, *args, launches: int = 1):
        # the entry points launch on the CURRENT device (and keep per-device state: shared-memory opt-ins, the SM
        # count, the static scratch): make this engine's device current for the call, whatever the caller's is
        if torch.cuda.current_device() != self.device.index:
            with torch.cuda.device(self.device):
                rc = getattr(self.lib, name)(*args)
        else:
            rc = getattr(self.lib, name)(*args)
        _lib.check(rc, name)
        self.launches += launches
        if self.trace is not None:   # tools/: (entry point, CUDA event recorded right after its launch)
            ev = torch.cuda.Event(enable_timing=True)
            ev.record(torch.cuda.current_stream(self.device))
            self.trace.append((name, ev))

    def _buf(self, key, shape, dtype) -> torch.Tensor:
        """Reusable scratch tensor (grown on demand) -- keeps the Lloyd loop allocation-free."""
        n = int(np.prod(shape))
        t = self._scratch.get(key)
        if t is None or t.dtype != dtype or t.numel() < n:
            t = torch.empty(max(n, 1), dtype=dtype, device=self.device)
            self._scratch[key] = t
        return t[:n].view(*shape)

    def to_device(self, a, dtype=None) -> torch.Tensor:
        if isinstance(a, torch.Tensor):
            t = a
        else:
            t = torch.from_numpy(np.ascontiguousarray(a))
        if dtype is not None and t.dtype != dtype:
            t = t.to(dtype)
        return t.to(self.device, non_blocking=True).contiguous()

    FLAG_SCRATCH = 4              # flag buffer = this many int32 per row: room for long candidate lists (pruned re-decision)
    F64_BLOCK_BYTES = 1 << 30     # float32 scratch of the float64-on-tensor-cores path (rows per block = this / 4d)
    STREAM_MIN_BYTES = 64 << 20   # host inputs at least this big are uploaded in overlapped row chunks
    STREAM_CHUNKS = 8

    def to_device_streamed(self, a) -> torch.Tensor:
        """Device tensor for a 2-D HOST array whose upload is deferred: ``start_upload`` later issues the copy in
        row chunks on a side stream and attaches ``_b200_pending = [(r0, r1, event), ...]``; a consumer makes its
        stream wait for a chunk's event and can compute on the first rows while the rest is still crossing
        PCIe (the first Lloyd pass and the NaN/inf scan do exactly that).  The start is deferred so that the
        caller's small host->device copies (initial centres) are queued on the copy engine BEFORE the big
        one -- H2D copies execute in issue order whatever their stream."""
        t = a if isinstance(a, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(a))
        t = t.contiguous()
        dst = torch.empty(t.shape, dtype=t.dtype, device=self.device)
        dst._b200_host = t
        dst._b200_pending = None
        return dst

    def start_upload(self, dst: torch.Tensor) -> None:
        t = getattr(dst, "_b200_host", None)
        if t is None:
            return
        n = t.shape[0]
        if not hasattr(self, "_copy_stream"):
            self._copy_stream = torch.cuda.Stream(device=self.device)
        cs = self._copy_stream
        cs.wait_stream(torch.cuda.current_stream(self.device))
        # (no dst.record_stream: every consumer waits for the chunk events on the compute stream, so the
        #  tensor's last use is ordered after the copies and the caching allocator may recycle it normally)
        pending = []
        step = -(-n // self.STREAM_CHUNKS)
        with torch.cuda.stream(cs):
            for r0 in range(0, n, step):
                r1 = min(n, r0 + step)
                dst[r0:r1].copy_(t[r0:r1], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(cs)
                pending.append((r0, r1, ev))
        dst._b200_pending = pending
        dst._b200_host = None
        dst._b200_src = t   # keep the (possibly temporary) host buffer alive until the copies are done

    def wait_pending(self, X: torch.Tensor) -> None:
        """Start (if needed) and make the current stream wait for every upload chunk of X (no host sync)."""
        self.start_upload(X)
        pending = getattr(X, "_b200_pending", None)
        if pending:
            cur = torch.cuda.current_stream(self.device)
            for _, _, ev in pending:
                cur.wait_event(ev)
            X._b200_pending = None

    @staticmethod
    def _check_matrix(X: torch.Tensor):
        if X.dim() != 2 or X.stride(1) != 1:
            raise ValueError("X must be a row-major 2-D tensor")
        return X.shape[0], X.shape[1], (X.stride(0) if X.shape[0] > 1 else X.shape[1])

    # ---- K1 -------------------------------------------------------------------------------
    def pack(self, centers64: torch.Tensor, out: Optional[CenterPack] = None) -> CenterPack:
        """b200km_pack_centers.  ``centers64``: (k, d) float64 on the device."""
        assert centers64.dtype == torch.float64 and centers64.is_contiguous()
        k, d = centers64.shape
        nbytes = self.lib.b200km_pack_bytes(k, d)
        if out is None or out.k != k or out.d != d:
            buf = torch.empty(nbytes + 256, dtype=torch.uint8, device=self.device)
            off = (-buf.data_ptr()) % 256
            out = CenterPack(buf[off : off + nbytes], k, d, centers64)
        out.centers64 = centers64
        # three kernels: the centring vector, the base pack, and the tcgen05 operand image when d <= 128
        self._call("b200km_pack_centers", centers64.data_ptr(), k, d, out.buf.data_ptr(), self._stream(),
                   launches=3 if d <= 2048 else 2)
        return out

    LLOYD_MAX_D = 128

    def lloyd_split(self, X: torch.Tensor, k: int, d: int) -> bool:
        """True where the Lloyd iteration runs as ``assign(..., lloyd=True)`` + ``update_inertia``: K1 decides the labels
        only and K2 sums the inertia from the rows it reads anyway (float32 X on the tensor-core kernels, d <= 128)."""
        return (X.dtype == torch.float32 and k > 16 and d <= self.LLOYD_MAX_D
                and self.lib.b200km_get_option(_lib.OPT_K1_VARIANT) in (_lib.K1_AUTO, _lib.K1_TCGEN05))

    def assign(self, X: torch.Tensor, pack: CenterPack, labels: Optional[torch.Tensor] = None,
               mind2: Optional[torch.Tensor] = None, want_mind2: bool = True, refine: bool = True,
               lloyd: bool = False) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
        """labels (int32) and squared min distances (X dtype) of every row of X.

        float32 X: tensor-core kernel, then the fp64 re-decision of the flagged near-tie rows, so the
        labels equal what float64 arithmetic gives (reference: sklearn ArgKmin is float64).
        ``lloyd=True`` (float32 X, with ``mind2`` as scratch): labels only -- ``mind2`` comes back unspecified
        (b200km_assign_lloyd_f32); the caller takes the inertia from ``update_inertia``.
        """
        n, d, ldx = self._check_matrix(X)
        if d != pack.d:
            raise ValueError(f"X has {d} features, centres have {pack.d}")
        sfx = _suffix(X)
        if labels is None:
            labels = torch.empty(n, dtype=torch.int32, device=self.device)
        if want_mind2 and mind2 is None:
            mind2 = torch.empty(n, dtype=X.dtype, device=self.device)
        mptr = mind2.data_ptr() if (want_mind2 and mind2 is not None) else None
        st = self._stream()
        for r0 in range(0, n, _MAX_ROWS_PER_CALL):
            nn = min(_MAX_ROWS_PER_CALL, n - r0)
            xs, ls = X[r0 : r0 + nn], labels[r0 : r0 + nn]
            mp = (mind2[r0 : r0 + nn].data_ptr() if mptr is not None else None)
            if sfx == "f32":
                if refine:
                    flen = nn * self.FLAG_SCRATCH if (mptr is not None and self.pruned_refine) else nn
                    flags = self._buf("flag_rows", (flen,), torch.int32)
                    fcnt = self._buf("flag_count", (1,), torch.int32)
                    if self.k1_events is not None:
                        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        e0.record()
                    self._call("b200km_assign_lloyd_f32" if (lloyd and mp is not None) else "b200km_assign_f32",
                               xs.data_ptr(), nn, d, ldx, pack.buf.data_ptr(), pack.k,
                               ls.data_ptr(), mp, flags.data_ptr(), fcnt.data_ptr(), 1, st, launches=1)
                    if self.k1_events is not None:
                        e1.record()
                        self.k1_events.append((e0, e1))
                    # float64 re-decision of the flagged rows; with the winner distances at hand the tensor cores
                    # first list each row's possible winners (4 launches, 2 of them gated off on the device)
                    if mp is not None and self.pruned_refine:
                        self._call("b200km_refine_pruned_len_f64", xs.data_ptr(), nn, d, ldx, pack.buf.data_ptr(), pack.k,
                                   flags.data_ptr(), flen, fcnt.data_ptr(), ls.data_ptr(), mp, st, launches=4)
                    else:
                        self._call("b200km_refine_f64", xs.data_ptr(), nn, d, ldx, pack.buf.data_ptr(), pack.k,
                                   flags.data_ptr(), fcnt.data_ptr(), ls.data_ptr(), mp, st, launches=1)
                else:
                    self._call("b200km_assign_f32", xs.data_ptr(), nn, d, ldx, pack.buf.data_ptr(), pack.k,
                               ls.data_ptr(), mp, None, None, 0, st, launches=1)
            elif self.f64_on_tensor_cores and pack.k > 16 and d <= 2048 and self.lib.b200km_get_option(_lib.OPT_K1_VARIANT) in (_lib.K1_AUTO, _lib.K1_TCGEN05):
                # float64 rows: centred + rounded block by block, decided on the tensor cores with the float32 path's
                # near-tie flags, re-decided in float64 where flagged, distances rebuilt in float64
                if self.k1_events is not None:
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                blk = max(1, min(nn, self.F64_BLOCK_BYTES // (4 * d)))
                xc = self._buf("x_centred", (blk, d), torch.float32)
                flen = blk * self.FLAG_SCRATCH if (refine and self.pruned_refine) else blk
                flags = self._buf("flag_rows", (flen,), torch.int32)
                fcnt = self._buf("flag_count", (1,), torch.int32)
                for b0 in range(0, nn, blk):
                    bn = min(blk, nn - b0)
                    xb, lb = xs[b0 : b0 + bn], ls[b0 : b0 + bn]
                    self._call("b200km_center_convert_f64", xb.data_ptr(), bn, d, ldx, pack.buf.data_ptr(), pack.k,
                               xc.data_ptr(), st, launches=1)
                    d32 = self._buf("dist32", (blk,), torch.float32) if (refine and self.pruned_refine) else None
                    self._call("b200km_assign_centred_f32", xc.data_ptr(), bn, d, d, pack.buf.data_ptr(), pack.k,
                               lb.data_ptr(), d32.data_ptr() if d32 is not None else None, flags.data_ptr(),
                               fcnt.data_ptr(), 1, st, launches=1)
                    if refine and d32 is not None:
                        self._call("b200km_refine_pruned_x64_len_f64", xb.data_ptr(), bn, d, ldx, xc.data_ptr(),
                                   d32.data_ptr(), pack.buf.data_ptr(), pack.k, flags.data_ptr(), flen, fcnt.data_ptr(),
                                   lb.data_ptr(), None, st, launches=4)
                    elif refine:
                        self._call("b200km_refine_x64_f64", xb.data_ptr(), bn, d, ldx, pack.buf.data_ptr(), pack.k,
                                   flags.data_ptr(), fcnt.data_ptr(), lb.data_ptr(), None, st, launches=1)
                    if mp is not None:
                        self._call("b200km_rowdist_f64", xb.data_ptr(), bn, d, ldx, pack.buf.data_ptr(), pack.k,
                                   lb.data_ptr(), mind2[r0 + b0 : r0 + b0 + bn].data_ptr(), st, launches=1)
                if self.k1_events is not None:
                    e1.record()
                    self.k1_events.append((e0, e1))
            else:
                if self.k1_events is not None:
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                self._call("b200km_assign_f64", xs.data_ptr(), nn, d, ldx, pack.buf.data_ptr(), pack.k,
                           ls.data_ptr(), mp, st, launches=1)
                if self.k1_events is not None:
                    e1.record()
                    self.k1_events.append((e0, e1))
        return labels, (mind2 if want_mind2 else None)

    def flagged_count(self) -> int:
        """Rows the last float32 assign sent to the fp64 re-decision (synchronises)."""
        t = self._scratch.get("flag_count")
        return int(t[0].item()) if t is not None else 0

    # ---- K2 / K3 --------------------------------------------------------------------------
    def acc_len(self, k: int, d: int) -> int:
        return int(self.lib.b200km_acc_len(k, d))

    def update(self, X: torch.Tensor, labels: torch.Tensor, mind2: Optional[torch.Tensor], k: int,
               acc: Optional[torch.Tensor] = None, sqrt: bool = False) -> torch.Tensor:
        """acc (float64, [k*d sums | k counts | inertia]) += contribution of these rows."""
        n, d, ldx = self._check_matrix(X)
        if acc is None:
            acc = torch.zeros(self.acc_len(k, d), dtype=torch.float64, device=self.device)
        sfx = _suffix(X)
        st = self._stream()
        for r0 in range(0, n, _MAX_ROWS_PER_CALL):
            nn = min(_MAX_ROWS_PER_CALL, n - r0)
            self._call(f"b200km_update_{sfx}", X[r0 : r0 + nn].data_ptr(), nn, d, ldx, labels[r0 : r0 + nn].data_ptr(),
                       (mind2[r0 : r0 + nn].data_ptr() if mind2 is not None else None), k, acc.data_ptr(),
                       1 if sqrt else 0, st)
        return acc

    def update_inertia(self, X: torch.Tensor, labels: torch.Tensor, centers64: torch.Tensor, k: int,
                       acc: torch.Tensor) -> torch.Tensor:
        """acc += [sums | counts | sum_i ||x_i - centers64[labels_i]||^2], the inertia summed in float64 from the rows."""
        n, d, ldx = self._check_matrix(X)
        assert centers64.dtype == torch.float64 and centers64.is_contiguous() and centers64.shape == (k, d)
        sfx = _suffix(X)
        st = self._stream()
        for r0 in range(0, n, _MAX_ROWS_PER_CALL):
            nn = min(_MAX_ROWS_PER_CALL, n - r0)
            self._call(f"b200km_update_inertia_{sfx}", X[r0 : r0 + nn].data_ptr(), nn, d, ldx,
                       labels[r0 : r0 + nn].data_ptr(), centers64.data_ptr(), k, acc.data_ptr(), st)
        return acc

    def label_counts(self, labels: torch.Tensor, k: int, counts: Optional[torch.Tensor] = None) -> torch.Tensor:
        """counts (float64, k) += histogram of ``labels`` (b200km_label_counts)."""
        if counts is None:
            counts = torch.zeros(k, dtype=torch.float64, device=self.device)
        if labels.numel():
            self._call("b200km_label_counts", labels.data_ptr(), labels.numel(), k, counts.data_ptr(), self._stream())
        return counts

    FUSED_MAX_K, FUSED_MAX_D = 16, 32

    def can_fuse(self, k: int, d: int) -> bool:
        """True when K1 and K2 run as ONE pass over X (the HBM-bound regime: tiny k and d)."""
        return k <= self.FUSED_MAX_K and d <= self.FUSED_MAX_D

    def assign_update(self, X: torch.Tensor, centers64: torch.Tensor, labels: torch.Tensor, mind2: torch.Tensor,
                      acc: torch.Tensor) -> None:
        """b200km_assign_update_*: labels, mind2 and acc += [sums | counts | inertia] in one pass."""
        n, d, ldx = self._check_matrix(X)
        k = centers64.shape[0]
        sfx = _suffix(X)
        st = self._stream()
        for r0 in range(0, n, _MAX_ROWS_PER_CALL):
            nn = min(_MAX_ROWS_PER_CALL, n - r0)
            if self.k1_events is not None:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
            self._call(f"b200km_assign_update_{sfx}", X[r0 : r0 + nn].data_ptr(), nn, d, ldx, centers64.data_ptr(), k,
                       labels[r0 : r0 + nn].data_ptr(), mind2[r0 : r0 + nn].data_ptr(), acc.data_ptr(), st)
            if self.k1_events is not None:
                e1.record()
                self.k1_events.append((e0, e1))

    _LS_SLOTS = 2                 # results in flight (the Lloyd driver keeps the next iteration queued behind this one)

    def lloyd_small_launch(self, X: torch.Tensor, centers64: torch.Tensor, labels: torch.Tensor, mind2: torch.Tensor,
                           acc: torch.Tensor, ticket: torch.Tensor, centers_new: torch.Tensor, slot: int = 0) -> float:
        """b200km_lloyd_small_*: one whole iteration (assign + sums + counts + inertia + divide + shift) in one launch.
        ``acc`` / ``ticket`` (int32[1]) must be zero when the kernel STARTS and come back zero (so launches may be
        queued back to back); single GPU, X resident (n < 2^30).  The kernel writes [shift, inertia] into page-locked
        host memory (one 64-byte slot per ``slot``) followed by this launch's tag; returns the tag for
        ``lloyd_small_wait``."""
        n, d, ldx = self._check_matrix(X)
        if not hasattr(self, "_ls_host"):
            self._ls_host = torch.zeros(8 * self._LS_SLOTS, dtype=torch.float64).pin_memory()
            self._ls_view = self._ls_host.numpy()
            self._ls_tag = 0.0
        self._ls_tag += 1.0
        tag = self._ls_tag
        if self.k1_events is not None:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        self._call(f"b200km_lloyd_small_{_suffix(X)}", X.data_ptr(), n, d, ldx, centers64.data_ptr(), centers64.shape[0],
                   labels.data_ptr(), mind2.data_ptr(), acc.data_ptr(), ticket.data_ptr(), centers_new.data_ptr(),
                   self._ls_host.data_ptr() + 64 * slot, tag, self._stream())
        if self.k1_events is not None:
            e1.record()
            self.k1_events.append((e0, e1))
        return tag

    def lloyd_small_wait(self, tag: float, slot: int = 0) -> Tuple[float, float]:
        """(shift, inertia) of the launch that carries ``tag``: the host polls the slot's tag word instead of
        synchronising the stream (falls back to a synchronisation after 2 s)."""
        view = self._ls_view
        o = 8 * slot
        deadline = None
        spins = 0
        while view[o + 2] != tag:
            spins += 1
            if (spins & 0xFFFF) == 0:
                import time

                now = time.perf_counter()
                if deadline is None:
                    deadline = now + 2.0
                elif now > deadline:      # a failed launch never writes the tag: let the synchronisation report it
                    torch.cuda.current_stream(self.device).synchronize()
                    if view[o + 2] != tag:
                        raise _lib.B200KernelError("b200km_lloyd_small: the iteration's result never arrived")
        return float(view[o]), float(view[o + 1])

    def lloyd_small(self, X: torch.Tensor, centers64: torch.Tensor, labels: torch.Tensor, mind2: torch.Tensor,
                    acc: torch.Tensor, ticket: torch.Tensor, centers_new: torch.Tensor) -> Tuple[float, float]:
        """One iteration, launched and awaited: (shift, inertia)."""
        return self.lloyd_small_wait(self.lloyd_small_launch(X, centers64, labels, mind2, acc, ticket, centers_new))

    def finalize(self, acc: torch.Tensor, k: int, d: int, centers_old: torch.Tensor,
                 centers_new: Optional[torch.Tensor] = None, shift: Optional[torch.Tensor] = None):
        if centers_new is None:
            centers_new = torch.empty(k, d, dtype=torch.float64, device=self.device)
        if shift is None:
            shift = torch.empty(1, dtype=torch.float64, device=self.device)
        self._call("b200km_finalize", acc.data_ptr(), k, d, centers_old.data_ptr(), centers_new.data_ptr(),
                   shift.data_ptr(), self._stream())
        return centers_new, shift

    def reduce_sum(self, v: torch.Tensor, sqrt: bool = False, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        if out is None:
            out = torch.zeros(1, dtype=torch.float64, device=self.device)
        if v.numel():
            self._call(f"b200km_reduce_sum_{_suffix(v)}", v.data_ptr(), v.numel(), 1 if sqrt else 0, out.data_ptr(),
                       self._stream())
        return out

    # ---- K4 .. K6 -------------------------------------------------------------------------
    MINDIST_MAX_M = 8
    MINDIST_K1_MIN_M = 48        # from this many new candidates on, one K1 pass beats m / 8 passes of K4
    # b200km_mindist_tc_f32 (float32, d <= 128): one pass for up to 128 candidates at 11 ms per 6e7 x 64 rows whatever
    # their number; a pass of the 8-candidate kernel is 3.7 ms (one thread per row), so the tensor-core pass pays from
    # 25 candidates on (40 candidates: 18.7 -> 11.3 ms)
    MINDIST_TC_MIN_M, MINDIST_TC_MAX_M = 25, 128

    def mindist_update(self, X: torch.Tensor, cand: torch.Tensor, mind2: torch.Tensor, init: bool) -> torch.Tensor:
        """Running min of squared distances to ``cand`` rows; returns the (1,) float64 local cost."""
        n, d, ldx = self._check_matrix(X)
        sfx = _suffix(X)
        m = cand.shape[0]
        cost = torch.zeros(1, dtype=torch.float64, device=self.device)
        st = self._stream()
        if (self.MINDIST_TC_MIN_M <= m <= self.MINDIST_TC_MAX_M and n and X.dtype == torch.float32
                and d <= self.CDIST_TC_MAX_D and hasattr(self.lib, "b200km_mindist_tc_f32")
                and self.lib.b200km_get_option(_lib.OPT_K1_VARIANT) in (_lib.K1_AUTO, _lib.K1_TCGEN05)
                and self.lib.b200km_get_option(_lib.OPT_K4_VARIANT) != 1):
            # 25 .. 128 new candidates: ONE pass on the tensor cores (K7's kernel with a min-merge epilogue) instead of
            # m / 8 passes of an FFMA kernel that is bound by its arithmetic from three candidates on
            pack = self.pack(cand.to(torch.float64).contiguous())
            for r0 in range(0, n, _MAX_ROWS_PER_CALL):
                nn = min(_MAX_ROWS_PER_CALL, n - r0)
                self._call("b200km_mindist_tc_f32", X[r0 : r0 + nn].data_ptr(), nn, d, ldx, pack.buf.data_ptr(), m,
                           mind2[r0 : r0 + nn].data_ptr(), 1 if init else 0, cost.data_ptr(), st)
            return cost
        if m >= self.MINDIST_K1_MIN_M and n:
            # many new candidates (a large oversampling factor): ONE K1 pass against all of them instead of m / 8
            # passes of K4, then the merge into the running minimum
            pack = self.pack(cand.to(torch.float64).contiguous())
            lab = self._buf("mindist_labels", (n,), torch.int32)
            other = self._buf("mindist_other", (n,), X.dtype)
            self.assign(X, pack, lab, other)
            for r0 in range(0, n, _MAX_ROWS_PER_CALL):
                nn = min(_MAX_ROWS_PER_CALL, n - r0)
                self._call(f"b200km_mindist_merge_{sfx}", other[r0 : r0 + nn].data_ptr(), nn,
                           mind2[r0 : r0 + nn].data_ptr(), 1 if init else 0, cost.data_ptr(), st)
            return cost
        cand = cand.to(X.dtype).contiguous()
        per = max(1, min(self.MINDIST_MAX_M, (96 * 1024) // (d * X.element_size())))
        first = init
        for j0 in range(0, m, per):
            cj = cand[j0 : j0 + per]
            cost.zero_()
            for r0 in range(0, n, _MAX_ROWS_PER_CALL):
                nn = min(_MAX_ROWS_PER_CALL, n - r0)
                self._call(f"b200km_mindist_update_{sfx}", X[r0 : r0 + nn].data_ptr(), nn, d, ldx, cj.data_ptr(),
                           cj.shape[0], mind2[r0 : r0 + nn].data_ptr(), 1 if first else 0, cost.data_ptr(), st)
            first = False
        return cost

    def mindist_pots(self, X: torch.Tensor, cand: torch.Tensor, mind2: torch.Tensor) -> torch.Tensor:
        """k-means++ trial potentials: (m,) float64, pots[j] = sum_i min(mind2[i], ||x_i - cand_j||^2)."""
        n, d, ldx = self._check_matrix(X)
        sfx = _suffix(X)
        cand = cand.to(X.dtype).contiguous()
        m = cand.shape[0]
        per = max(1, min(self.MINDIST_MAX_M, (96 * 1024) // (d * X.element_size())))
        pots = torch.zeros(m, dtype=torch.float64, device=self.device)
        st = self._stream()
        for j0 in range(0, m, per):
            cj = cand[j0 : j0 + per]
            for r0 in range(0, n, _MAX_ROWS_PER_CALL):
                nn = min(_MAX_ROWS_PER_CALL, n - r0)
                self._call(f"b200km_mindist_pots_{sfx}", X[r0 : r0 + nn].data_ptr(), nn, d, ldx, cj.data_ptr(),
                           cj.shape[0], mind2[r0 : r0 + nn].data_ptr(), pots[j0 : j0 + per].data_ptr(), st)
        return pots

    def bernoulli(self, mind2: torch.Tensor, scale: float, seed: int, round_: int, row_offset: int,
                  draws: Optional[torch.Tensor] = None, capacity: int = 4096) -> torch.Tensor:
        """Global indices (int64, unsorted) of the rows picked this round."""
        n = mind2.numel()
        sfx = _suffix(mind2)
        st = self._stream()
        while True:
            picked = self._buf("picked", (capacity,), torch.int64)
            cnt = self._buf("pick_count", (1,), torch.int32)
            cnt.zero_()
            self._call(f"b200km_bernoulli_pick_{sfx}", mind2.data_ptr(), n, float(scale), int(seed) & (2**64 - 1),
                       int(round_) & 0xFFFFFFFF, int(row_offset), (draws.data_ptr() if draws is not None else None),
                       picked.data_ptr(), cnt.data_ptr(), capacity, st)
            c = int(cnt.item())
            if c <= capacity:
                return picked[:c].clone()
            capacity = c  # rare: more picks than room -- redo with the exact size (same draws)

    def bernoulli_buf(self, mind2: torch.Tensor, scale: float, seed: int, round_: int, row_offset: int,
                      draws: Optional[torch.Tensor] = None, capacity: int = 1024) -> torch.Tensor:
        """The same pick WITHOUT a host synchronisation: a device int64 vector ``[count, pick_0, ..., pick_{cap-1}]``
        (``count`` may exceed ``capacity``: the caller then asks again with more room -- same draws)."""
        n = mind2.numel()
        buf = self._buf("picked_buf", (capacity + 1,), torch.int64)
        buf[:1].zero_()
        self._call(f"b200km_bernoulli_pick_{_suffix(mind2)}", mind2.data_ptr(), n, float(scale),
                   int(seed) & (2**64 - 1), int(round_) & 0xFFFFFFFF, int(row_offset),
                   (draws.data_ptr() if draws is not None else None), buf[1:].data_ptr(),
                   buf.data_ptr(),          # the int32 counter = the low half of buf[0] (little endian)
                   capacity, self._stream())
        return buf

    def gather(self, X: torch.Tensor, idx_local: torch.Tensor) -> torch.Tensor:
        n, d, ldx = self._check_matrix(X)
        m = idx_local.numel()
        out = torch.empty(m, d, dtype=X.dtype, device=self.device)
        if m:
            idx_local = idx_local.to(device=self.device, dtype=torch.int64).contiguous()
            self._call(f"b200km_gather_rows_{_suffix(X)}", X.data_ptr(), n, d, ldx, idx_local.data_ptr(), m,
                       out.data_ptr(), self._stream())
        return out

    # ---- K7 + helpers ---------------------------------------------------------------------
    CDIST_TC_MAX_D = 128
    CDIST_TC_MAX_K = 1 << 16     # centres of one tensor-core call (the pack holds five copies of them)

    def _cdist_tc(self, X: torch.Tensor, Y: torch.Tensor, mode: int, gamma: float = 0.0) -> Optional[torch.Tensor]:
        """K7 on the tensor cores (b200km_cdist_pack_f32) where it applies -- float32, <= 128 features, the one-CTA
        pack layout -- else None (the caller takes the FFMA tile kernel).  The kernel writes (rows x centres) with
        the centres as the packed side; when Y has more rows than X (SpectralClustering's B block: 100 kept rows
        against millions) the roles are swapped -- distances and the rbf kernel are symmetric -- and the result is
        the transposed view of what was written."""
        n, d, _ = self._check_matrix(X)
        k = Y.shape[0]
        if (X.dtype != torch.float32 or Y.dtype != torch.float32 or d > self.CDIST_TC_MAX_D
                or not self.cdist_on_tensor_cores or n == 0 or k == 0
                or self.lib.b200km_get_option(_lib.OPT_K1_VARIANT) == _lib.K1_TCGEN05_PAIR):
            return None
        swap = k > n
        if swap:
            X, Y, n, k = Y, X, k, n
            if X.stride(1) != 1:
                X = X.contiguous()
        if k > self.CDIST_TC_MAX_K:
            return None
        ldx = X.stride(0) if n > 1 else d
        pack = self.pack(Y.to(torch.float64).contiguous())
        out = torch.empty(n, k, dtype=torch.float32, device=self.device)
        for r0 in range(0, n, _MAX_ROWS_PER_CALL):
            nn = min(_MAX_ROWS_PER_CALL, n - r0)
            self._call("b200km_cdist_pack_f32", X[r0 : r0 + nn].data_ptr(), nn, d, ldx, pack.buf.data_ptr(), pack.k,
                       out[r0 : r0 + nn].data_ptr(), out.stride(0), mode, float(gamma), self._stream())
        return out.t() if swap else out

    def cdist(self, X: torch.Tensor, Y: torch.Tensor, squared: bool = False, xnorm2=None, ynorm2=None) -> torch.Tensor:
        n, d, ldx = self._check_matrix(X)
        Y = Y.contiguous()
        k = Y.shape[0]
        if xnorm2 is None and ynorm2 is None:    # (caller-supplied norms are honoured as given: the FFMA kernel below)
            res = self._cdist_tc(X, Y, _lib.CDIST_SQUARED if squared else _lib.CDIST_DIST)
            if res is not None:
                return res
        out = torch.empty(n, k, dtype=X.dtype, device=self.device)
        if n and k:
            self._call(f"b200km_cdist_{_suffix(X)}", X.data_ptr(), n, d, ldx, Y.data_ptr(), k,
                       (xnorm2.data_ptr() if xnorm2 is not None else None),
                       (ynorm2.data_ptr() if ynorm2 is not None else None), out.data_ptr(), 1 if squared else 0,
                       self._stream())
        return out

    KERNEL_KINDS = {"rbf": 2, "linear": 3, "polynomial": 4, "sigmoid": 5}

    def pairwise_kernel(self, X: torch.Tensor, Y: torch.Tensor, kind: str, gamma: float = 1.0, coef0: float = 1.0,
                        degree: int = 3) -> torch.Tensor:
        """(n, k) kernel matrix between the rows of X and Y (b200km_pairwise_kernel_*)."""
        n, d, ldx = self._check_matrix(X)
        Y = Y.contiguous()
        k = Y.shape[0]
        if kind == "rbf":
            res = self._cdist_tc(X, Y, _lib.CDIST_RBF, gamma)
            if res is not None:
                return res
        out = torch.empty(n, k, dtype=X.dtype, device=self.device)
        if n and k:
            self._call(f"b200km_pairwise_kernel_{_suffix(X)}", X.data_ptr(), n, d, ldx, Y.data_ptr(), k, out.data_ptr(),
                       self.KERNEL_KINDS[kind], float(gamma), float(coef0), int(degree), self._stream())
        return out

    def row_norms(self, X: torch.Tensor, squared: bool = False) -> torch.Tensor:
        n, d, ldx = self._check_matrix(X)
        out = torch.empty(n, dtype=X.dtype, device=self.device)
        if n:
            self._call(f"b200km_row_norms_{_suffix(X)}", X.data_ptr(), n, d, ldx, out.data_ptr(), 1 if squared else 0,
                       self._stream())
        return out

    def nonfinite_flag(self, X: torch.Tensor, bad: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Device flag (int32[1]) that becomes non-zero when X holds a NaN / inf; no host sync."""
        n, d, ldx = self._check_matrix(X)
        if bad is None:
            bad = torch.zeros(1, dtype=torch.int32, device=self.device)
        if n and d:
            self._call(f"b200km_count_nonfinite_{_suffix(X)}", X.data_ptr(), n, d, ldx, bad.data_ptr(), self._stream())
        return bad

    def has_nonfinite(self, X: torch.Tensor) -> bool:
        """Does X -- or, in a multi-rank fit, ANY rank's shard -- hold a NaN / inf?  The verdict is shared so that
        every rank raises the reference's ValueError together instead of one rank leaving the others in a collective."""
        return any_rank_flag(self.nonfinite_flag(X))

    def make_blobs(self, n: int, d: int, centers: torch.Tensor, cluster_std: float = 1.0, seed: int = 0,
                   row_offset: int = 0, dtype=torch.float32, out: Optional[torch.Tensor] = None,
                   y: Optional[torch.Tensor] = None) -> torch.Tensor:
        if out is None:
            out = torch.empty(n, d, dtype=dtype, device=self.device)
        centers = centers.to(device=self.device, dtype=out.dtype).contiguous()
        self._call(f"b200km_make_blobs_{_suffix(out)}", out.data_ptr(), n, d, d, centers.data_ptr(), centers.shape[0],
                   float(cluster_std), int(seed) & (2**64 - 1), int(row_offset), (y.data_ptr() if y is not None else None),
                   self._stream())
        return out


def any_rank_flag(flag: torch.Tensor) -> bool:
    """``flag`` (int32[1], non-zero = this rank saw it) -> the OR over all ranks, as a Python bool (synchronises)."""
    from . import _comm

    comm = _comm.default_comm()
    if comm.world > 1:
        f = (flag != 0).to(torch.float64)   # the comm shim reduces floating tensors
        comm.allreduce_sum_(f)
        return bool(f.item() > 0)
    return bool(int(flag.item()))


_engines = {}
_override = None


def set_engine_override(engine) -> None:
    """TEST HOOK ONLY: make ``get_engine`` return ``engine`` (an object with the CudaEngine interface)
    so the host drivers and the collective plumbing can be exercised on CPU with gloo.  The product
    never sets it; ``None`` restores the CUDA engine."""
    global _override
    _override = engine


def get_engine(device=None) -> CudaEngine:
    """Per-device engine singleton.  Raises (never falls back) without CUDA / the library."""
    if _override is not None:
        return _override
    if isinstance(device, torch.device) and device.type == "cuda" and device.index is None:
        device = None
    key = str(device) if device is not None else (f"cuda:{torch.cuda.current_device()}" if torch.cuda.is_available() else "none")
    eng = _engines.get(key)
    if eng is None:
        eng = CudaEngine(device)
        _engines[key] = eng
    return eng
