"""DistributedDataParallel-compatible wrapper (target layer T3).

Behavioural contract taken from stock ``torch.nn.parallel.DistributedDataParallel``
(which BASELINE.json:5 names as the compatibility target; the mounted reference has no
distributed code):

* wraps an ``nn.Module``; ``state_dict()`` keys carry the ``module.`` prefix;
* rank-0 parameters and buffers are broadcast at construction;
* gradients are averaged over ranks, bucket by bucket, while backward is still running;
* ``no_sync()`` accumulates locally; the next synchronised backward reduces the sum;
* gradients are views into flat per-bucket buffers (``gradient_as_bucket_view``).

What is different (B200-first):

* the flat bucket buffers live in *symmetric memory* (every rank maps every peer's
  bucket), and a bucket is reduced by one hand-written kernel that reads the peers'
  copies over NVSwitch, sums in fp32, scales by 1/world and writes the (optionally
  bf16-cast) result - no NCCL call on this path;
* reduction runs on a dedicated comm stream, fenced with CUDA events, so it is
  capturable inside a CUDA graph together with forward/backward/optimizer;
* an optional per-bucket callback (``on_bucket_reduced``) lets the fused AdamW kernel
  run on the comm stream as soon as its bucket is reduced (K13 in SURVEY.md §2.E.3).
"""
from __future__ import annotations

import weakref
from contextlib import contextmanager
from typing import Callable, Dict, Iterable, List, Optional

import torch
import torch.nn as nn

from . import process_group as pg
from .comm import CommBackend, CommHandle, default_comm


class GradBucket:
    """A contiguous slice of gradient memory that is reduced as one message."""

    __slots__ = (
        "index", "params", "offsets", "numel", "dtype", "buffer", "views", "pending",
        "handle", "launched", "padded_numel",
    )

    def __init__(self, index: int, dtype: torch.dtype):
        self.index = index
        self.params: List[nn.Parameter] = []
        self.offsets: List[int] = []
        self.numel = 0
        self.padded_numel = 0
        self.dtype = dtype
        self.buffer: Optional[torch.Tensor] = None
        self.views: List[torch.Tensor] = []
        self.pending = 0
        self.handle: Optional[CommHandle] = None
        self.launched = False


def _align(n: int, a: int) -> int:
    return (n + a - 1) // a * a


def _register_grad_destination(p: torch.Tensor, view: torch.Tensor) -> None:
    """Tell the linear-layer kernels where this parameter's gradient lives (ops/gemm.py::register_grad_destination)."""
    try:
        from ..ops import gemm as G
    except Exception:  # native library not importable (CPU-only box)
        return
    G.register_grad_destination(p, view)


def flat_view(flat: torch.Tensor, off: int, p: torch.Tensor) -> torch.Tensor:
    """View of ``flat[off : off + p.numel()]`` with ``p``'s shape AND memory layout.

    A channels-last 4-D parameter (convolution weights of an NHWC model) keeps its channels-last
    strides inside the flat buffer.  A plain ``.view(p.shape)`` would silently turn it into an NCHW
    tensor and make cuDNN convert the weight (and its gradient) on every step."""
    n = p.numel()
    if p.dim() == 4 and p.is_contiguous(memory_format=torch.channels_last) and not p.is_contiguous():
        co, ci, kh, kw = p.shape
        return flat[off : off + n].view(co, kh, kw, ci).permute(0, 3, 1, 2)
    return flat[off : off + n].view(p.shape)


def assign_buckets(
    params: Iterable[nn.Parameter],
    bucket_cap_bytes: int,
    first_bucket_cap_bytes: Optional[int] = None,
    align_elems: int = 64,
) -> List[GradBucket]:
    """Assign parameters to buckets in *reverse* registration order.

    Gradients become ready roughly in reverse order of ``module.parameters()``, so
    bucket 0 holds the last layers and can be reduced first.  One bucket holds one
    dtype.  Every parameter starts at a multiple of ``align_elems`` elements so that
    16-byte vector accesses in the reduce / optimizer kernels never straddle tensors.
    """
    plist = [p for p in params if p.requires_grad]
    buckets: List[GradBucket] = []
    open_by_dtype: Dict[torch.dtype, GradBucket] = {}
    first_cap = first_bucket_cap_bytes if first_bucket_cap_bytes is not None else bucket_cap_bytes
    for p in reversed(plist):
        b = open_by_dtype.get(p.dtype)
        cap = first_cap if (b is not None and b.index == 0) or (b is None and not buckets) else bucket_cap_bytes
        esize = p.element_size()
        if b is not None and b.numel > 0 and (b.numel + p.numel()) * esize > cap:
            b = None
        if b is None:
            b = GradBucket(len(buckets), p.dtype)
            buckets.append(b)
            open_by_dtype[p.dtype] = b
        b.params.append(p)
        b.offsets.append(b.numel)
        b.numel = _align(b.numel + p.numel(), align_elems)
    for b in buckets:
        b.padded_numel = _align(max(b.numel, align_elems), 1024)
    return buckets


class DistributedDataParallel(nn.Module):
    """Data-parallel wrapper with bucketed, overlapped gradient averaging.

    Parameters
    ----------
    module:             the model replica of this rank (already on its device).
    bucket_cap_mb:      bucket size; 25 MB is the stock default
                        (torch ``nn/parallel/distributed.py``).
    comm:               a :class:`CommBackend`, or "auto" / "p2p" / "torch".
    reduce_dtype:       dtype of the *reduced* gradients handed to the optimizer.  With
                        fp32 gradients and ``torch.bfloat16`` here the cast is fused into
                        the reduce kernel (BASELINE.json:5).
    broadcast_buffers:  re-broadcast module buffers (BN statistics) from rank 0 at every
                        synchronised forward, as stock DDP does.
    flatten_params:     re-point ``param.data`` into one flat buffer per bucket so a fused
                        optimizer can update a whole bucket with one kernel.
    """

    def __init__(
        self,
        module: nn.Module,
        device_ids=None,  # accepted for API compatibility; one device per process
        output_device=None,
        bucket_cap_mb: float = 25.0,
        first_bucket_cap_mb: Optional[float] = None,
        comm="auto",
        reduce_dtype: Optional[torch.dtype] = None,
        broadcast_buffers: bool = True,
        gradient_as_bucket_view: bool = True,
        flatten_params: bool = False,
        init_sync: bool = True,
    ):
        super().__init__()
        self.module = module
        params = [p for p in module.parameters() if p.requires_grad]
        if not params:
            raise ValueError("DistributedDataParallel needs at least one parameter that requires grad")
        self.device = params[0].device
        if isinstance(comm, CommBackend):
            self.comm = comm
        else:
            self.comm = default_comm(self.device, prefer=comm) if pg.is_initialized() else _SingleProcessComm()
        self.world_size = self.comm.world_size
        self.rank = self.comm.rank
        self.broadcast_buffers = broadcast_buffers
        self.require_backward_grad_sync = True
        self.reduce_dtype = reduce_dtype
        self.gradient_as_bucket_view = gradient_as_bucket_view
        self.on_bucket_reduced: Optional[Callable[[GradBucket], None]] = None
        self._callback_queued = False
        self._grads_detached = False  # True after zero_grad(set_to_none=True): buckets are not pre-zeroed
        self._comm_stream = torch.cuda.Stream(device=self.device) if self.device.type == "cuda" else None
        # per-bucket communication timing (SURVEY.md section 5 "metrics"): off by default, see comm_timing() / comm_stats()
        self._comm_timing = False
        self._comm_records: list = []
        self._next_bucket = 0  # buckets are launched strictly in index order (see _grad_ready)
        self._param_listeners: List[Callable] = []
        # lets checkpoint loaders that only see the bare module find the wrapper (utils/checkpoint.py); a plain
        # attribute on the module object, not a submodule / parameter / buffer, so state_dict() is unaffected
        object.__setattr__(module, "_ddl_wrapper", weakref.ref(self))

        self.buckets = assign_buckets(
            params,
            int(bucket_cap_mb * 1024 * 1024),
            None if first_bucket_cap_mb is None else int(first_bucket_cap_mb * 1024 * 1024),
        )
        self._param_to_bucket: Dict[nn.Parameter, tuple] = {}
        self.param_buffers: List[Optional[torch.Tensor]] = []
        for b in self.buckets:
            b.buffer = self.comm.alloc_flat(b.padded_numel, b.dtype, self.device)
            b.buffer.zero_()
            flat_p = None
            if flatten_params:
                flat_p = torch.zeros(b.padded_numel, dtype=b.dtype, device=self.device)
            self.param_buffers.append(flat_p)
            for i, (p, off) in enumerate(zip(b.params, b.offsets)):
                view = flat_view(b.buffer, off, p)
                b.views.append(view)
                self._param_to_bucket[p] = (b, i)
                if flat_p is not None:
                    pv = flat_view(flat_p, off, p)
                    pv.copy_(p.data)
                    p.data = pv
                if gradient_as_bucket_view:
                    p.grad = view
                    if p.is_cuda:
                        _register_grad_destination(p, view)
                p.register_post_accumulate_grad_hook(self._make_hook(p))
        if init_sync:
            self.sync_params_and_buffers()

    # ------------------------------------------------------------------ sync
    @torch.no_grad()
    def sync_params_and_buffers(self) -> None:
        """Rank 0's parameters and buffers overwrite everyone else's (K5)."""
        if self.world_size == 1:
            return
        tensors = [p.data for p in self.module.parameters()] + [b.data for b in self.module.buffers()]
        self._coalesced_broadcast(tensors)

    @torch.no_grad()
    def _sync_buffers(self) -> None:
        bufs = [b.data for b in self.module.buffers()]
        if bufs and self.world_size > 1:
            self._coalesced_broadcast(bufs)

    def _coalesced_broadcast(self, tensors: List[torch.Tensor], chunk_bytes: int = 256 << 20) -> None:
        by_dtype: Dict[torch.dtype, List[torch.Tensor]] = {}
        for t in tensors:
            by_dtype.setdefault(t.dtype, []).append(t)
        for dtype, ts in by_dtype.items():
            group: List[torch.Tensor] = []
            size = 0
            for t in ts + [None]:
                if t is not None and (size + t.numel() * t.element_size() <= chunk_bytes or not group):
                    group.append(t)
                    size += t.numel() * t.element_size()
                    continue
                flat = torch.cat([g.reshape(-1) for g in group])
                self.comm.broadcast_(flat, src=0)
                off = 0
                for g in group:
                    g.copy_(flat[off : off + g.numel()].view(g.shape))
                    off += g.numel()
                group, size = ([t], t.numel() * t.element_size()) if t is not None else ([], 0)

    # ------------------------------------------------------------------ forward / backward
    def forward(self, *args, **kwargs):
        if self.require_backward_grad_sync and torch.is_grad_enabled():
            self._prepare_for_backward()
            if self.broadcast_buffers and self.world_size > 1:
                self._sync_buffers()
        return self.module(*args, **kwargs)

    def _prepare_for_backward(self) -> None:
        for b in self.buckets:
            b.pending = len(b.params)
            b.launched = False
            b.handle = None
        self._callback_queued = False
        self._next_bucket = 0

    # ------------------------------------------------------------------ weights changed behind the optimizer's back
    def register_param_listener(self, fn: Callable) -> None:
        """``fn(state_or_None)`` runs after parameters were overwritten from outside (checkpoint load)."""
        self._param_listeners.append(fn)

    def notify_params_changed(self, state=None) -> None:
        for fn in self._param_listeners:
            fn(state)

    def _make_hook(self, p: nn.Parameter):
        def hook(param: nn.Parameter):
            self._grad_ready(param)

        return hook

    def _grad_ready(self, p: nn.Parameter) -> None:
        b, i = self._param_to_bucket[p]
        if not self.require_backward_grad_sync:
            # no_sync(): leave .grad where autograd put it (the bucket view, or its own tensor after
            # zero_grad(set_to_none=True)); later backward passes accumulate into it
            return
        if not self._callback_queued:
            torch.autograd.Variable._execution_engine.queue_callback(self._finalize_backward)
            self._callback_queued = True
        b.pending -= 1
        # Launch in index order only, as the stock reducer does.  The peer-memory kernels pair up across ranks by
        # per-block epoch counters, not by message identity: if two ranks finished buckets in a different order
        # (rank-dependent control flow, unused parameters) kernels of DIFFERENT buckets would meet.  A bucket that
        # completes early waits (pending == 0) until every lower index has been launched.
        while self._next_bucket < len(self.buckets):
            nb = self.buckets[self._next_bucket]
            if nb.pending != 0:
                break
            if not nb.launched:
                self._launch_bucket(nb)
            self._next_bucket += 1

    @torch.no_grad()
    def _gather_grads(self, b: GradBucket) -> None:
        """Bring every gradient of the bucket into its slot of the flat buffer.

        After ``zero_grad(set_to_none=True)`` autograd hands each parameter a gradient tensor of its
        own; they are moved here with ONE multi-tensor copy per bucket (instead of one memset per
        bucket plus one accumulate kernel per parameter), and ``.grad`` is pointed at the bucket view."""
        src, dst = [], []
        for p, v in zip(b.params, b.views):
            g = p.grad
            if g is None:
                if self._grads_detached:
                    v.zero_()  # unused parameter: contributes zeros, like find_unused_parameters=True
            elif g.data_ptr() != v.data_ptr():
                src.append(g)
                dst.append(v)
            if self.gradient_as_bucket_view and (g is not None or self._grads_detached):
                p.grad = v
        if src:
            torch._foreach_copy_(dst, src)

    def _launch_bucket(self, b: GradBucket) -> None:
        b.launched = True
        if self.device.type == "cuda":
            from ..ops import functional as OF

            OF.join_side_work()  # weight gradients still running on the side stream (ops/functional.py)
        self._gather_grads(b)
        rec = self._comm_record_start(b) if self._comm_timing else None
        if self._comm_stream is not None:
            # the comm stream must see every gradient write of this bucket
            self._comm_stream.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(self._comm_stream):
                if rec is not None:
                    rec["start"].record(self._comm_stream)
                b.handle = self.comm.all_reduce_mean_(b.buffer, self._reduce_out(b))
                if rec is not None:
                    rec["end"].record(self._comm_stream)
                if self.on_bucket_reduced is not None:
                    b.handle.wait()
                    self.on_bucket_reduced(b)
        else:
            b.handle = self.comm.all_reduce_mean_(b.buffer, self._reduce_out(b))

    # ------------------------------------------------------------------ per-bucket communication metrics
    def comm_timing(self, on: bool = True) -> None:
        """Record, for every bucket all-reduce from now on, its size and duration (CUDA events on the comm stream; wall
        clock from launch to completion on the CPU / gloo path).  Not under CUDA-graph capture (events inside a graph
        cannot be timed): :meth:`engine.Trainer.profile_comm` runs a few eager steps for that."""
        self._comm_timing = bool(on)
        self._comm_records = []

    def _comm_record_start(self, b: GradBucket):
        import time

        if self._comm_stream is not None:
            if torch.cuda.is_current_stream_capturing():
                return None
            rec = {"bucket": b.index, "bytes": b.padded_numel * b.buffer.element_size(),
                   "start": torch.cuda.Event(enable_timing=True), "end": torch.cuda.Event(enable_timing=True)}
        else:
            rec = {"bucket": b.index, "bytes": b.padded_numel * b.buffer.element_size(), "t0": time.perf_counter(), "t1": None,
                   "b": b}
        self._comm_records.append(rec)
        return rec

    def comm_stats(self, reset: bool = True):
        """One dict per recorded bucket all-reduce: bucket index, bytes, microseconds, bus bandwidth (algorithm bandwidth x
        2 (W - 1) / W) and - on NVLink - its fraction of the nominal 900 GB/s per direction."""
        import time

        out = []
        W = self.comm.world_size if hasattr(self.comm, "world_size") else pg.world_size()
        f = 2.0 * (W - 1) / W if W > 1 else 0.0
        for r in self._comm_records:
            if "start" in r:
                r["end"].synchronize()
                us = r["start"].elapsed_time(r["end"]) * 1e3
            else:
                us = ((r["t1"] if r["t1"] is not None else time.perf_counter()) - r["t0"]) * 1e6
            bus = r["bytes"] / max(us, 1e-3) * 1e-3 * f  # GB/s
            d = {"bucket": r["bucket"], "bytes": r["bytes"], "us": us, "busbw_gbs": bus}
            if "start" in r:
                d["frac_of_900"] = bus / 900.0
            out.append(d)
        if reset:
            self._comm_records = []
        return out

    def _reduce_out(self, b: GradBucket) -> Optional[torch.Tensor]:
        return None  # in place; subclasses / fused optimizers may redirect

    def _finalize_backward(self) -> None:
        self._callback_queued = False
        for b in self.buckets:
            if not b.launched:
                # some parameter of this bucket got no gradient this iteration: its slot in
                # the (zeroed) bucket contributes zeros, like find_unused_parameters=True
                self._launch_bucket(b)
        if self._comm_stream is not None:
            if self.on_bucket_reduced is None:
                with torch.cuda.stream(self._comm_stream):
                    for b in self.buckets:
                        if b.handle is not None:
                            b.handle.wait()
            torch.cuda.current_stream(self.device).wait_stream(self._comm_stream)
        else:
            import time

            for b in self.buckets:
                if b.handle is not None:
                    b.handle.wait()
                    for r in self._comm_records:
                        if r.get("b") is b and r["t1"] is None:
                            r["t1"] = time.perf_counter()
                    if self.on_bucket_reduced is not None:
                        self.on_bucket_reduced(b)
        for b in self.buckets:
            b.handle = None

    # ------------------------------------------------------------------ user API
    @contextmanager
    def no_sync(self):
        """Skip gradient averaging inside the context (gradient accumulation)."""
        old = self.require_backward_grad_sync
        self.require_backward_grad_sync = False
        try:
            yield
        finally:
            self.require_backward_grad_sync = old

    @torch.no_grad()
    def zero_grad(self, set_to_none: bool = False) -> None:  # type: ignore[override]
        """``set_to_none=False``: zero the flat buckets (one memset per bucket), ``.grad`` stays the
        bucket view and autograd accumulates into it.  ``set_to_none=True``: no memset at all; autograd
        produces fresh gradient tensors and each bucket collects them with one multi-tensor copy when
        it is launched (the cheaper choice when every step is synchronised)."""
        self._grads_detached = bool(set_to_none)
        for b in self.buckets:
            if set_to_none:
                for p in b.params:
                    p.grad = None
            else:
                b.buffer.zero_()
                if self.gradient_as_bucket_view:
                    for p, v in zip(b.params, b.views):
                        p.grad = v

    def flat_grads(self) -> List[torch.Tensor]:
        return [b.buffer for b in self.buckets]

    def extra_repr(self) -> str:
        return f"world_size={self.world_size}, buckets={len(self.buckets)}, comm={self.comm.name}"


class _SingleProcessComm(CommBackend):
    """No process group: world of one (lets the wrapper be used in notebooks/tests)."""

    name = "single"

    def all_reduce_mean_(self, buf, out=None):
        from .comm import _TorchHandle

        if out is not None and out.data_ptr() != buf.data_ptr():
            out.copy_(buf)
        return _TorchHandle(None)

    def broadcast_(self, buf, src=0):
        return None

    def all_gather_into(self, out, shard):
        from .comm import _TorchHandle

        out.copy_(shard.reshape(-1)[: out.numel()])
        return _TorchHandle(None)

    def reduce_scatter_mean_into(self, out_shard, full):
        from .comm import _TorchHandle

        out_shard.copy_(full.reshape(-1)[: out_shard.numel()])
        return _TorchHandle(None)

    def barrier(self):
        return None


# the stock spelling
DDP = DistributedDataParallel
