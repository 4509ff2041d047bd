"""Training engine: the public "train this model data-parallel" API.

``Trainer`` wires together the pieces a ``train_fn`` would otherwise assemble by hand:
model -> bf16 compute copy with fp32 masters -> :class:`DistributedDataParallel` (flat
symmetric-memory buckets, peer-memory all-reduce overlapped with backward) ->
:class:`FusedAdamW` -> optional whole-step CUDA graph (forward, backward, bucket
reductions on the comm stream, optimizer; one ``cudaGraphLaunch`` per step instead of
thousands of kernel launches).  Inputs arrive from pinned host memory through a
double-buffered copy stream; the loss of every step is copied back to pinned host memory.

On CPU (gloo plumbing config, BASELINE.json:7) the same class runs eagerly in fp32.
The reference has no training loop; requirements are BASELINE.json:5,7-10.
"""
from __future__ import annotations

from typing import Any, Callable, Dict, Iterable, List, Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

from .optim import FusedAdamW
from .parallel import process_group as pg
from .parallel.ddp import DistributedDataParallel


def _default_loss(model: nn.Module, batch, fused: bool) -> torch.Tensor:
    if isinstance(batch, dict):
        loss, _ = model(**batch)
        return loss
    x, y = batch
    logits = model(x)
    if fused and logits.is_cuda:
        from .ops import functional as OF

        return OF.cross_entropy(logits, y)
    return F.cross_entropy(logits.float(), y)


class Trainer:
    """One object per rank.

    Parameters
    ----------
    model:        an ``nn.Module`` on CPU (moved/cast here).  Models from :mod:`.models`
                  switch to their fused implementation when ``impl="fused"``.
    lr, betas, eps, weight_decay: AdamW hyper-parameters.
    dtype:        compute dtype on GPU (bf16; masters and moments stay fp32).
    use_graph:    capture the whole step in a CUDA graph after ``graph_warmup`` eager steps.
    comm:         "auto" | "p2p" | "torch" - gradient path (see :func:`..parallel.comm.default_comm`).
    overlap_optimizer: run each bucket's AdamW on the comm stream right after its reduction.
    """

    def __init__(self, model: nn.Module, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 0.01,
                 device: Optional[torch.device] = None, dtype: torch.dtype = torch.bfloat16, impl: str = "fused",
                 use_graph: bool = True, graph_warmup: int = 3, comm: str = "auto", bucket_cap_mb: float = 25.0,
                 overlap_optimizer: bool = True, loss_fn: Optional[Callable[[nn.Module, Any], torch.Tensor]] = None,
                 broadcast_buffers: bool = False, channels_last: bool = True, parallelism: str = "ddp",
                 reshard_after_forward: bool = False):
        self.device = device or pg.device()
        self.cuda = self.device.type == "cuda"
        self.fused = impl == "fused" and self.cuda
        if hasattr(model, "set_impl"):
            model.set_impl("fused" if self.fused else "torch")
        if self.cuda:
            from .models import prepare_for_fused

            model = prepare_for_fused(model, self.device, dtype, channels_last)
        else:
            model = model.to(self.device)
        self.model = model
        self.parallelism = parallelism
        if parallelism == "fsdp":
            # ZeRO-3: sharded parameters/optimizer state; the AdamW update is fused into the
            # gradient reduce-scatter, so the wrapper is also the optimizer
            from .parallel.fsdp import FullyShardedDataParallel

            self.ddp = FullyShardedDataParallel(model, lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, comm=comm,
                                                reshard_after_forward=reshard_after_forward)
            self.opt = self.ddp
        elif parallelism == "ddp":
            self.ddp = DistributedDataParallel(model, bucket_cap_mb=bucket_cap_mb, comm=comm, flatten_params=True,
                                               broadcast_buffers=broadcast_buffers)
            self.opt = FusedAdamW(self.ddp, lr=lr, betas=betas, eps=eps, weight_decay=weight_decay,
                                  overlap=overlap_optimizer and self.cuda)
        else:
            raise ValueError(f"parallelism must be 'ddp' or 'fsdp', got {parallelism!r}")
        self.loss_fn = loss_fn
        self.use_graph = use_graph and self.cuda
        self.graph_warmup = graph_warmup
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self.static_batch = None
        self.static_loss: Optional[torch.Tensor] = None
        self.steps_done = 0
        self._copy_stream = torch.cuda.Stream(self.device) if self.cuda else None
        self._staging: List[Any] = []
        self._stage_events: List[torch.cuda.Event] = []
        self._loss_ring: Optional[torch.Tensor] = None

    # ---- one optimisation step on device-resident inputs ----------------------------------
    def _step_impl(self, batch) -> torch.Tensor:
        # no memset: every step is synchronised, so each bucket collects its fresh gradients with one
        # multi-tensor copy when it is launched (parallel/ddp.py::_gather_grads)
        self.ddp.zero_grad(set_to_none=True)
        loss = self.loss_fn(self.ddp, batch) if self.loss_fn else _default_loss(self.ddp, batch, self.fused)
        loss.backward()
        if self.fused:
            from .ops import functional as OF

            OF.join_side_work()
        self.opt.step()
        if self.fused:
            from .ops import functional as OF

            OF.advance_seed(self.device)
        return loss.detach()

    def _alloc_like(self, batch):
        if isinstance(batch, dict):
            return {k: torch.empty_like(v, device=self.device) for k, v in batch.items()}
        return tuple(torch.empty_like(v, device=self.device) for v in batch)

    @staticmethod
    def _copy_into(dst, src, non_blocking=True):
        if isinstance(src, dict):
            for k, v in src.items():
                dst[k].copy_(v, non_blocking=non_blocking)
        else:
            for d, s in zip(dst, src):
                d.copy_(s, non_blocking=non_blocking)

    def _ensure_static(self, batch) -> None:
        if self.static_batch is None:
            self.static_batch = self._alloc_like(batch)
            self._staging = [self._alloc_like(batch) for _ in range(2)]
            self._stage_events = [torch.cuda.Event() for _ in range(2)]
            self._consumed = [torch.cuda.Event() for _ in range(2)]
            self._loss_ring = torch.zeros(1024, dtype=torch.float32).pin_memory()

    def _training_state(self) -> List[torch.Tensor]:
        """Every tensor a training step mutates: parameters, optimizer state, step counters, module buffers (BatchNorm
        statistics) and the dropout seed word."""
        ts: List[torch.Tensor] = [b for b in self.model.buffers()]
        if self.parallelism == "fsdp":
            for u in self.ddp.units:
                ts += [u.shard_param, u.master, u.exp_avg, u.exp_avg_sq]
            ts.append(self.ddp.hyper)
        else:
            ts += [t for t in self.ddp.param_buffers if t is not None]
            ts += self.opt.master + self.opt.exp_avg + self.opt.exp_avg_sq + [self.opt.hyper]
        if self.fused:
            from .ops import functional as OF

            ts.append(OF.seed_word(self.device))
        return ts

    def _capture(self) -> None:
        # The warm-up steps before capture (allocator / autotune / lazy-init warm-up) run real optimizer steps on the
        # first batch; their effect is rolled back so that a graphed run takes exactly the steps an eager run takes.
        state = self._training_state()
        snap = [t.detach().clone() for t in state]
        py_step = getattr(self.opt, "_step", None)
        s = torch.cuda.Stream(self.device)
        s.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(s):
            for _ in range(self.graph_warmup):
                self._step_impl(self.static_batch)
        torch.cuda.current_stream(self.device).wait_stream(s)
        torch.cuda.synchronize(self.device)
        with torch.no_grad():
            for t, c in zip(state, snap):
                t.copy_(c)
        if py_step is not None:
            self.opt._step = py_step
        if self.parallelism == "fsdp":
            for u in self.ddp.units:
                u.gathered = False
        del snap
        torch.cuda.synchronize(self.device)
        pg.barrier()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.static_loss = self._step_impl(self.static_batch)
        torch.cuda.synchronize(self.device)
        pg.barrier()

    # ---- communication metrics --------------------------------------------------------------
    def profile_comm(self, batch, steps: int = 3) -> dict:
        """Per-bucket all-reduce time and bus bandwidth of this model's gradient buckets (SURVEY.md section 5 "metrics").

        Runs ``steps`` EAGER steps (timed events cannot live inside the whole-step CUDA graph) on ``batch`` and rolls the
        training state back afterwards, like the warm-up of the graph capture.  Returns ``{"buckets": [{bucket, bytes,
        us, busbw_gbs, frac_of_900}, ...] (mean over steps), "comm_us_per_step", "bytes_per_step"}``; DDP only."""
        if self.parallelism != "ddp":
            return {"buckets": [], "note": "per-bucket timing is implemented for the DDP wrapper"}
        dev_batch = batch
        if self.cuda:
            self._ensure_static(batch)
            self._copy_into(self.static_batch, batch, non_blocking=False)
            dev_batch = self.static_batch
        state = self._training_state()
        snap = [t.detach().clone() for t in state]
        py_step = getattr(self.opt, "_step", None)
        self.ddp.comm_timing(True)
        per_bucket: dict = {}
        try:
            for _ in range(steps):
                self._step_impl(dev_batch)
                if self.cuda:
                    torch.cuda.synchronize(self.device)
                for r in self.ddp.comm_stats():
                    per_bucket.setdefault(r["bucket"], []).append(r)
        finally:
            self.ddp.comm_timing(False)
            with torch.no_grad():
                for t, c in zip(state, snap):
                    t.copy_(c)
            if py_step is not None:
                self.opt._step = py_step
        buckets = []
        for k in sorted(per_bucket):
            rs = per_bucket[k]
            d = {"bucket": k, "bytes": rs[0]["bytes"], "us": sum(r["us"] for r in rs) / len(rs),
                 "busbw_gbs": sum(r["busbw_gbs"] for r in rs) / len(rs)}
            if "frac_of_900" in rs[0]:
                d["frac_of_900"] = d["busbw_gbs"] / 900.0
            buckets.append(d)
        return {"buckets": buckets, "comm_us_per_step": sum(b["us"] for b in buckets),
                "bytes_per_step": sum(b["bytes"] for b in buckets), "world": pg.world_size()}

    # ---- public API ------------------------------------------------------------------------
    def step(self, batch) -> torch.Tensor:
        """One step.  ``batch`` holds host (ideally pinned) or device tensors.  Returns the
        loss as a device scalar (no host synchronisation here)."""
        if not self.cuda:
            self.steps_done += 1
            return self._step_impl(batch)
        self._ensure_static(batch)
        slot = self.steps_done % 2
        cur = torch.cuda.current_stream(self.device)
        # host -> staging on the copy stream, staging -> static inputs on the compute stream
        if self.steps_done >= 2:  # this staging slot was last read by step k-2's D2D copy
            self._copy_stream.wait_event(self._consumed[slot])
        with torch.cuda.stream(self._copy_stream):
            self._copy_into(self._staging[slot], batch)
            self._stage_events[slot].record(self._copy_stream)
        cur.wait_event(self._stage_events[slot])
        self._copy_into(self.static_batch, self._staging[slot])
        self._consumed[slot].record(cur)
        if self.use_graph:
            if self.graph is None:
                self._capture()
            self.graph.replay()
            loss = self.static_loss
        else:
            loss = self._step_impl(self.static_batch)
        self._loss_ring[self.steps_done % 1024].copy_(loss, non_blocking=True)
        self.steps_done += 1
        return loss

    # ---- checkpoints --------------------------------------------------------------------------
    def save_checkpoint(self, path: str, extra: Optional[Dict[str, Any]] = None):
        """Rank-0 ``state_dict`` checkpoint with optimizer state and step (collective under FSDP)."""
        from .utils.checkpoint import save_checkpoint

        if self.cuda:
            torch.cuda.synchronize(self.device)
        return save_checkpoint(path, self.ddp, self.opt, step=self.steps_done, extra=extra)

    def load_checkpoint(self, path: str, strict: bool = True, with_optimizer: bool = True) -> Dict[str, Any]:
        """Load weights (+ optimizer state and step when present and wanted) into the live trainer.  A weights-only
        load keeps the moments but re-seeds the fp32 masters from the loaded values."""
        from .utils.checkpoint import load_checkpoint

        meta = load_checkpoint(path, self.ddp, self.opt if with_optimizer else None, strict=strict)
        if with_optimizer:
            self.steps_done = int(meta.get("step", self.steps_done))
        if self.cuda:
            torch.cuda.synchronize(self.device)
        return meta

    def loss_history(self, last: int = 1) -> List[float]:
        """Host values of the last ``last`` losses (synchronises)."""
        if not self.cuda:
            return []
        torch.cuda.synchronize(self.device)
        n = min(last, self.steps_done, 1024)
        return [float(self._loss_ring[(self.steps_done - n + i) % 1024]) for i in range(n)]

    def fit(self, loader: Iterable, steps: Optional[int] = None, log_every: int = 0, logger=None,
            log_comm: bool = False) -> Dict[str, Any]:
        """Run ``steps`` optimisation steps over ``loader``.  With a ``logger`` (``utils.timing.JsonlLogger``) every
        ``log_every``-th step is logged with its loss and the throughput since the previous record; ``log_comm=True`` first logs
        one ``comm`` record with the per-bucket all-reduce time / bytes / bus bandwidth of this model (:meth:`profile_comm`, run
        on the first batch, state rolled back)."""
        import time

        losses = []
        it = iter(loader)
        first = None
        if log_comm and logger is not None and pg.world_size() > 1:
            first = next(it, None)
            if first is not None:
                logger.log(kind="comm", **self.profile_comm(first))
        t_last, n_last = time.perf_counter(), 0

        def batches():
            if first is not None:
                yield first
            yield from it

        for i, batch in enumerate(batches()):
            if steps is not None and i >= steps:
                break
            loss = self.step(batch)
            if not self.cuda:
                losses.append(float(loss))
            if log_every and (i + 1) % log_every == 0 and logger is not None:
                val = self.loss_history(1)[0] if self.cuda else losses[-1]  # (the read-back synchronises: the rate below is real)
                now = time.perf_counter()
                x0 = batch[next(iter(batch))] if isinstance(batch, dict) else batch[0]
                per_step = int(x0.shape[0]) * pg.world_size()
                logger.log(kind="step", step=i + 1, loss=val, samples_per_s=round(per_step * (i + 1 - n_last) / max(now - t_last, 1e-9), 2))
                t_last, n_last = now, i + 1
        if self.cuda:
            losses = self.loss_history(min(self.steps_done, 1024))
        return {"steps": self.steps_done, "losses": losses}
