"""FSDP / ZeRO-3 mode (target layer T4; BASELINE.json:5 and config BASELINE.json:10).

Full sharding: every rank owns 1/W of each unit's flat parameter vector together with the
fp32 master copy and both AdamW moments of that slice.  A step looks like:

    forward  : unit i pre-hook   -> all-gather unit i's bf16 shards into its full buffer
               (kernel pulls the W shards straight out of the peers' symmetric memory;
                unit i+1 is prefetched on the comm stream while unit i computes)
    backward : last gradient of unit i lands in its (symmetric) full-gradient buffer ->
               ONE kernel per rank pulls this rank's slice from all peers, reduces it in
               fp32 registers and applies AdamW to the master/moment shard, writing the new
               bf16 parameter shard (K7: reduce-scatter fused with the sharded AdamW - the
               reduced gradient never exists in memory)
    step()   : only advances the device-resident step counter.

``reshard_after_forward=True`` additionally shares two full-parameter buffers between all
units (memory of the two largest units instead of the whole model) and re-gathers a unit
right before its backward, which is the classic ZeRO-3 memory/traffic trade.

On CPU / gloo (tests, BASELINE.json:7-style plumbing) the same schedule runs with
``torch.distributed`` collectives and tensor-op AdamW.  The reference has no FSDP
(SURVEY.md §2.C); the surface follows torch's wrapper where it makes sense
(``module`` attribute, ``full_state_dict`` ~ FULL_STATE_DICT gathering on rank 0).
"""
from __future__ import annotations

import os
from contextlib import contextmanager
from typing import Dict, Iterable, List, Optional, Sequence

import torch
import torch.nn as nn

from . import process_group as pg
from .comm import CommBackend, default_comm
from .ddp import _SingleProcessComm, _align, _register_grad_destination, flat_view


class _Unit:
    """One all-gather / reduce-scatter granule: a list of parameters laid out in one flat vector."""

    def __init__(self, index: int, name: str, params: List[nn.Parameter]):
        self.index, self.name, self.params = index, name, params
        self.offsets: List[int] = []
        n = 0
        for p in params:
            self.offsets.append(n)
            n = _align(n + p.numel(), 64)
        self.numel = n
        self.padded = 0
        self.shard_numel = 0
        # tensors (filled by the wrapper)
        self.shard_param = None   # [padded / W]   compute dtype, symmetric on GPU
        self.master = self.exp_avg = self.exp_avg_sq = None  # fp32 [padded / W]
        self.full_param = None    # [padded]       compute dtype
        self.full_grad = None     # [padded]       compute dtype, symmetric on GPU
        self.pending = 0
        self.gathered_event = None
        self.gathered = False
        self.reduced = False


def default_units(module: nn.Module, unit_classes: Optional[Sequence[type]] = None) -> List[List[nn.Parameter]]:
    """Group parameters into units: every instance of ``unit_classes`` (default: members of any
    ``nn.ModuleList``, i.e. the transformer layers) is a unit, everything else is the root unit."""
    taken = set()
    units: List[List[nn.Parameter]] = []
    names: List[str] = []
    for name, m in module.named_modules():
        is_unit = isinstance(m, tuple(unit_classes)) if unit_classes else False
        if not unit_classes:
            # children of ModuleLists
            parent = name.rsplit(".", 1)[0] if "." in name else ""
            try:
                is_unit = isinstance(module.get_submodule(parent), nn.ModuleList) if name else False
            except AttributeError:
                is_unit = False
        if is_unit:
            ps = [p for p in m.parameters() if p.requires_grad and id(p) not in taken]
            if ps:
                for p in ps:
                    taken.add(id(p))
                units.append(ps)
                names.append(name)
    root = [p for p in module.parameters() if p.requires_grad and id(p) not in taken]
    if root:
        units.insert(0, root)
        names.insert(0, "<root>")
    return units, names


class FullyShardedDataParallel(nn.Module):
    def __init__(self, module: nn.Module, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 0.01,
                 unit_classes: Optional[Sequence[type]] = None, comm="auto", reshard_after_forward: bool = False,
                 prefetch: bool = True):
        super().__init__()
        self.module = module
        params = [p for p in module.parameters() if p.requires_grad]
        if not params:
            raise ValueError("no trainable parameters")
        self.device = params[0].device
        self.dtype = params[0].dtype
        self.cuda = self.device.type == "cuda"
        if isinstance(comm, CommBackend):
            self.comm = comm
        else:
            self.comm = default_comm(self.device, prefer=comm) if pg.is_initialized() else _SingleProcessComm()
        self.world_size, self.rank = self.comm.world_size, self.comm.rank
        self.reshard_after_forward = reshard_after_forward
        self.prefetch = prefetch and self.cuda
        self.defaults = dict(lr=lr, betas=tuple(betas), eps=eps, weight_decay=weight_decay)
        self.hyper = torch.tensor([lr, betas[0], betas[1], eps, weight_decay, 1.0, 0.0, 0.0], dtype=torch.float32, device=self.device)
        self._step = 1
        # bandwidth-bound kernels that also stream the optimizer state: one CTA per SM
        # one comm CTA per SM; on a single GPU the fused kernel is a plain AdamW over ALL parameters and wants two
        # CTAs per SM to cover the HBM latency (measured: 2.8 ms -> see BASELINE.md for 335 M parameters)
        self.comm_blocks = int(os.environ.get("DDL_FSDP_COMM_BLOCKS", "296" if self.world_size == 1 else "148"))
        self._blk = {"num_blocks": 148} if hasattr(self.comm, "reduce_scatter_adamw_") else {}
        self._comm_stream = torch.cuda.Stream(self.device) if self.cuda else None
        self.require_backward_grad_sync = True

        groups, names = default_units(module, unit_classes)
        self.units: List[_Unit] = [_Unit(i, n, ps) for i, (ps, n) in enumerate(zip(groups, names))]
        self._param_unit: Dict[nn.Parameter, _Unit] = {}
        W = self.world_size
        shared_full: List[Optional[torch.Tensor]] = [None, None]
        max_padded = 0
        for u in self.units:
            u.padded = _align(max(u.numel, 64), 1024 * W)
            u.shard_numel = u.padded // W
            max_padded = max(max_padded, u.padded)
        for u in self.units:
            flat = torch.zeros(u.padded, dtype=self.dtype, device=self.device)
            for p, off in zip(u.params, u.offsets):
                flat_view(flat, off, p).copy_(p.data)
            # rank 0's values are the truth (same contract as DDP's init broadcast)
            if W > 1:
                self._broadcast(flat)
            lo = self.rank * u.shard_numel
            u.shard_param = self.comm.alloc_flat(u.shard_numel, self.dtype, self.device)
            u.shard_param.copy_(flat[lo : lo + u.shard_numel])
            u.master = flat[lo : lo + u.shard_numel].float().clone()
            u.exp_avg = torch.zeros(u.shard_numel, dtype=torch.float32, device=self.device)
            u.exp_avg_sq = torch.zeros(u.shard_numel, dtype=torch.float32, device=self.device)
            u.full_grad = self.comm.alloc_flat(u.padded, self.dtype, self.device)
            u.full_grad.zero_()
            if reshard_after_forward and u.index > 0:
                slot = u.index % 2
                if shared_full[slot] is None:
                    shared_full[slot] = torch.zeros(max_padded, dtype=self.dtype, device=self.device)
                u.full_param = shared_full[slot][: u.padded]
            else:
                u.full_param = flat
            u.grad_views, u.param_index = [], {}
            for i, (p, off) in enumerate(zip(u.params, u.offsets)):
                p.data = flat_view(u.full_param, off, p)
                u.grad_views.append(flat_view(u.full_grad, off, p))
                u.param_index[id(p)] = i
                p.grad = u.grad_views[i]
                if self.cuda:
                    _register_grad_destination(p, u.grad_views[i])
                self._param_unit[p] = u
                p.register_post_accumulate_grad_hook(self._grad_hook)
            if self.cuda:
                u.gathered_event = torch.cuda.Event()
        self._install_hooks(names)
        self._callback_queued = False
        self._next_unit = len(self.units) - 1
        self._grads_detached = False

    # ------------------------------------------------------------------ plumbing
    def _broadcast(self, flat: torch.Tensor) -> None:
        chunk = 32 << 20
        for i in range(0, flat.numel(), chunk):
            piece = flat[i : i + chunk]
            self.comm.broadcast_(piece, src=0)

    def _install_hooks(self, names: List[str]) -> None:
        for u in self.units:
            if u.name == "<root>":
                continue
            m = self.module.get_submodule(u.name)
            m.register_forward_pre_hook(lambda mod, args, u=u: self._pre_forward(u))
            if self.reshard_after_forward:
                m.register_full_backward_pre_hook(lambda mod, gout, u=u: self._pre_backward(u))

    # ------------------------------------------------------------------ gather
    def _launch_gather(self, u: _Unit) -> None:
        if u.gathered:
            return
        if self.cuda:
            self._comm_stream.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(self._comm_stream):
                self.comm.all_gather_into(u.full_param, u.shard_param, **self._blk).wait()
                u.gathered_event.record(self._comm_stream)
        else:
            self.comm.all_gather_into(u.full_param, u.shard_param).wait()
        u.gathered = True

    def _wait_gather(self, u: _Unit) -> None:
        if self.cuda:
            torch.cuda.current_stream(self.device).wait_event(u.gathered_event)

    def _pre_forward(self, u: _Unit) -> None:
        self._launch_gather(u)
        self._wait_gather(u)
        if self.prefetch and u.index + 1 < len(self.units):
            self._launch_gather(self.units[u.index + 1])

    def _pre_backward(self, u: _Unit) -> None:
        # shared buffers: the unit's parameters were overwritten by later units' gathers
        u.gathered = False
        self._launch_gather(u)
        self._wait_gather(u)
        if self.prefetch and u.index - 1 >= 1:
            nxt = self.units[u.index - 1]
            nxt.gathered = False
            self._launch_gather(nxt)

    # ------------------------------------------------------------------ forward / backward
    def forward(self, *args, **kwargs):
        for u in self.units:
            u.gathered = False
            u.reduced = False
            u.pending = len(u.params)
        self._callback_queued = False
        self._next_unit = len(self.units) - 1
        root = self.units[0] if self.units and self.units[0].name == "<root>" else None
        if root is not None:
            self._launch_gather(root)
            self._wait_gather(root)
        elif self.units:
            self._launch_gather(self.units[0])
        if self.prefetch and root is not None and len(self.units) > 1:
            self._launch_gather(self.units[1])
        return self.module(*args, **kwargs)

    def _grad_hook(self, p: nn.Parameter) -> None:
        u = self._param_unit[p]
        if not self.require_backward_grad_sync:
            return  # no_sync(): .grad stays where autograd put it and keeps accumulating
        if not self._callback_queued:
            torch.autograd.Variable._execution_engine.queue_callback(self._finalize_backward)
            self._callback_queued = True
        u.pending -= 1
        # Reduce units strictly in backward order (highest index first), whatever order their hooks fired in: the
        # peer-memory kernels of different ranks pair up by launch order, not by unit identity.
        while self._next_unit >= 0:
            nu = self.units[self._next_unit]
            if nu.pending != 0:
                break
            if not nu.reduced:
                self._reduce_and_update(nu)
            self._next_unit -= 1

    @torch.no_grad()
    def _gather_grads(self, u: _Unit) -> None:
        """Move gradients that autograd allocated itself (after ``zero_grad(set_to_none=True)``) into the unit's
        flat buffer with one multi-tensor copy; parameters without a gradient contribute zeros."""
        src, dst = [], []
        for p, v in zip(u.params, u.grad_views):
            g = p.grad
            if g is None:
                if self._grads_detached:
                    v.zero_()
            elif g.data_ptr() != v.data_ptr():
                src.append(g)
                dst.append(v)
            if g is not None or self._grads_detached:
                p.grad = v
        if src:
            torch._foreach_copy_(dst, src)

    def _reduce_and_update(self, u: _Unit) -> None:
        u.reduced = True
        if self.cuda:
            from ..ops import functional as OF

            OF.join_side_work()
        self._gather_grads(u)
        if self.cuda:
            self._comm_stream.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(self._comm_stream):
                self._rs_adamw(u)
        else:
            self._rs_adamw(u)

    def _rs_adamw(self, u: _Unit) -> None:
        if hasattr(self.comm, "reduce_scatter_adamw_") and self.cuda:
            self.comm.reduce_scatter_adamw_(u.full_grad, u.master, u.exp_avg, u.exp_avg_sq, u.shard_param, self.hyper,
                                            num_blocks=self.comm_blocks)
            return
        # plumbing path: unfused reduce-scatter, tensor-op AdamW
        g = torch.empty(u.shard_numel, dtype=u.full_grad.dtype, device=self.device)
        self.comm.reduce_scatter_mean_into(g, u.full_grad).wait()
        lr, (b1, b2), eps, wd = (self.defaults[k] for k in ("lr", "betas", "eps", "weight_decay"))
        t = self._step
        gf = g.float()
        u.exp_avg.mul_(b1).add_(gf, alpha=1 - b1)
        u.exp_avg_sq.mul_(b2).addcmul_(gf, gf, value=1 - b2)
        denom = (u.exp_avg_sq.sqrt() / (1 - b2 ** t) ** 0.5).add_(eps)
        u.master.mul_(1 - lr * wd).addcdiv_(u.exp_avg / (1 - b1 ** t), denom, value=-lr)
        u.shard_param.copy_(u.master)

    def _finalize_backward(self) -> None:
        self._callback_queued = False
        for u in reversed(self.units):
            if not u.reduced:
                self._reduce_and_update(u)
        if self.cuda:
            torch.cuda.current_stream(self.device).wait_stream(self._comm_stream)

    @contextmanager
    def no_sync(self):
        """Gradient accumulation: backward passes inside the context only accumulate into ``.grad``; the next
        synchronised backward reduce-scatters the sum and applies the update (same surface as stock FSDP)."""
        old = self.require_backward_grad_sync
        self.require_backward_grad_sync = False
        try:
            yield
        finally:
            self.require_backward_grad_sync = old

    # ------------------------------------------------------------------ optimizer surface
    @torch.no_grad()
    def zero_grad(self, set_to_none: bool = False) -> None:  # type: ignore[override]
        """``set_to_none=True`` skips the memsets: autograd produces fresh gradients and each unit collects them
        with one multi-tensor copy right before its reduce-scatter (same contract as the DDP wrapper)."""
        self._grads_detached = bool(set_to_none)
        for u in self.units:
            if set_to_none:
                for p in u.params:
                    p.grad = None
            else:
                u.full_grad.zero_()
                for p, v in zip(u.params, u.grad_views):
                    p.grad = v

    @torch.no_grad()
    def step(self) -> None:
        """The parameter update already ran inside backward (fused with the reduce-scatter);
        this only advances the step counter that the bias correction reads."""
        if self.cuda:
            from ..ops import _lib as L

            L.call("ddl_adamw_tick", self.hyper.data_ptr(), L.stream_ptr(self.device))
        else:
            self.hyper[5] += 1
        self._step += 1

    def set_lr(self, lr: float) -> None:
        self.defaults["lr"] = lr
        self.hyper[0:1].fill_(lr)

    # ------------------------------------------------------------------ checkpoints
    @torch.no_grad()
    def gather_all(self) -> None:
        for u in self.units:
            u.gathered = False
            self._launch_gather(u)
            self._wait_gather(u)
        if self.cuda:
            torch.cuda.synchronize(self.device)

    @torch.no_grad()
    def full_state_dict(self) -> Dict[str, torch.Tensor]:
        """Collective: every rank calls; returns the full (unsharded) state dict on CPU.
        Gathers into temporary buffers so that it also works when units share their
        full-parameter storage (``reshard_after_forward``)."""
        saved = []
        for u in self.units:
            tmp = torch.empty(u.padded, dtype=self.dtype, device=self.device)
            self.comm.all_gather_into(tmp, u.shard_param).wait()
            for p, off in zip(u.params, u.offsets):
                saved.append((p, p.data))
                p.data = flat_view(tmp, off, p)
        if self.cuda:
            torch.cuda.synchronize(self.device)
        sd = {k: v.detach().cpu().clone() for k, v in self.module.state_dict().items()}
        for p, d in saved:
            p.data = d
        return sd

    @torch.no_grad()
    def load_full_state_dict(self, state: Dict[str, torch.Tensor], strict: bool = True):
        """Every rank passes the same full state dict; each keeps its own shard (compute-dtype shard + fp32 master;
        the master takes full-precision values from ``state`` where it has them).  Works unit by unit on private
        temporaries, so it is also correct when the units share their full-parameter storage
        (``reshard_after_forward``)."""
        names = {p: n for n, p in self.module.named_parameters()}
        saved, tmps = [], []
        for u in self.units:
            tmp = torch.empty(u.padded, dtype=self.dtype, device=self.device)
            self.comm.all_gather_into(tmp, u.shard_param).wait()  # current values: what a non-strict load keeps
            tmps.append(tmp)
            for p, off in zip(u.params, u.offsets):
                saved.append((p, p.data))
                p.data = flat_view(tmp, off, p)
        if self.cuda:
            torch.cuda.synchronize(self.device)
        res = self.module.load_state_dict(state, strict=strict)
        for u, tmp in zip(self.units, tmps):
            lo = self.rank * u.shard_numel
            full32 = tmp.float()
            for p, off in zip(u.params, u.offsets):
                src = state.get(names.get(p, ""))
                if src is not None and src.dtype == torch.float32 and src.shape == p.shape:
                    flat_view(full32, off, p).copy_(src)
            u.shard_param.copy_(tmp[lo : lo + u.shard_numel])
            u.master.copy_(full32[lo : lo + u.shard_numel])
            u.gathered = False
        for p, d in saved:
            p.data = d
        if not self.reshard_after_forward:
            self.gather_all()  # private full buffers: make them current right away (eval without a forward hook)
        return res

    # ---- optimizer state (the wrapper is its own optimizer) ------------------------------------
    def _gather_fp32(self, shard: torch.Tensor, padded: int) -> torch.Tensor:
        """All ranks' fp32 shards -> one full vector (checkpoint path: control-plane collective, not the peer kernels)."""
        import torch.distributed as dist

        if self.world_size == 1:
            return shard.clone()
        full = torch.empty(padded, dtype=torch.float32, device=shard.device)
        dist.all_gather_into_tensor(full, shard.contiguous())
        return full

    @torch.no_grad()
    def optim_state_dict(self) -> Dict[str, object]:
        """Collective: every rank calls.  Full (unsharded) AdamW state keyed by parameter NAME - fp32 master weights,
        both moments - plus the step counter, so a checkpoint moves between world sizes and between FSDP and DDP
        runs.  (``state_dict()`` of this wrapper is the module's, as for any ``nn.Module``.)"""
        names = {p: n for n, p in self.module.named_parameters()}
        state: Dict[str, Dict[str, torch.Tensor]] = {}
        for u in self.units:
            fulls = {k: self._gather_fp32(getattr(u, k), u.padded) for k in ("master", "exp_avg", "exp_avg_sq")}
            for p, off in zip(u.params, u.offsets):
                state[names[p]] = {k: flat_view(v, off, p).cpu().contiguous().clone() for k, v in fulls.items()}
        return {"state": state, "step": float(self.hyper[5].item()), "defaults": dict(self.defaults), "format": "fsdp/by-name/v1"}

    @torch.no_grad()
    def load_optim_state_dict(self, sd: Dict[str, object]) -> None:
        """Every rank passes the same dict (from :meth:`optim_state_dict`); each keeps its shard.  Also rewrites the
        compute-dtype parameter shard from the restored master."""
        names = {p: n for n, p in self.module.named_parameters()}
        st = sd["state"]
        for u in self.units:
            lo = self.rank * u.shard_numel
            for key in ("master", "exp_avg", "exp_avg_sq"):
                full = torch.zeros(u.padded, dtype=torch.float32, device=self.device)
                for p, off in zip(u.params, u.offsets):
                    flat_view(full, off, p).copy_(st[names[p]][key])
                getattr(u, key).copy_(full[lo : lo + u.shard_numel])
            u.shard_param.copy_(u.master)
            u.gathered = False
        self.defaults.update(sd.get("defaults", {}))
        d = self.defaults
        self.hyper.copy_(torch.tensor([d["lr"], d["betas"][0], d["betas"][1], d["eps"], d["weight_decay"], sd["step"], 0.0, 0.0]))
        self._step = int(sd["step"])
        if not self.reshard_after_forward:
            self.gather_all()

    def extra_repr(self) -> str:
        return f"world_size={self.world_size}, units={len(self.units)}, comm={self.comm.name}, reshard_after_forward={self.reshard_after_forward}"


FSDP = FullyShardedDataParallel
