"""Fused AdamW over the DDP wrapper's flat buckets (K13).

One native kernel per gradient bucket (``csrc/optim.cu``): reads the averaged gradient in
place from the bucket, updates fp32 master weights + moments, writes the bf16 (or fp32)
compute copy the model's parameters are views of.  With ``overlap=True`` each bucket is
updated on the comm stream right after its all-reduce, while backward still runs.

Hyper-parameters live in a device tensor so that LR schedules and the step count work
under CUDA-graph replay.  Math is ``torch.optim.AdamW`` (decoupled weight decay, bias
correction) - that is what the tests compare against.  On CPU (the gloo plumbing config,
BASELINE.json:7) the same update runs as plain tensor ops.
"""
from __future__ import annotations

from typing import Any, Dict, List, Optional

import torch

from .ops import _lib as L
from .parallel.ddp import DistributedDataParallel, GradBucket, flat_view


class FusedAdamW:
    def __init__(self, ddp: DistributedDataParallel, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.01, overlap: bool = False):
        if not isinstance(ddp, DistributedDataParallel):
            raise TypeError("FusedAdamW works on a DistributedDataParallel wrapper (world size 1 is fine): "
                            "it updates the wrapper's flat buckets")
        if any(pb is None for pb in ddp.param_buffers):
            raise ValueError("construct the wrapper with flatten_params=True")
        self.ddp = ddp
        self.device = ddp.device
        self.defaults = dict(lr=lr, betas=tuple(betas), eps=eps, weight_decay=weight_decay)
        self.hyper = torch.tensor([lr, betas[0], betas[1], eps, weight_decay, 1.0, 0.0, 0.0], dtype=torch.float32,
                                  device=self.device)
        self.master: List[torch.Tensor] = []
        self.exp_avg: List[torch.Tensor] = []
        self.exp_avg_sq: List[torch.Tensor] = []
        for b, flat_p in zip(ddp.buckets, ddp.param_buffers):
            self.master.append(flat_p.detach().float().clone())
            self.exp_avg.append(torch.zeros(b.padded_numel, dtype=torch.float32, device=self.device))
            self.exp_avg_sq.append(torch.zeros(b.padded_numel, dtype=torch.float32, device=self.device))
        self.overlap = overlap
        if overlap:
            ddp.on_bucket_reduced = self._step_bucket
        self._cpu_step = 1
        # weights loaded into the wrapped module later (load_checkpoint / load_safetensors without optimizer state)
        # must also replace the fp32 masters, otherwise the next step rewrites the parameters from the stale ones
        ddp.register_param_listener(self.resync_master)

    @torch.no_grad()
    def resync_master(self, state: Optional[Dict[str, torch.Tensor]] = None) -> None:
        """Rebuild the fp32 master copy from the current parameters.  ``state`` (name -> tensor, e.g. the fp32 state
        dict that was just loaded) supplies full-precision values where available; moments and step are kept."""
        names = {p: n for n, p in self.ddp.module.named_parameters()}
        for b, w, flat_p in zip(self.ddp.buckets, self.master, self.ddp.param_buffers):
            w.copy_(flat_p.float())
            if state:
                for p, off in zip(b.params, b.offsets):
                    src = state.get(names.get(p, ""))
                    if src is not None and src.shape == p.shape and src.dtype == torch.float32:
                        flat_view(w, off, p).copy_(src)

    # ---- hyper-parameters (device resident) -------------------------------------------
    def set_lr(self, lr: float) -> None:
        self.defaults["lr"] = lr
        self.hyper[0:1].fill_(lr)

    @property
    def param_groups(self):  # minimal torch.optim surface (schedulers read/write "lr")
        return [self.defaults]

    # ---- update -----------------------------------------------------------------------
    def _step_bucket(self, b: GradBucket) -> None:
        i = b.index
        flat_p = self.ddp.param_buffers[i]
        if self.device.type == "cuda":
            L.call("ddl_adamw", b.buffer.data_ptr(), L.dtype_code(b.buffer.dtype), self.master[i].data_ptr(),
                   self.exp_avg[i].data_ptr(), self.exp_avg_sq[i].data_ptr(), flat_p.data_ptr(), L.dtype_code(flat_p.dtype),
                   b.padded_numel, self.hyper.data_ptr(), 1.0, L.stream_ptr(self.device))
        else:
            self._cpu_update(b.buffer, i, flat_p)

    def _cpu_update(self, grad, i, flat_p) -> None:
        lr, (b1, b2), eps, wd = (self.defaults[k] for k in ("lr", "betas", "eps", "weight_decay"))
        t = self._cpu_step
        g = grad.float()
        m, v, w = self.exp_avg[i], self.exp_avg_sq[i], self.master[i]
        m.mul_(b1).add_(g, alpha=1 - b1)
        v.mul_(b2).addcmul_(g, g, value=1 - b2)
        denom = (v.sqrt() / (1 - b2 ** t) ** 0.5).add_(eps)
        w.mul_(1 - lr * wd).addcdiv_(m / (1 - b1 ** t), denom, value=-lr)
        flat_p.copy_(w)

    @torch.no_grad()
    def step(self) -> None:
        if not self.overlap:
            for b in self.ddp.buckets:
                self._step_bucket(b)
        if self.device.type == "cuda":
            L.call("ddl_adamw_tick", self.hyper.data_ptr(), L.stream_ptr(self.device))
        else:
            self._cpu_step += 1
            self.hyper[5] += 1

    def zero_grad(self, set_to_none: bool = False) -> None:
        self.ddp.zero_grad(set_to_none=set_to_none)

    # ---- checkpoint (per-parameter entries, independent of the bucket layout) ----------
    def state_dict(self) -> Dict[str, Any]:
        index = {p: i for i, p in enumerate(self.ddp.module.parameters())}
        state = {}
        for b, w, m, v in zip(self.ddp.buckets, self.master, self.exp_avg, self.exp_avg_sq):
            for p, off in zip(b.params, b.offsets):
                n = p.numel()
                state[index[p]] = {
                    "master": flat_view(w, off, p).cpu().contiguous().clone(),
                    "exp_avg": flat_view(m, off, p).cpu().contiguous().clone(),
                    "exp_avg_sq": flat_view(v, off, p).cpu().contiguous().clone(),
                }
        return {"state": state, "step": float(self.hyper[5].item()), "defaults": dict(self.defaults)}

    def load_state_dict(self, sd: Dict[str, Any]) -> None:
        index = {p: i for i, p in enumerate(self.ddp.module.parameters())}
        for b, w, m, v, flat_p in zip(self.ddp.buckets, self.master, self.exp_avg, self.exp_avg_sq, self.ddp.param_buffers):
            for p, off in zip(b.params, b.offsets):
                n = p.numel()
                e = sd["state"][index[p]]
                flat_view(w, off, p).copy_(e["master"])
                flat_view(m, off, p).copy_(e["exp_avg"])
                flat_view(v, off, p).copy_(e["exp_avg_sq"])
            flat_p.copy_(w)
        self.defaults.update(sd.get("defaults", {}))
        d = self.defaults
        self.hyper.copy_(torch.tensor([d["lr"], d["betas"][0], d["betas"][1], d["eps"], d["weight_decay"], sd["step"], 0.0, 0.0]))
        self._cpu_step = int(sd["step"])
