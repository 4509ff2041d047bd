"""Checkpoint save / load (target layer T8, SURVEY.md §1.2, §5 "Checkpoint / resume").

What the reference does: it only *saves* - a whole-module pickle
(notebooks/cv/onnx_experiments.py:198), a TorchScript archive (:215) and an ONNX file
with embedded weights (:33-42) - and never resumes.  BASELINE.json:5 asks for "rank-0
torch state_dict / HF safetensors checkpoints"; both are here, with load + resume.

* :func:`save_checkpoint` / :func:`load_checkpoint`: rank 0 writes
  ``{"model": state_dict, "optimizer": ..., "step": ..., "extra": ...}`` with
  ``torch.save``; every rank loads it (map_location = its own device).  The ``module.``
  prefix a DDP wrapper adds is stripped on save and tolerated on load, so a checkpoint
  is interchangeable between wrapped and bare models (and with stock DDP).
* :func:`save_safetensors` / :func:`load_safetensors`: Hugging Face layout (flat
  name -> tensor, contiguous, metadata ``{"format": "pt"}``), so a BERT checkpoint
  written here opens with ``transformers`` and vice versa.
* FSDP: :meth:`FullyShardedDataParallel.full_state_dict` gathers to rank 0 first.
"""
from __future__ import annotations

import os
import tempfile
from typing import Any, Dict, Optional

import torch
import torch.nn as nn

from ..parallel import process_group as pg

_PREFIX = "module."


def _optimizer_state(optimizer: Any):
    """FSDP wrappers are their own optimizer (and an nn.Module): their optimizer state has its own method."""
    if hasattr(optimizer, "optim_state_dict"):
        return optimizer.optim_state_dict()  # collective
    return optimizer.state_dict()


def _notify_weights_loaded(model: nn.Module, state: Dict[str, torch.Tensor]) -> None:
    """Weights were written into a module that a DDP wrapper + FusedAdamW may already own: refresh the fp32 masters."""
    seen = set()
    for m in (model, unwrap(model)):
        w = m if hasattr(m, "notify_params_changed") else None
        if w is None:
            ref = getattr(m, "_ddl_wrapper", None)
            w = ref() if ref is not None else None
        if w is not None and id(w) not in seen:
            seen.add(id(w))
            w.notify_params_changed(state)


def unwrap(model: nn.Module) -> nn.Module:
    """Return the user's module from under DDP/FSDP wrappers."""
    seen = 0
    while hasattr(model, "module") and isinstance(getattr(model, "module"), nn.Module) and seen < 4:
        model = model.module
        seen += 1
    return model


def strip_prefix(state: Dict[str, torch.Tensor], prefix: str = _PREFIX) -> Dict[str, torch.Tensor]:
    return {(k[len(prefix):] if k.startswith(prefix) else k): v for k, v in state.items()}


def model_state_dict(model: nn.Module) -> Dict[str, torch.Tensor]:
    """State dict with wrapper prefixes removed; FSDP wrappers gather the full tensors."""
    if hasattr(model, "full_state_dict"):
        return strip_prefix(model.full_state_dict())
    return strip_prefix(model.state_dict())


def _atomic_write(path: str, writer) -> None:
    d = os.path.dirname(os.path.abspath(path)) or "."
    os.makedirs(d, exist_ok=True)
    fd, tmp = tempfile.mkstemp(dir=d, prefix="tmp_ckpt_", suffix=".part")
    os.close(fd)
    try:
        writer(tmp)
        os.replace(tmp, path)
    finally:
        if os.path.exists(tmp):
            os.remove(tmp)


def save_checkpoint(
    path: str,
    model: nn.Module,
    optimizer: Optional[Any] = None,
    step: int = 0,
    extra: Optional[Dict[str, Any]] = None,
    rank0_only: bool = True,
) -> Optional[str]:
    """Write a torch ``state_dict`` checkpoint from rank 0; all ranks synchronise after."""
    state = model_state_dict(model)  # collective for FSDP: every rank must call
    opt_state = _optimizer_state(optimizer) if optimizer is not None else None  # collective for FSDP as well
    wrote = None
    if not rank0_only or pg.rank() == 0:
        blob = {
            "model": {k: v.detach().to("cpu") for k, v in state.items()},
            "optimizer": opt_state,
            "step": int(step),
            "extra": extra or {},
            "format": "b200-ddl/state_dict/v1",
        }
        _atomic_write(path, lambda tmp: torch.save(blob, tmp))
        wrote = path
    pg.barrier()
    return wrote


def load_checkpoint(
    path: str,
    model: nn.Module,
    optimizer: Optional[Any] = None,
    map_location: Optional[Any] = None,
    strict: bool = True,
) -> Dict[str, Any]:
    """Load a checkpoint on every rank.  Returns ``{"step": ..., "extra": ...}``.

    Accepts this framework's format, a bare ``state_dict`` file, and state dicts written
    from under stock DDP (``module.`` prefixed keys).
    """
    blob = torch.load(path, map_location=map_location or "cpu", weights_only=False)
    state = blob["model"] if isinstance(blob, dict) and "model" in blob and isinstance(blob["model"], dict) else blob
    state = strip_prefix(state)
    target = unwrap(model)
    if hasattr(model, "load_full_state_dict"):
        model.load_full_state_dict(state, strict=strict)
    else:
        target.load_state_dict(state, strict=strict)
    if optimizer is not None and isinstance(blob, dict) and blob.get("optimizer") is not None:
        if hasattr(optimizer, "load_optim_state_dict"):
            optimizer.load_optim_state_dict(blob["optimizer"])
        else:
            optimizer.load_state_dict(blob["optimizer"])
    else:
        # weights only: a live optimizer keeps its moments but must take the loaded values as its masters
        _notify_weights_loaded(model, state)
    meta = {"step": 0, "extra": {}}
    if isinstance(blob, dict):
        meta["step"] = int(blob.get("step", 0))
        meta["extra"] = blob.get("extra", {})
    return meta


def save_safetensors(path: str, model: nn.Module, metadata: Optional[Dict[str, str]] = None,
                     rank0_only: bool = True) -> Optional[str]:
    """Hugging Face style ``model.safetensors`` written by rank 0."""
    from safetensors.torch import save_file

    state = model_state_dict(model)
    wrote = None
    if not rank0_only or pg.rank() == 0:
        # safetensors refuses shared storage and non-contiguous tensors: clone views
        tensors = {k: v.detach().to("cpu").contiguous().clone() for k, v in state.items()}
        md = {"format": "pt"}
        md.update(metadata or {})
        _atomic_write(path, lambda tmp: save_file(tensors, tmp, metadata=md))
        wrote = path
    pg.barrier()
    return wrote


def load_safetensors(path: str, model: nn.Module, strict: bool = True, device: str = "cpu"):
    from safetensors.torch import load_file

    state = strip_prefix(load_file(path, device=device))
    if hasattr(model, "load_full_state_dict"):
        return model.load_full_state_dict(state, strict=strict)  # refreshes its own master shards
    res = unwrap(model).load_state_dict(state, strict=strict)
    _notify_weights_loaded(model, state)
    return res


def save_pretrained(directory: str, model: nn.Module, config: Optional[Dict[str, Any]] = None) -> None:
    """HF directory layout: ``model.safetensors`` (+ ``config.json`` when given)."""
    import json

    os.makedirs(directory, exist_ok=True)
    save_safetensors(os.path.join(directory, "model.safetensors"), model)
    if config is not None and pg.rank() == 0:
        with open(os.path.join(directory, "config.json"), "w") as f:
            json.dump(config, f, indent=2)
