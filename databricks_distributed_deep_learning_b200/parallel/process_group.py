"""Process-group bootstrap (target layer T1 in SURVEY.md §1.2).

``torch.distributed`` is the *plumbing* only: rendezvous, barriers, small object
exchange (IPC handles, shapes) and the CPU/gloo configuration of BASELINE.json:7.
On GPU the gradient traffic itself never goes through it: the hand-written peer-memory
kernels in ``csrc/collectives.cu`` do (see :mod:`.symm` and :mod:`.collectives`).
The reference has no equivalent (zero collective call sites).
"""
from __future__ import annotations

import datetime
import os
from typing import Any, List, Optional

import torch
import torch.distributed as dist

_STATE = {"device": None, "owns_pg": False}


def env_rank() -> int:
    return int(os.environ.get("RANK", "0"))


def env_world_size() -> int:
    return int(os.environ.get("WORLD_SIZE", "1"))


def env_local_rank() -> int:
    return int(os.environ.get("LOCAL_RANK", os.environ.get("RANK", "0")))


def is_initialized() -> bool:
    return dist.is_available() and dist.is_initialized()


def init_process_group(
    backend: Optional[str] = None,
    use_gpu: Optional[bool] = None,
    timeout_s: float = 1800.0,
) -> torch.device:
    """Initialise the default group from the env contract and pick this rank's device.

    Returns the device this rank owns (``cuda:LOCAL_RANK`` or ``cpu``).  Safe to call
    when a group already exists (e.g. under ``torchrun`` after user init).
    """
    if use_gpu is None:
        use_gpu = torch.cuda.is_available()
    if use_gpu:
        local = env_local_rank()
        torch.cuda.set_device(local)
        device = torch.device("cuda", local)
    else:
        device = torch.device("cpu")
    _STATE["device"] = device
    if is_initialized():
        return device
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29511")
    if backend is None:
        # "cpu:gloo,cuda:nccl": objects/handles over gloo, device barriers over nccl
        backend = "cpu:gloo,cuda:nccl" if use_gpu else "gloo"
    kwargs = {}
    if use_gpu:
        kwargs["device_id"] = device
    dist.init_process_group(
        backend=backend,
        rank=env_rank(),
        world_size=env_world_size(),
        timeout=datetime.timedelta(seconds=timeout_s),
        **kwargs,
    )
    _STATE["owns_pg"] = True
    return device


def destroy_process_group() -> None:
    if is_initialized() and _STATE["owns_pg"]:
        try:
            dist.destroy_process_group()
        finally:
            _STATE["owns_pg"] = False


def rank() -> int:
    return dist.get_rank() if is_initialized() else env_rank()


def world_size() -> int:
    return dist.get_world_size() if is_initialized() else env_world_size()


def local_rank() -> int:
    return env_local_rank()


def device() -> torch.device:
    d = _STATE["device"]
    if d is None:
        d = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
    return d


def barrier() -> None:
    if is_initialized() and world_size() > 1:
        if device().type == "cuda":
            dist.barrier(device_ids=[device().index])
        else:
            dist.barrier()


def all_gather_object(obj: Any) -> List[Any]:
    """Gather one picklable object per rank (used for IPC handles, shapes, stats)."""
    if not is_initialized() or world_size() == 1:
        return [obj]
    out: List[Any] = [None] * world_size()
    dist.all_gather_object(out, obj)
    return out


def broadcast_object(obj: Any, src: int = 0) -> Any:
    if not is_initialized() or world_size() == 1:
        return obj
    box = [obj]
    dist.broadcast_object_list(box, src=src)
    return box[0]


def max_over_ranks(value: float) -> float:
    """Max of a host scalar over all ranks (timing rule: report the slowest rank)."""
    if not is_initialized() or world_size() == 1:
        return float(value)
    return float(max(all_gather_object(float(value))))
