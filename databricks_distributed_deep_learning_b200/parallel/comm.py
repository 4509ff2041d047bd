"""Communication back ends used by the DDP / FSDP wrappers.

Two implementations of one small interface:

* :class:`TorchCollectiveComm` - ``torch.distributed`` (gloo on CPU for the
  BASELINE.json:7 plumbing config; NCCL only as the *measured baseline*).
* :class:`PeerMemoryComm` (in :mod:`.collectives`) - the product: hand-written sm_100a
  kernels that load/store peer GPU memory over NVLink 5 / NVSwitch
  (``csrc/collectives.cu``), no NCCL call on the gradient path (BASELINE.json:5).

The reference has no communication layer.
"""
from __future__ import annotations

from typing import Any, Optional

import torch
import torch.distributed as dist

from . import process_group as pg


class CommHandle:
    """Completion token of one asynchronous collective."""

    def wait(self) -> None:  # pragma: no cover - interface
        raise NotImplementedError


class _TorchHandle(CommHandle):
    def __init__(self, work, post=None):
        self._work = work
        self._post = post

    def wait(self) -> None:
        if self._work is not None:
            self._work.wait()
            self._work = None
        if self._post is not None:
            self._post()
            self._post = None


class CommBackend:
    """Interface the reducers program against."""

    name = "abstract"
    #: buffers returned by alloc_flat are usable as collective operands directly
    world_size: int = 1
    rank: int = 0

    def alloc_flat(self, numel: int, dtype: torch.dtype, device: torch.device) -> torch.Tensor:
        return torch.zeros(numel, dtype=dtype, device=device)

    def all_reduce_mean_(self, buf: torch.Tensor, out: Optional[torch.Tensor] = None) -> CommHandle:
        """Average ``buf`` over ranks.  Result lands in ``out`` (may alias ``buf``; may be a
        narrower dtype - the cast is fused on the peer-memory path)."""
        raise NotImplementedError

    def broadcast_(self, buf: torch.Tensor, src: int = 0) -> None:
        raise NotImplementedError

    def all_gather_into(self, out: torch.Tensor, shard: torch.Tensor) -> CommHandle:
        raise NotImplementedError

    def reduce_scatter_mean_into(self, out_shard: torch.Tensor, full: torch.Tensor) -> CommHandle:
        raise NotImplementedError

    def barrier(self) -> None:
        pg.barrier()


class TorchCollectiveComm(CommBackend):
    """gloo / NCCL through ``torch.distributed`` (baseline + CPU plumbing)."""

    name = "torch"

    def __init__(self, group: Any = None):
        self.group = group
        self.world_size = dist.get_world_size(group) if pg.is_initialized() else 1
        self.rank = dist.get_rank(group) if pg.is_initialized() else 0

    def all_reduce_mean_(self, buf, out=None):
        if self.world_size == 1:
            if out is not None and out.data_ptr() != buf.data_ptr():
                out.copy_(buf)
            return _TorchHandle(None)
        work = dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
        ws = self.world_size

        def post():
            buf.div_(ws)
            if out is not None and out.data_ptr() != buf.data_ptr():
                out.copy_(buf)

        return _TorchHandle(work, post)

    def broadcast_(self, buf, src=0):
        if self.world_size > 1:
            dist.broadcast(buf, src=src, group=self.group)

    def all_gather_into(self, out, shard):
        if self.world_size == 1:
            out.copy_(shard.reshape(-1)[: out.numel()])
            return _TorchHandle(None)
        work = dist.all_gather_into_tensor(out, shard.contiguous(), group=self.group, async_op=True)
        return _TorchHandle(work)

    def reduce_scatter_mean_into(self, out_shard, full):
        if self.world_size == 1:
            out_shard.copy_(full.reshape(-1)[: out_shard.numel()])
            return _TorchHandle(None)
        ws = self.world_size
        if full.device.type == "cpu":
            # gloo has no reduce_scatter: all_reduce then slice (plumbing path only)
            work = dist.all_reduce(full, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
            n = out_shard.numel()
            r = self.rank

            def post():
                out_shard.copy_(full.reshape(-1)[r * n : (r + 1) * n]).div_(ws)

            return _TorchHandle(work, post)
        work = dist.reduce_scatter_tensor(out_shard, full, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
        return _TorchHandle(work, lambda: out_shard.div_(ws))


def default_comm(device: torch.device, prefer: str = "auto") -> CommBackend:
    """Pick the back end: peer-memory kernels on CUDA, torch collectives on CPU.

    ``prefer`` = "auto" | "p2p" | "torch".  "p2p" on a CUDA device fails loudly if the
    native library or peer access is missing - no silent NCCL fallback on the product
    path.
    """
    if prefer not in ("auto", "p2p", "torch"):
        raise ValueError(prefer)
    if device.type == "cuda" and prefer in ("auto", "p2p"):
        from .collectives import PeerMemoryComm

        return PeerMemoryComm(device)
    if prefer == "p2p":
        raise RuntimeError("peer-memory collectives need a CUDA device")
    return TorchCollectiveComm()
