"""Python face of the peer-memory collectives (T2, K1-K5/K7 in SURVEY.md §2.E.3).

:class:`PeerMemoryComm` implements the :class:`..comm.CommBackend` interface on top of
``csrc/collectives.cu``: buffers come from the symmetric heap, every collective is one
kernel launch per rank on the *current CUDA stream* (so it is stream-ordered, capturable
in a CUDA graph, and ``CommHandle.wait`` is a no-op), and the all-reduce variant is chosen
per message size by :class:`VariantSelector` from a measured bus-bandwidth table (K4).
"""
from __future__ import annotations

import ctypes as C
import json
import os
from typing import Dict, List, Optional, Tuple

import torch

from ..ops import _lib as L
from . import process_group as pg
from .comm import CommBackend, CommHandle
from .symm import SymmBuffer, SymmetricHeap, get_heap

_ALGO_NAMES = {"auto": L.ALGO_AUTO, "oneshot": L.ALGO_ONESHOT, "twoshot": L.ALGO_TWOSHOT, "nvls": L.ALGO_NVLS}


class _StreamOrdered(CommHandle):
    def wait(self) -> None:  # the kernel is already ordered on the launching stream
        return None


class VariantSelector:
    """K4: message bytes -> (algorithm, num_blocks), from a measured table.

    The table is a list of ``{"max_bytes": B, "algo": name, "blocks": n}`` rows in
    ascending ``max_bytes``; the default below is replaced by the one
    ``benchmarks/allreduce_sweep.py`` writes to ``profiles/allreduce_table_w{world}.json``.
    """

    # shape of the measured 2- and 8-GPU tables (profiles/allreduce_table_w{2,8}.json): one-shot while
    # latency dominates, then the switch reduction where the host exposes it, else two-shot
    DEFAULT = [
        {"max_bytes": 16 << 10, "algo": "oneshot", "blocks": 4},
        {"max_bytes": 1 << 20, "algo": "nvls", "blocks": 16},
        {"max_bytes": 8 << 20, "algo": "nvls", "blocks": 32},
        {"max_bytes": 1 << 62, "algo": "nvls", "blocks": 8},
    ]

    def __init__(self, world: int, nvls: bool, table: Optional[List[dict]] = None):
        self.world, self.nvls = world, nvls
        self.table = table or self._load(world) or self.DEFAULT

    @staticmethod
    def _load(world: int) -> Optional[List[dict]]:
        root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
        p = os.environ.get("DDL_ALLREDUCE_TABLE", os.path.join(root, "profiles", f"allreduce_table_w{world}.json"))
        if os.path.exists(p):
            try:
                with open(p) as f:
                    return json.load(f)["table"]
            except Exception:
                return None
        return None

    def pick_row(self, nbytes: int) -> dict:
        for row in self.table:
            if nbytes <= row["max_bytes"]:
                return row
        return {"algo": "twoshot", "blocks": 64}

    def pick(self, nbytes: int, same_dtype: bool) -> Tuple[int, int]:
        for row in self.table:
            if nbytes <= row["max_bytes"]:
                algo = row["algo"]
                blocks = int(row["blocks"])
                if algo == "hybrid":  # callers that can run it ask pick_row(); everybody else gets the switch reduction
                    algo = "nvls"
                if algo == "nvls" and not (self.nvls and same_dtype):
                    # the switch is not doing the reduction: the SMs have to move (W-1)/W of the bytes
                    algo, blocks = "twoshot", max(blocks, 64 if nbytes >= (4 << 20) else 32)
                return _ALGO_NAMES[algo], blocks
        return L.ALGO_TWOSHOT, 64


class PeerMemoryComm(CommBackend):
    """Hand-written NVLink collectives; one instance == one signal-pad channel.

    Use distinct instances (channels) for collectives that may run concurrently on
    different streams (e.g. DDP gradient buckets vs. FSDP parameter gathers).
    """

    name = "p2p"

    def __init__(self, device: torch.device, heap: Optional[SymmetricHeap] = None, force_algo: Optional[str] = None,
                 num_blocks: Optional[int] = None):
        if not pg.is_initialized():
            raise RuntimeError("PeerMemoryComm needs an initialised process group for the handle exchange")
        self.device = device
        self.heap = heap or get_heap(device)
        self.world_size, self.rank = pg.world_size(), pg.rank()
        lib = L.lib()
        pad_bytes = lib.ddl_comm_signal_pad_bytes()
        self._pad = self.heap.alloc(pad_bytes, torch.uint8)
        self._epoch = torch.zeros(lib.ddl_comm_epoch_bytes(), dtype=torch.uint8, device=device)
        self._ct = L.CommT()
        for r in range(self.world_size):
            self._ct.signal_pads[r] = self._pad.peer_ptrs[r]
        self._ct.epoch = self._epoch.data_ptr()
        self._ct.rank, self._ct.world = self.rank, self.world_size
        self.selector = VariantSelector(self.world_size, nvls=self.heap.multicast_supported)
        self.force_algo = force_algo or os.environ.get("DDL_ALLREDUCE_ALGO")
        self.num_blocks = num_blocks
        self._registry: Dict[int, SymmBuffer] = {}
        torch.cuda.synchronize(device)
        pg.barrier()

    # ---- buffers ---------------------------------------------------------------------
    def alloc_symm(self, numel: int, dtype: torch.dtype) -> SymmBuffer:
        buf = self.heap.alloc(numel, dtype)
        self._registry[buf.tensor.data_ptr()] = buf
        return buf

    def alloc_flat(self, numel, dtype, device):
        return self.alloc_symm(numel, dtype).tensor

    def lookup(self, t: torch.Tensor) -> Tuple[SymmBuffer, int]:
        """Find the symmetric allocation a tensor (or a view into one) lives in."""
        p = t.data_ptr()
        hit = self._registry.get(p)
        if hit is not None:
            return hit, 0
        for base, buf in self._registry.items():
            if base <= p < base + buf.nbytes:
                return buf, p - base
        raise RuntimeError("tensor is not in symmetric memory; allocate it with comm.alloc_flat / alloc_symm")

    # ---- collectives -----------------------------------------------------------------
    def all_reduce_mean_(self, buf: torch.Tensor, out: Optional[torch.Tensor] = None, algo: Optional[str] = None,
                         num_blocks: Optional[int] = None, scale: Optional[float] = None) -> CommHandle:
        if self.world_size == 1 and scale is None:
            if out is not None and out.data_ptr() != buf.data_ptr():
                out.copy_(buf)  # mean over one rank: at most a cast
            return _StreamOrdered()
        sb, off = self.lookup(buf)
        if out is None or out.data_ptr() == buf.data_ptr():
            ob, ooff, out_dtype = sb, off, buf.dtype
        else:
            ob, ooff = self.lookup(out)
            out_dtype = out.dtype
        n = buf.numel()
        nbytes = n * buf.element_size()
        same = out_dtype == buf.dtype
        name = algo or self.force_algo
        if name is None and same and buf.dtype == torch.bfloat16 and ob is sb and ooff == off and scale is None:
            row = self.selector.pick_row(nbytes)
            if row.get("algo") == "hybrid" and self.heap.multicast_supported and self.world_size in (2, 4, 8):
                return self.all_reduce_hybrid_(buf, int(row.get("permille", 600)), int(row.get("blocks", 8)), int(row.get("p2p_blocks", 48)))
        if name == "hybrid":
            return self.all_reduce_hybrid_(buf, scale=scale)
        if name:
            a, nb = _ALGO_NAMES[name], num_blocks or self.num_blocks or 32
        else:
            a, nb = self.selector.pick(nbytes, same)
            nb = num_blocks or self.num_blocks or nb
        if a == L.ALGO_ONESHOT and nbytes > L.lib().ddl_allreduce_oneshot_max_bytes(L.dtype_code(buf.dtype), nb):
            a = L.ALGO_TWOSHOT
        L.call(
            "ddl_allreduce", C.byref(self._ct), sb.ptr_array(off), ob.ptr_array(ooff), n, L.dtype_code(buf.dtype),
            L.dtype_code(out_dtype), float(scale if scale is not None else 1.0 / self.world_size), a, sb.mc(off), ob.mc(ooff), nb,
            L.stream_ptr(self.device),
        )
        return _StreamOrdered()

    def all_reduce_hybrid_(self, buf: torch.Tensor, permille: int = 600, nvls_blocks: int = 8, p2p_blocks: int = 48,
                           scale: Optional[float] = None) -> CommHandle:
        """In-place bf16 mean (or ``scale`` x sum): ``permille`` / 1000 of the message through the switch (NVLS), the rest
        through two-shot peer loads / stores, in one kernel (``ddl_allreduce_hybrid``)."""
        if buf.dtype != torch.bfloat16 or not self.heap.multicast_supported:
            raise RuntimeError("hybrid all-reduce needs a bf16 buffer in multicast-mapped symmetric memory")
        sb, off = self.lookup(buf)
        L.call("ddl_allreduce_hybrid", C.byref(self._ct), sb.ptr_array(off), buf.numel(),
               float(scale if scale is not None else 1.0 / self.world_size), sb.mc(off), int(permille), int(nvls_blocks), int(p2p_blocks),
               L.stream_ptr(self.device))
        return _StreamOrdered()

    def broadcast_(self, buf: torch.Tensor, src: int = 0) -> None:
        try:
            sb, off = self.lookup(buf)
            staged = None
        except RuntimeError:
            # not symmetric: stage through a scratch allocation (init-time only)
            nbytes = (buf.numel() * buf.element_size() + 15) // 16 * 16
            staged = self._staging(nbytes)
            staged.tensor[: buf.numel() * buf.element_size()].copy_(buf.reshape(-1).view(torch.uint8))
            sb, off = staged, 0
        nbytes = (buf.numel() * buf.element_size() + 15) // 16 * 16
        L.call("ddl_broadcast", C.byref(self._ct), sb.ptr_array(off), int(src), nbytes, self.num_blocks or 32, L.stream_ptr(self.device))
        if staged is not None:
            buf.reshape(-1).view(torch.uint8).copy_(staged.tensor[: buf.numel() * buf.element_size()])

    def _staging(self, nbytes: int) -> SymmBuffer:
        cur = getattr(self, "_stage", None)
        if cur is None or cur.nbytes < nbytes:
            self._stage = self.alloc_symm(max(nbytes, 64 << 20), torch.uint8)
        return self._stage

    def all_gather_into(self, out: torch.Tensor, shard: torch.Tensor, num_blocks: Optional[int] = None) -> CommHandle:
        sb, off = self.lookup(shard)
        L.call("ddl_allgather", C.byref(self._ct), sb.ptr_array(off), out.data_ptr(), shard.numel() * shard.element_size(),
               num_blocks or self.num_blocks or 32, L.stream_ptr(self.device))
        return _StreamOrdered()

    def reduce_scatter_mean_into(self, out_shard: torch.Tensor, full: torch.Tensor, num_blocks: Optional[int] = None) -> CommHandle:
        sb, off = self.lookup(full)
        L.call("ddl_reducescatter", C.byref(self._ct), sb.ptr_array(off), out_shard.data_ptr(), out_shard.numel(),
               L.dtype_code(full.dtype), L.dtype_code(out_shard.dtype), 1.0 / self.world_size, num_blocks or self.num_blocks or 32,
               L.stream_ptr(self.device))
        return _StreamOrdered()

    def reduce_scatter_adamw_(self, full_grad: torch.Tensor, master: torch.Tensor, m: torch.Tensor, v: torch.Tensor,
                              param_shard_out: Optional[torch.Tensor], hyper: torch.Tensor, num_blocks: Optional[int] = None) -> None:
        """K7: gradient reduce-scatter fused with the sharded AdamW update."""
        sb, off = self.lookup(full_grad)
        L.call("ddl_reducescatter_adamw", C.byref(self._ct), sb.ptr_array(off), master.numel(), L.dtype_code(full_grad.dtype),
               1.0 / self.world_size, master.data_ptr(), m.data_ptr(), v.data_ptr(), L.ptr(param_shard_out), hyper.data_ptr(),
               num_blocks or self.num_blocks or 32, L.stream_ptr(self.device))

    def barrier(self) -> None:
        L.call("ddl_barrier", C.byref(self._ct), L.stream_ptr(self.device))

    def device_barrier(self) -> None:
        self.barrier()
