"""Symmetric memory: every rank maps every peer's heap (target layer T1).

A :class:`SymmetricHeap` is a set of slabs.  Each rank ``cudaMalloc``s a slab of the same
size, exports a CUDA IPC handle, the 64-byte handles are exchanged over the bootstrap
process group (objects only - no NCCL data path), and every rank opens every peer's
handle.  Because all ranks perform the same allocation sequence, a buffer lives at the
same offset in every rank's slab: a kernel receives the 8 peer base pointers plus one
offset and can ``ld``/``st`` any peer's copy over NVLink 5 / NVSwitch
(``csrc/collectives.cu``).

Providers
---------
``auto``   (default) ``torch`` when world size > 1 (it brings the NVLS multicast mapping),
           falling back to ``ipc`` if the rendezvous is not available; ``ipc`` for one rank.
``ipc``    the native path above (``csrc/runtime.cu``).
``torch``  allocation + handle exchange through ``torch.distributed._symmetric_memory``;
           used when the NVLS multicast mapping is wanted (K3), because it also binds a
           multicast object where the host exposes one.  Only the *mapping* comes from
           torch; the kernels that touch the memory are ours.

The reference has nothing comparable.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass
from typing import List, Optional

import torch

from ..ops import _lib as L
from . import process_group as pg

_ALIGN = 1024


class _RawCudaMemory:
    """Wraps a raw device pointer for ``torch.as_tensor`` (CUDA array interface v3)."""

    def __init__(self, ptr: int, nbytes: int, owner=None):
        self.__cuda_array_interface__ = {
            "shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3, "strides": None,
        }
        self._owner = owner


@dataclass
class SymmBuffer:
    """One allocation in the symmetric heap."""

    tensor: torch.Tensor          # this rank's view (flat, dtype as requested)
    peer_ptrs: List[int]          # address of the same buffer in every rank's slab, mapped here
    multicast_ptr: Optional[int]  # NVLS multicast address of the buffer, if bound
    nbytes: int

    def ptr_array(self, byte_offset: int = 0):
        arr = (C.c_void_p * 8)()
        for i, p in enumerate(self.peer_ptrs):
            arr[i] = p + byte_offset
        return arr

    def mc(self, byte_offset: int = 0) -> Optional[int]:
        return None if self.multicast_ptr is None else self.multicast_ptr + byte_offset


class _Slab:
    def __init__(self, local_ptr: int, peer_bases: List[int], size: int, raw: torch.Tensor, mc_base: Optional[int], keep=None):
        self.local_ptr, self.peer_bases, self.size, self.raw, self.mc_base = local_ptr, peer_bases, size, raw, mc_base
        self.used = 0
        self.keep = keep


class SymmetricHeap:
    def __init__(self, device: torch.device, slab_bytes: Optional[int] = None, provider: Optional[str] = None):
        if device.type != "cuda":
            raise RuntimeError("symmetric memory needs a CUDA device")
        self.device = device
        self.rank, self.world = pg.rank(), pg.world_size()
        if self.world > 8:
            raise RuntimeError("one NVSwitch domain: world size must be <= 8")
        self.slab_bytes = int(slab_bytes or int(os.environ.get("DDL_SYMM_SLAB_MB", "512")) << 20)
        self.provider = provider or os.environ.get("DDL_SYMM_PROVIDER", "auto")
        if self.provider not in ("auto", "ipc", "torch"):
            raise ValueError(self.provider)
        if self.provider == "auto":
            # the multicast (NVLS) mapping only comes with the torch provider; a single rank has no use for it
            self.provider = "torch" if self.world > 1 else "ipc"
            self._auto = True
        else:
            self._auto = False
        self.slabs: List[_Slab] = []
        self.multicast_supported = False

    # ---- slab creation (collective: every rank must call in the same order) -------------
    def _new_slab(self, min_bytes: int) -> _Slab:
        size = max(self.slab_bytes, (min_bytes + _ALIGN - 1) // _ALIGN * _ALIGN)
        if self.provider == "torch":
            try:
                slab = self._new_slab_torch(size)
                failed = False
            except Exception:  # noqa: BLE001 - symmetric-memory rendezvous not available on this host
                if not (self._auto and not self.slabs):
                    raise
                failed = True
            # every rank must take the same path: agree before falling back
            if self._auto and not self.slabs and any(pg.all_gather_object(failed)):
                self.provider = "ipc"
                slab = self._new_slab_ipc(size)
        else:
            slab = self._new_slab_ipc(size)
        self.slabs.append(slab)
        return slab

    def _new_slab_ipc(self, size: int) -> _Slab:
        lib = L.lib()
        torch.cuda.set_device(self.device)
        p = C.c_void_p()
        L.check(lib.ddl_symm_alloc(size, C.byref(p)), "ddl_symm_alloc")
        local = int(p.value)
        handle = C.create_string_buffer(64)
        L.check(lib.ddl_ipc_export(local, handle), "ddl_ipc_export")
        handles = pg.all_gather_object(bytes(handle.raw))
        bases: List[int] = []
        for r, h in enumerate(handles):
            if r == self.rank:
                bases.append(local)
                continue
            q = C.c_void_p()
            L.check(lib.ddl_ipc_open(C.create_string_buffer(h, 64), C.byref(q)), f"ddl_ipc_open(rank {r})")
            bases.append(int(q.value))
        raw = torch.as_tensor(_RawCudaMemory(local, size), device=self.device)
        pg.barrier()
        return _Slab(local, bases, size, raw, None)

    def _new_slab_torch(self, size: int) -> _Slab:
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm

        t = symm.empty(size, dtype=torch.uint8, device=self.device)
        hdl = symm.rendezvous(t, dist.group.WORLD.group_name)
        t.zero_()
        torch.cuda.synchronize(self.device)
        bases = [int(p) for p in hdl.buffer_ptrs]
        mc = int(hdl.multicast_ptr) if getattr(hdl, "multicast_ptr", 0) else None
        self.multicast_supported = mc is not None
        pg.barrier()
        return _Slab(t.data_ptr(), bases, size, t, mc, keep=hdl)

    # ---- allocation ------------------------------------------------------------------
    def alloc(self, numel: int, dtype: torch.dtype) -> SymmBuffer:
        """Collective (all ranks, same order, same sizes).  Memory is zero-initialised."""
        esize = torch.empty((), dtype=dtype).element_size()
        nbytes = (numel * esize + _ALIGN - 1) // _ALIGN * _ALIGN
        slab = None
        for s in self.slabs:
            if s.size - s.used >= nbytes:
                slab = s
                break
        if slab is None:
            slab = self._new_slab(nbytes)
        off = slab.used
        slab.used += nbytes
        view = slab.raw[off : off + numel * esize].view(dtype)
        return SymmBuffer(
            tensor=view,
            peer_ptrs=[b + off for b in slab.peer_bases],
            multicast_ptr=None if slab.mc_base is None else slab.mc_base + off,
            nbytes=nbytes,
        )


_HEAPS = {}


def get_heap(device: torch.device) -> SymmetricHeap:
    key = (device.type, device.index)
    if key not in _HEAPS:
        _HEAPS[key] = SymmetricHeap(device)
    return _HEAPS[key]
