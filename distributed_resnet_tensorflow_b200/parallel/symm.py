"""Symmetric device memory: the same-sized buffer on every rank, mapped into every peer.

Two providers:
  * ``torch``  - ``torch.distributed._symmetric_memory`` (CUDA VMM allocations exchanged through the
    process-group store); also yields an NVLS *multicast* address when the fabric supports it, which
    the collectives use for ``multimem.ld_reduce`` / ``multimem.st``.
  * ``ipc``    - plain ``cudaMalloc`` + CUDA IPC handles exchanged with ``all_gather_object``
    (csrc/comm.cu: drn_ipc_alloc / drn_ipc_open). Peer loads/stores only, no multicast.

The reference gets the equivalent "shared variable space" from TF's gRPC parameter servers
(``tf.train.Server`` + ``replica_device_setter``, resnet_cifar_main.py:371-396); here the parameter
and gradient arenas of all replicas are simply addressable from every GPU over NVLink.
"""
from __future__ import annotations

import ctypes
import os
from typing import List, Optional

import torch
import torch.distributed as dist

from .._native import lib as _nl


class _RawCuda:
    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1",
                                         "data": (ptr, False), "version": 3}


class SymmetricBuffer:
    """``nbytes`` of zero-initialised device memory on each rank of ``group``."""

    def __init__(self, nbytes: int, device: torch.device, group=None, provider: Optional[str] = None):
        self.nbytes = int(nbytes)
        self.device = torch.device(device)
        self.group = group if group is not None else dist.group.WORLD
        self.rank = dist.get_rank(self.group)
        self.world_size = dist.get_world_size(self.group)
        self.multicast_ptr = 0
        self.provider = provider or os.environ.get("DRN_SYMM", "auto")
        self._keep = []
        err = None
        if self.provider in ("auto", "torch"):
            try:
                self._init_torch()
                self.provider = "torch"
            except Exception as e:  # pragma: no cover - depends on driver/fabric
                err = e
                if self.provider == "torch":
                    raise
                self.provider = "ipc"
        if self.provider == "ipc":
            self._init_ipc()
        self._err = err
        self._ptr_array = (ctypes.c_void_p * self.world_size)(*self.peer_ptrs)

    def _init_torch(self) -> None:
        import torch.distributed._symmetric_memory as symm_mem
        t = symm_mem.empty(self.nbytes, dtype=torch.uint8, device=self.device)
        t.zero_()
        hdl = symm_mem.rendezvous(t, self.group)
        self.local = t
        self.peer_ptrs: List[int] = [int(p) for p in hdl.buffer_ptrs]
        try:
            self.multicast_ptr = int(hdl.multicast_ptr) if hdl.has_multicast_support else 0
        except Exception:
            self.multicast_ptr = int(getattr(hdl, "multicast_ptr", 0) or 0)
        self._keep.append(hdl)
        assert self.peer_ptrs[self.rank] == t.data_ptr()

    def _init_ipc(self) -> None:
        ptr = ctypes.c_void_p()
        handle = (ctypes.c_ubyte * 64)()
        _nl.call("drn_ipc_alloc", ctypes.c_size_t(self.nbytes), ctypes.byref(ptr), handle)
        mine = bytes(handle)
        gathered: List[Optional[bytes]] = [None] * self.world_size
        dist.all_gather_object(gathered, mine, group=self.group)
        ptrs = []
        for r, h in enumerate(gathered):
            if r == self.rank:
                ptrs.append(int(ptr.value))
                continue
            out = ctypes.c_void_p()
            buf = (ctypes.c_ubyte * 64).from_buffer_copy(h)
            _nl.call("drn_ipc_open", buf, ctypes.byref(out))
            ptrs.append(int(out.value))
        self.peer_ptrs = ptrs
        self.local = torch.as_tensor(_RawCuda(int(ptr.value), self.nbytes), device=self.device)
        dist.barrier(group=self.group)

    # ---- views --------------------------------------------------------------------------------
    def view(self, dtype: torch.dtype, offset_bytes: int = 0, numel: Optional[int] = None):
        item = torch.empty((), dtype=dtype).element_size()
        end = self.nbytes if numel is None else offset_bytes + numel * item
        return self.local[offset_bytes:end].view(dtype)

    def ptr_array(self, offset_bytes: int = 0):
        """ctypes ``void*[world]`` of every rank's pointer (+offset) for the C collectives."""
        if offset_bytes == 0:
            return self._ptr_array
        return (ctypes.c_void_p * self.world_size)(*[p + offset_bytes for p in self.peer_ptrs])

    def mc(self, offset_bytes: int = 0) -> int:
        return self.multicast_ptr + offset_bytes if self.multicast_ptr else 0
