"""Peer-mapped device buffers for the GPUs of one box (csrc/p2p.cu).

`DevBuffer` is a cudaMalloc'ed region owned by this process; `exchange()` shares its CUDA IPC
handle with the other ranks of a torch.distributed group and maps theirs, so every rank ends up
with one raw device pointer per rank for each named buffer.  Kernels dereference those pointers
directly over NVLink/NVSwitch (frame pull in csrc/preproc.cu, detection push in csrc/detect.cu).

torch.distributed is only the plumbing here (handle exchange, barriers); no tensor data moves
through NCCL on the product path.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from ..ops import lib as L


class _RawView:
    """Minimal __cuda_array_interface__ holder so torch can wrap a raw device pointer."""

    def __init__(self, ptr: int, shape, typestr: str, owner=None):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 3}
        self._owner = owner


_TYPESTR = {torch.uint8: "|u1", torch.int32: "<i4", torch.float32: "<f4", torch.float16: "<f2", torch.uint32: "<u4"}


def tensor_from_ptr(ptr: int, shape, dtype: torch.dtype, device, owner=None) -> torch.Tensor:
    with torch.cuda.device(device):
        return torch.as_tensor(_RawView(ptr, shape, _TYPESTR[dtype], owner), device=device)


class DevBuffer:
    def __init__(self, nbytes: int, device):
        self.device = torch.device(device)
        self.nbytes = int((nbytes + 255) // 256 * 256)
        self.lib = L.load()
        p = C.c_void_p()
        with torch.cuda.device(self.device):
            L.check(self.lib.dvc_p2p_alloc(self.nbytes, C.byref(p)), "p2p_alloc")
        self.ptr = int(p.value)

    def handle(self) -> bytes:
        buf = C.create_string_buffer(64)
        with torch.cuda.device(self.device):
            L.check(self.lib.dvc_p2p_get_handle(self.ptr, C.cast(buf, C.c_void_p)), "p2p_get_handle")
        return bytes(buf.raw)

    def tensor(self, dtype: torch.dtype, shape, offset_bytes: int = 0) -> torch.Tensor:
        return tensor_from_ptr(self.ptr + offset_bytes, shape, dtype, self.device, owner=self)

    def free(self):
        if self.ptr:
            with torch.cuda.device(self.device):
                self.lib.dvc_p2p_free(self.ptr)
            self.ptr = 0


def open_handle(handle: bytes, device) -> int:
    lib = L.load()
    p = C.c_void_p()
    buf = C.create_string_buffer(handle, 64)
    with torch.cuda.device(device):
        L.check(lib.dvc_p2p_open(C.cast(buf, C.c_void_p), C.byref(p)), "p2p_open")
    return int(p.value)


def exchange(named: Dict[str, Optional[DevBuffer]], device, group=None) -> Dict[str, List[Optional[int]]]:
    """All-gather IPC handles. Returns name -> [ptr usable on this rank, for each rank] (None where a
    rank did not provide that buffer)."""
    import torch.distributed as dist
    rank = dist.get_rank(group)
    world = dist.get_world_size(group)
    mine = {k: (v.handle() if v is not None else None) for k, v in named.items()}
    gathered: List[Optional[dict]] = [None] * world
    dist.all_gather_object(gathered, mine, group=group)
    out: Dict[str, List[Optional[int]]] = {}
    for name in named:
        ptrs: List[Optional[int]] = []
        for r in range(world):
            h = gathered[r].get(name) if gathered[r] else None
            if h is None:
                ptrs.append(None)
            elif r == rank:
                ptrs.append(named[name].ptr)
            else:
                ptrs.append(open_handle(h, device))
        out[name] = ptrs
    return out


def device_ptr_array(ptrs: Sequence[int], device) -> torch.Tensor:
    return torch.as_tensor(np.asarray(ptrs, dtype=np.int64), device=device)
