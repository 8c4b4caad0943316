"""Symmetric heap over NVLink peer memory (host side of csrc/comm.cuh).

Each rank owns one heap (flags in the first 64 KiB, payload after) that every peer has mapped; kernels push
data straight into peer memory and NCCL is not on any hot path.  Regions are handed out by a bump allocator
that every rank drives identically, so an offset means the same thing everywhere ("symmetric").

Two ways of getting the mappings (``DAPB_HEAP_BACKEND=auto|vmm|ipc``):

* ``vmm`` -- the heap is a ``cuMemCreate`` allocation made and exchanged through
  ``torch.distributed._symmetric_memory`` (fd passing, peer mappings).  When the fabric supports it the
  allocation is also bound to an NVLS **multicast** object and ``comm.mc_heap`` carries that mapping: one
  ``multimem.st`` then reaches all ranks and ``multimem.ld_reduce`` sums in the switch (csrc/comm.cuh).
* ``ipc`` -- plain ``cudaMalloc`` + ``cudaIpc*`` handles exchanged through ``all_gather_object``; unicast only
  (every primitive keeps a pure P2P path).

``auto`` tries ``vmm`` and falls back to ``ipc`` (the decision is agreed on by all ranks).
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass
from typing import Dict, Optional

import torch

from .. import _native
from .dist import Dist, current


@dataclass
class Region:
    """A symmetric allocation: payload offset + a private range of flag words + an epoch counter."""
    payload_off: int
    nbytes: int
    flag_base: int
    n_flags: int
    epoch: int = 0

    def next_epoch(self) -> int:
        self.epoch += 1
        return self.epoch


class SymmetricHeap:
    def __init__(self, nbytes: int = 64 << 20, dist: Optional[Dist] = None):
        self.dist = dist or current()
        self.lib = _native.load()
        self.nbytes = int(nbytes)
        self._payload_top = 0
        self._flag_top = 0
        self._regions: Dict[str, Region] = {}
        self._peer_ptrs = [None] * self.dist.world
        self.mc_ptr = 0
        self._vmm = None                    # (tensor, handle) keep-alive of the torch symmetric-memory allocation
        want = os.environ.get("DAPB_HEAP_BACKEND", "auto")
        if self.dist.world > 1 and want in ("auto", "vmm"):
            self._try_vmm(required=(want == "vmm"))
        self.backend = "vmm" if self._vmm is not None else "ipc"
        if self._vmm is None:
            self._init_ipc()
        self.comm = _native.DapbComm()
        self.comm.rank = self.dist.rank
        self.comm.world = self.dist.world
        self.comm.heap_bytes = self.nbytes
        for r in range(self.dist.world):
            self.comm.heap[r] = self._peer_ptrs[r]
        use_mc = os.environ.get("DAPB_MULTICAST", "1") != "0"
        self.comm.mc_heap = self.mc_ptr if (self.mc_ptr and use_mc) else None

    @property
    def multicast(self) -> bool:
        return bool(self.comm.mc_heap)

    def _try_vmm(self, required: bool):
        """cuMemCreate-backed heap with peer + multicast mappings via torch's symmetric-memory rendezvous."""
        import torch.distributed as tdist
        err = None
        try:
            import torch.distributed._symmetric_memory as symm_mem
            t = symm_mem.empty(self.nbytes, dtype=torch.uint8, device=self.dist.device)
            hdl = symm_mem.rendezvous(t, tdist.group.WORLD)
            t.zero_()
            torch.cuda.synchronize()
            ptrs = [int(p) for p in hdl.buffer_ptrs]
            if len(ptrs) != self.dist.world or ptrs[self.dist.rank] != t.data_ptr():
                raise RuntimeError("unexpected symmetric-memory pointer table")
            mc = int(hdl.multicast_ptr or 0)
        except Exception as e:                  # noqa: BLE001 -- any failure means "use the IPC heap"
            err = repr(e)[:200]
        oks = self.dist.all_gather_object(err is None)
        if not all(oks):
            if required:
                raise RuntimeError(f"DAPB_HEAP_BACKEND=vmm but the symmetric-memory rendezvous failed: {err}")
            return
        self._vmm = (t, hdl)
        self.base = ptrs[self.dist.rank]
        self._peer_ptrs = ptrs
        mcs = self.dist.all_gather_object(mc != 0)
        self.mc_ptr = mc if all(mcs) else 0
        self.dist.barrier()

    def _init_ipc(self):
        base = C.c_void_p()
        _native.check(self.lib.dapb_symm_alloc(self.nbytes, C.byref(base)), "symm_alloc")
        self.base = base.value
        self._peer_ptrs[self.dist.rank] = self.base
        if self.dist.world > 1:
            handle = (C.c_ubyte * 64)()
            _native.check(self.lib.dapb_ipc_get_handle(self.base, handle), "ipc_get_handle")
            handles = self.dist.all_gather_object(bytes(handle))
            for r, h in enumerate(handles):
                if r == self.dist.rank:
                    continue
                buf = (C.c_ubyte * 64).from_buffer_copy(h)
                out = C.c_void_p()
                _native.check(self.lib.dapb_ipc_open_handle(buf, C.byref(out)), f"ipc_open_handle(rank {r})")
                self._peer_ptrs[r] = out.value
            self.dist.barrier()

    # ------------------------------------------------------------------ allocation
    def region(self, name: str, nbytes: int, n_flags: int) -> Region:
        """Named symmetric region (idempotent).  Every rank must request regions in the same order."""
        if name in self._regions:
            reg = self._regions[name]
            if reg.nbytes < nbytes or reg.n_flags < n_flags:
                raise ValueError(f"region {name!r} exists with a smaller size")
            return reg
        off = (self._payload_top + 255) & ~255
        if _native.FLAG_BYTES + off + nbytes > self.nbytes:
            raise MemoryError(f"symmetric heap exhausted: need {nbytes} B at {off}, heap {self.nbytes} B")
        if (self._flag_top + n_flags) * 8 > _native.FLAG_BYTES:
            raise MemoryError("symmetric heap flag words exhausted")
        reg = Region(off, nbytes, self._flag_top, n_flags)
        self._payload_top = off + nbytes
        self._flag_top += n_flags
        self._regions[name] = reg
        return reg

    def payload_ptr(self, reg: Region, rank: Optional[int] = None) -> int:
        r = self.dist.rank if rank is None else rank
        return self._peer_ptrs[r] + _native.FLAG_BYTES + reg.payload_off

    def tensor(self, reg: Region, shape, dtype) -> torch.Tensor:
        """Local view of a region as a torch tensor (zero-copy, device memory owned by the heap)."""
        n = 1
        for s in shape:
            n *= int(s)
        itemsize = torch.empty((), dtype=dtype).element_size()
        assert n * itemsize <= reg.nbytes
        ptr = self.payload_ptr(reg)
        arr = _CudaArray(ptr, tuple(int(s) for s in shape), dtype)
        return torch.as_tensor(arr, device=self.dist.device)

    def comm_ref(self):
        return C.byref(self.comm)

    def close(self):
        if self.base is None:
            return
        torch.cuda.synchronize()
        if self._vmm is not None:
            self.dist.barrier()
            self._vmm = None                # torch unmaps and releases the allocation with the last reference
            self.base = None
            return
        if self.dist.world > 1:
            self.dist.barrier()
            for r, p in enumerate(self._peer_ptrs):
                if r != self.dist.rank and p:
                    self.lib.dapb_ipc_close_handle(p)
            self.dist.barrier()
        self.lib.dapb_symm_free(self.base)
        self.base = None


_TYPESTR = {torch.float32: "<f4", torch.float64: "<f8", torch.int32: "<i4", torch.int64: "<i8",
            torch.uint8: "|u1", torch.float16: "<f2", torch.int16: "<i2"}


class _CudaArray:
    """Minimal __cuda_array_interface__ carrier so torch can wrap raw heap memory."""

    def __init__(self, ptr: int, shape, dtype):
        if dtype == torch.bfloat16:
            raise TypeError("wrap bf16 regions as int16 and .view(torch.bfloat16)")
        self.__cuda_array_interface__ = {"shape": shape, "typestr": _TYPESTR[dtype], "data": (ptr, False),
                                         "version": 2, "strides": None}


_DEFAULT: Optional[SymmetricHeap] = None


def default_heap(nbytes: int = 1 << 30) -> SymmetricHeap:
    """Process-wide heap (1 GiB of the 180 GB by default), created on first use (collective across ranks!).
    ``DAPB_HEAP_MB`` overrides the size."""
    global _DEFAULT
    if _DEFAULT is None:
        import os
        mb = os.environ.get("DAPB_HEAP_MB")
        _DEFAULT = SymmetricHeap(int(mb) << 20 if mb else nbytes)
    return _DEFAULT


def reset_default_heap():
    global _DEFAULT
    if _DEFAULT is not None:
        _DEFAULT.close()
    _DEFAULT = None
