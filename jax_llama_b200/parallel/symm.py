"""Symmetric heap of one tensor-parallel group: the same buffer on every GPU, mapped three ways.

* ``local_ptr``  — this rank's copy (ordinary device memory, also visible as the torch tensor ``heap.t``);
* ``peer_ptrs``  — every rank's copy mapped into this process (plain ``ld``/``st`` over NVLink);
* ``mc_ptr``     — the NVLS multicast address of the buffer: ``multimem.st`` writes all copies with one store,
  ``multimem.ld_reduce`` returns the sum over all copies reduced inside the NVSwitch (csrc/nvls.cuh).

The mapping (``cuMemCreate`` + ``cuMulticastCreate/BindMem`` + handle exchange) is done by
``torch.distributed._symmetric_memory``; every instruction that touches it is ours.  When the platform has no
multicast support ``mc_ptr`` is 0 and the kernels fall back to one unicast store per peer.  What this replaces:
the placement XLA/GSPMD derives from /root/reference/jax_llama/partition.py:62-78 plus the NCCL buffers it
allocates behind the scenes.
"""
from __future__ import annotations

import ctypes as C
import logging
import os
from typing import List, Optional, Tuple

import torch
import torch.distributed as dist

logger = logging.getLogger(__name__)

MAX_TP = 8


class SymmetricHeap:
    """Bump allocator over one symmetric allocation (sub-buffers keep the same offset on every rank)."""

    def __init__(self, tp: int, rank: int, device: torch.device, nbytes: int,
                 group: Optional[dist.ProcessGroup] = None):
        self.tp, self.rank, self.device = tp, rank, device
        self.nbytes = (int(nbytes) + 4095) // 4096 * 4096
        self.top = 0
        self.mc_ptr = 0
        self.backend = "local"
        if tp == 1:
            self.t = torch.zeros(self.nbytes, dtype=torch.uint8, device=device)
            self.peer_ptrs: List[int] = [self.t.data_ptr()]
        else:
            import torch.distributed._symmetric_memory as symm_mem

            pg = group if group is not None else dist.group.WORLD
            with torch.cuda.device(device):
                self.t = symm_mem.empty(self.nbytes, dtype=torch.uint8, device=device)
                self.hdl = symm_mem.rendezvous(self.t, pg)
            self.t.zero_()
            self.peer_ptrs = [int(p) for p in self.hdl.buffer_ptrs]
            if os.environ.get("JLLAMA_NO_NVLS") != "1":
                try:
                    self.mc_ptr = int(self.hdl.multicast_ptr or 0)
                except Exception:  # noqa: BLE001  (older builds raise when the platform has no multicast)
                    self.mc_ptr = 0
            self.backend = "symm_mem+nvls" if self.mc_ptr else "symm_mem"
            torch.cuda.synchronize(device)
            dist.barrier(group=pg)                 # every copy is zeroed before anybody stores into a peer
        self.local_ptr = self.t.data_ptr()
        if self.peer_ptrs[rank if tp > 1 else 0] != self.local_ptr:
            raise RuntimeError("symmetric heap: own peer pointer does not match the local tensor")
        logger.info("symmetric heap: %d MiB, backend %s", self.nbytes >> 20, self.backend)

    @property
    def has_multicast(self) -> bool:
        return self.mc_ptr != 0

    def alloc(self, nbytes: int, align: int = 1024) -> int:
        """Offset of a fresh sub-buffer.  Every rank must perform the same sequence of allocations."""
        off = (self.top + align - 1) // align * align
        if off + nbytes > self.nbytes:
            raise RuntimeError(f"symmetric heap exhausted: need {nbytes} B at {off}, capacity {self.nbytes}")
        self.top = off + nbytes
        return off

    def view(self, off: int, shape: Tuple[int, ...], dtype: torch.dtype) -> torch.Tensor:
        n = 1
        for s in shape:
            n *= int(s)
        return self.t[off:off + n * torch.empty((), dtype=dtype).element_size()].view(dtype).view(*shape)

    def peer_table(self, off: int = 0):
        """ctypes array of 8 pointers: rank r's copy of the sub-buffer at `off` (0 beyond the group)."""
        ptrs = [(p + off) for p in self.peer_ptrs] + [0] * (MAX_TP - len(self.peer_ptrs))
        return (C.c_void_p * MAX_TP)(*ptrs)

    def mc(self, off: int = 0) -> int:
        return self.mc_ptr + off if self.mc_ptr else 0
