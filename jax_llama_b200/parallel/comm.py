"""Process bootstrap and NVLink peer memory for tensor parallelism.

The reference never names a collective: GSPMD derives NCCL all-reduces / all-gathers from the parameter
sharding.  Here communication is a first-class component:

* `init_distributed` — one process per GPU (torchrun env), `torch.distributed` used for bootstrap and as
  the measured NCCL baseline only;
* `PeerComm` — every rank allocates an *inbox* and a flag array with `cudaMalloc`, exports them with CUDA
  IPC, and maps every peer's pair.  The fused kernels in `csrc/comm.cu` store partial results straight into
  peers' inboxes over NVLink and synchronise with release/acquire flags that carry a monotonically
  increasing epoch (a device-side counter + the call's index since the last bump, so CUDA-graph replays need
  no host involvement).
"""
from __future__ import annotations

import ctypes as C
import logging
import os
from typing import Any, Dict, Optional, Tuple

import torch
import torch.distributed as dist

from ..ops import _lib
from ..ops import kernels as K

logger = logging.getLogger(__name__)

MAX_TP = 8

_EXTRA_SIGNATURES = {
    "jl_ipc_alloc": (C.c_int, [C.c_size_t, C.POINTER(C.c_void_p), C.c_void_p]),
    "jl_ipc_open": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    "jl_ipc_close": (C.c_int, [C.c_void_p]),
    "jl_ipc_free": (C.c_int, [C.c_void_p]),
    "jl_gemv_allreduce": (C.c_int, [C.c_void_p] * 4 + [C.c_int] * 5 + [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int,
                                                                          C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "jl_gemv_allreduce_flags": (C.c_int, [C.c_int]),
    "jl_allgather_small": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                     C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_int, C.c_void_p]),
    "jl_comm_bump": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p]),
    "jl_flag_exchange": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "jl_wait_flags": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "jl_memcpy_d2d": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "jl_rs_reduce_mm": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                  C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                  C.c_void_p, C.c_float, C.c_void_p]),
}


def _L() -> C.CDLL:
    lib = _lib.lib()
    if not getattr(lib, "_comm_ready", False):
        for name, (res, args) in _EXTRA_SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        lib._comm_ready = True
    return lib


def init_distributed(backend: Optional[str] = None) -> Tuple[int, int, int]:
    """Join the job described by RANK / WORLD_SIZE / LOCAL_RANK / MASTER_*.  Returns (rank, world, local)."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local)
            dist.init_process_group(backend, device_id=torch.device("cuda", local))
        else:
            dist.init_process_group(backend)
    return rank, world, local


class PeerComm:
    """Peer-mapped inbox + flags of one tensor-parallel group, and the fused collectives that use them."""

    def __init__(self, tp: int, rank: int, device: torch.device, inbox_bytes: int = 8 << 20,
                 n_flags: int = 1 << 16, group: Optional[dist.ProcessGroup] = None):
        if not 1 <= tp <= MAX_TP:
            raise ValueError(f"tensor-parallel degree {tp} outside 1..{MAX_TP}")
        self.tp, self.rank, self.device, self.group = tp, rank, device, group
        self.parity_bytes = (inbox_bytes // 2) // 256 * 256
        self.n_flags = n_flags
        L = _L()
        self._own_inbox, h_inbox = self._alloc(2 * self.parity_bytes)
        self._own_flags, h_flags = self._alloc(n_flags * 4)
        if tp > 1:
            gathered: list = [None] * tp
            dist.all_gather_object(gathered, (rank, h_inbox, h_flags), group=group)
        else:
            gathered = [(rank, h_inbox, h_flags)]
        self._opened = []
        inbox, flags = [0] * MAX_TP, [0] * MAX_TP
        for r, hi, hf in gathered:
            if r == rank:
                inbox[r], flags[r] = self._own_inbox, self._own_flags
            else:
                inbox[r], flags[r] = self._open(hi), self._open(hf)
        self.inbox_ptrs = (C.c_void_p * MAX_TP)(*inbox)
        self.flag_ptrs = (C.c_void_p * MAX_TP)(*flags)
        # the small all-gather gets the top 64 KB of each parity half to itself: the all-reduce region is read
        # with tag matching (decode_mega.cu) and must never see unrelated payload bytes
        self.gather_region = 64 << 10
        self.gather_ptrs = (C.c_void_p * MAX_TP)(*[(p + self.parity_bytes - self.gather_region) if p else 0
                                                   for p in inbox])
        self.seq = torch.zeros(1, dtype=torch.int32, device=device)     # device epoch counter
        self.call_index = 0                                               # collectives since the last bump
        self._flag_blocks: Dict[Tuple, int] = {}
        self._flag_top = 0
        if tp > 1:
            dist.barrier(group=group)          # everyone has mapped everyone before the first kernel runs
        del L

    # ------------------------------------------------------------------ IPC helpers
    def _alloc(self, nbytes: int) -> Tuple[int, bytes]:
        ptr = C.c_void_p()
        handle = C.create_string_buffer(64)
        _lib.check(_L().jl_ipc_alloc(nbytes, C.byref(ptr), handle), "jl_ipc_alloc")
        return ptr.value, handle.raw

    def _open(self, handle: bytes) -> int:
        ptr = C.c_void_p()
        _lib.check(_L().jl_ipc_open(C.create_string_buffer(handle, 64), C.byref(ptr)), "jl_ipc_open")
        self._opened.append(ptr.value)
        return ptr.value

    def close(self) -> None:
        L = _L()
        torch.cuda.synchronize(self.device)
        for p in self._opened:
            L.jl_ipc_close(p)
        self._opened = []
        for p in (self._own_inbox, self._own_flags):
            if p:
                L.jl_ipc_free(p)
        self._own_inbox = self._own_flags = 0

    def flag_block(self, key: Tuple, count: int) -> int:
        """Disjoint flag range per call-site layout (different slice counts must not alias)."""
        if key not in self._flag_blocks:
            if self._flag_top + count > self.n_flags:
                raise RuntimeError("out of communication flags")
            self._flag_blocks[key] = self._flag_top
            self._flag_top += count
        return self._flag_blocks[key]

    # ------------------------------------------------------------------ epochs
    def begin_sequence(self) -> None:
        """Start counting collectives of one step / forward (host side)."""
        self.call_index = 0

    def end_sequence(self) -> None:
        """Advance the device epoch counter past the collectives issued since `begin_sequence`."""
        if self.call_index:
            _lib.check(_L().jl_comm_bump(self.seq.data_ptr(), self.call_index, K._stream()), "jl_comm_bump")
        self.call_index = 0

    # ------------------------------------------------------------------ fused collectives
    def gemv_allreduce(self, x: torch.Tensor, w: torch.Tensor, residual: torch.Tensor,
                       out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """out = residual + all_reduce_sum(x_local · w_localᵀ) for M ≤ 8 rows, in one kernel per rank."""
        n, k = w.shape
        m = x.numel() // k
        if out is None:
            out = torch.empty((m, n), dtype=x.dtype, device=x.device)
        L = _L()
        need = self.tp * m * n * 4
        usable = self.parity_bytes - self.gather_region      # the top of each parity half belongs to allgather_small
        if need > usable:
            raise ValueError(f"inbox too small for a {m}x{n} all-reduce ({need} > {usable} bytes)")
        off = self.flag_block(("gemv_ar", n), int(L.jl_gemv_allreduce_flags(n)))
        _lib.check(L.jl_gemv_allreduce(x.data_ptr(), w.data_ptr(), residual.data_ptr(), out.data_ptr(), m, n, k,
                                       self.tp, self.rank, self.inbox_ptrs, self.flag_ptrs, self.parity_bytes // 4,
                                       off, self.seq.data_ptr(), self.call_index, K.num_sms(x.device), K._stream()),
                   "jl_gemv_allreduce")
        self.call_index += 1
        return out

    def allgather_small(self, src: torch.Tensor, out: torch.Tensor) -> torch.Tensor:
        """out[r] = src of rank r for a small, 16-byte-multiple buffer (top-k candidates)."""
        nbytes = src.numel() * src.element_size()
        if nbytes % 16:
            raise ValueError("allgather_small needs a multiple of 16 bytes")
        L = _L()
        if self.tp * nbytes > self.gather_region:
            raise ValueError(f"allgather_small payload {nbytes} B x {self.tp} exceeds the {self.gather_region} B region")
        off = self.flag_block(("ag",), MAX_TP)
        _lib.check(L.jl_allgather_small(src.data_ptr(), out.data_ptr(), nbytes, nbytes, self.tp, self.rank,
                                        self.gather_ptrs, self.flag_ptrs, self.parity_bytes, off,
                                        self.seq.data_ptr(), self.call_index, K._stream()), "jl_allgather_small")
        self.call_index += 1
        return out

    # ------------------------------------------------------------------ sequence-parallel prefill
    def symmetric_alloc(self, nbytes: int) -> Tuple[int, Any]:
        """cudaMalloc + IPC exchange: (local pointer, ctypes table of every rank's pointer to its copy)."""
        nbytes = (nbytes + 255) // 256 * 256
        local, handle = self._alloc(nbytes)
        if self.tp > 1:
            gathered: list = [None] * self.tp
            dist.all_gather_object(gathered, (self.rank, handle), group=self.group)
        else:
            gathered = [(self.rank, handle)]
        ptrs = [0] * MAX_TP
        for r, h in gathered:
            ptrs[r] = local if r == self.rank else self._open(h)
        self._symmetric = getattr(self, "_symmetric", []) + [local]
        if self.tp > 1:
            dist.barrier(group=self.group)
        return local, (C.c_void_p * MAX_TP)(*ptrs)

    def nvls_state(self, M: int, H: int, tag: int = 0) -> "NvlsPrefill":
        """Symmetric (multicast-mapped) buffers of the prefill collectives for one (tokens, hidden) shape; `tag`
        distinguishes independent chains of the same shape (the two halves of a pipelined prompt)."""
        key = (M, H, tag)
        states = self.__dict__.setdefault("_nvls_states", {})
        if key not in states:
            states[key] = NvlsPrefill(self, M, H, tag)
        return states[key]

    def nvls_end_forward(self) -> None:
        """End of a forward: every NVLS prefill state that issued collectives advances its device epoch base."""
        for st in self.__dict__.get("_nvls_states", {}).values():
            if st.rel:
                st.end_forward()


class NvlsPrefill:
    """Prefill collectives over an NVLS multicast mapping (csrc/sp_prefill.cu, `rs_reduce_mm_kernel`).

    part[2]    [M][H] bf16   this rank's partial product of a row-parallel GEMM (wo / w2 alternate between the two);
                             the GEMM stores it locally, peers read it only through ``multimem.ld_reduce``
    a_full[2]  [M][H] bf16   gathered, normalised activations, double-buffered by parity; row owners fill every rank's
                             copy with ``multimem.st``

    Two ways to use it, both without NCCL:
    * sequence parallel (token count a multiple of 128·tp): `rs_reduce` on the rank's row shard, `ag_gemm`;
    * `allreduce_add` for any token count: every rank pulls the in-switch sum of all rows into its replicated stream.

    No kernel that moves data ever spins on a peer: the "my GEMM is done" exchange and the wait for the owners' "rows
    are in every copy" flags are one-warp kernels (`jl_flag_exchange`, `jl_wait_flags`) between them, so two of these
    chains can run on two streams (the engine pipelines two halves of the prompt) without one chain's waiting kernel
    keeping the other's producer off the SMs.
    """

    def __init__(self, comm: PeerComm, M: int, H: int, tag: int = 0):
        from .symm import SymmetricHeap
        self.comm, self.M, self.H = comm, M, H
        tp = comm.tp
        self.sp_ok = M % (tp * 128) == 0
        self.mloc = M // tp if self.sp_ok else 0
        one = (M * H * 2 + 1023) // 1024 * 1024
        self.heap = SymmetricHeap(tp, comm.rank, comm.device, 4 * one + 4096, group=comm.group)
        if tp > 1 and not self.heap.has_multicast:
            raise RuntimeError("this platform gave the symmetric heap no NVLS multicast address")
        self.part_off = [self.heap.alloc(one), self.heap.alloc(one)]
        self.a_off = [self.heap.alloc(one), self.heap.alloc(one)]
        self.part = [self.heap.view(o, (M, H), torch.bfloat16) for o in self.part_off]
        self.a_full = [self.heap.view(o, (M, H), torch.bfloat16) for o in self.a_off]
        self.ticket = torch.zeros(1, dtype=torch.int32, device=comm.device)
        self.gemm_off = comm.flag_block(("nvls_gemm", M, H, tag), MAX_TP)
        self.ready_off = comm.flag_block(("nvls_ready", M, H, tag), MAX_TP)
        self.parity = 0
        self.slot = 0
        self.use_nccl_baseline = False          # bench only: time the NCCL all-reduce in place of the in-switch pull
        # Flag values are (device epoch base) + (position of the collective inside this forward): the base lives on the
        # device and is advanced by one tiny kernel at the end of every forward (`end_forward`), so a forward captured
        # in a CUDA graph stays correct on every replay; `rel` only counts inside one forward.
        self.base = torch.zeros(1, dtype=torch.int32, device=comm.device)
        self.rel = 0
        self.last_ready = 0

    def next_part(self) -> torch.Tensor:
        """Output buffer of the next row-parallel GEMM (its partial product stays on this rank)."""
        self.slot ^= 1
        return self.part[self.slot]

    def _exchange_gemm_done(self) -> None:
        cm = self.comm
        self.rel += 1
        _lib.check(_L().jl_flag_exchange(cm.flag_ptrs, self.gemm_off, self.rel, self.base.data_ptr(), cm.tp, cm.rank,
                                         K._stream()), "jl_flag_exchange")

    def _wait_ready(self) -> None:
        cm = self.comm
        _lib.check(_L().jl_wait_flags(cm._own_flags + 4 * self.ready_off, self.last_ready, self.base.data_ptr(), cm.tp,
                                      cm.rank, K._stream()), "jl_wait_flags")

    def end_forward(self) -> None:
        """Advance the device epoch base past the collectives of this forward (on the current stream)."""
        if self.rel:
            _lib.check(_L().jl_comm_bump(self.base.data_ptr(), self.rel, K._stream()), "jl_comm_bump")
        self.rel = 0
        self.last_ready = 0

    def allreduce_add(self, x: torch.Tensor) -> torch.Tensor:
        """x += Σ_ranks part (the buffer handed out by the last `next_part`), all rows, replicated result."""
        cm = self.comm
        if self.use_nccl_baseline:
            y = self.part[self.slot].clone()
            dist.all_reduce(y, group=cm.group)
            return x.add_(y)
        self._exchange_gemm_done()
        _lib.check(_L().jl_rs_reduce_mm(x.data_ptr(), self.heap.mc(self.part_off[self.slot]), None, None, None, self.M,
                                        self.H, 2, cm.tp, cm.rank, cm.flag_ptrs, self.gemm_off, self.ready_off,
                                        0, 0, None, self.ticket.data_ptr(), 0.0, K._stream()), "jl_rs_reduce_mm")
        return x

    def allreduce_add_norm(self, x: torch.Tensor, norm_w: torch.Tensor, eps: float) -> torch.Tensor:
        """Batched decode (M <= 128 rows): x += Σ_ranks part, and the rows RMS-normalised with `norm_w` come back in a
        local buffer — flag exchange, in-switch reduction, residual add and the next norm in ONE launch (the prompt-sized
        form keeps the exchange in its own one-warp kernel, see the class docstring)."""
        cm = self.comm
        if self.M > 128 or self.use_nccl_baseline:
            raise ValueError("allreduce_add_norm is the batched-decode form (M <= 128)")
        if not hasattr(self, "xn_local"):
            self.xn_local = [torch.empty((self.M, self.H), dtype=torch.bfloat16, device=cm.device) for _ in range(2)]
            self.xn_slot = 0
        self.xn_slot ^= 1
        out = self.xn_local[self.xn_slot]
        self.rel += 1
        _lib.check(_L().jl_rs_reduce_mm(x.data_ptr(), self.heap.mc(self.part_off[self.slot]), norm_w.data_ptr(), None,
                                        out.data_ptr(), self.M, self.H, 1, cm.tp, cm.rank, cm.flag_ptrs, self.gemm_off,
                                        self.ready_off, self.rel, 0, self.base.data_ptr(), self.ticket.data_ptr(),
                                        float(eps), K._stream()), "jl_rs_reduce_mm")
        return out

    def rs_reduce(self, x_shard: torch.Tensor, norm_w: torch.Tensor, eps: float, has_partials: bool) -> None:
        """Owner step: x_shard += in-switch Σ of this rank's rows; RMSNorm → every rank's a_full[next parity]; publish."""
        cm = self.comm
        self.parity ^= 1
        if has_partials:
            self._exchange_gemm_done()
        self.rel += 1
        self.last_ready = self.rel
        row0 = cm.rank * self.mloc * self.H * 2
        part_mc = self.heap.mc(self.part_off[self.slot] + row0) if has_partials else None
        _lib.check(_L().jl_rs_reduce_mm(x_shard.data_ptr(), part_mc, norm_w.data_ptr(),
                                        self.heap.mc(self.a_off[self.parity] + row0), None, self.mloc, self.H,
                                        2 if has_partials else 0, cm.tp, cm.rank, cm.flag_ptrs, self.gemm_off,
                                        self.ready_off, 0, self.last_ready, self.base.data_ptr(), self.ticket.data_ptr(),
                                        float(eps), K._stream()), "jl_rs_reduce_mm")

    def ag_gemm(self, w: torch.Tensor, mode: int) -> torch.Tensor:
        """y = (gathered normalised rows) · wᵀ: a one-warp kernel waits for every owner's "ready" flag, then the plain
        tcgen05 GEMM reads this rank's copy of the gathered activations."""
        self._wait_ready()
        return K.linear(self.a_full[self.parity], w, mode=mode)

    def read_rows(self, rows: torch.Tensor) -> torch.Tensor:
        """Rows (global token indices) of the current gathered, normalised activations (every rank holds all rows)."""
        self._wait_ready()
        return self.a_full[self.parity][rows]
