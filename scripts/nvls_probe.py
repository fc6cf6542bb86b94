"""Box probe (torchrun, one rank per GPU): does the symmetric heap come up with an NVLS multicast address, and do
our multimem.ld_reduce / multimem.st / multimem.red instructions give the right answers on it?  Also times a
tagged-packet broadcast round trip and an in-switch reduction so the decode all-reduce can be sized from data."""
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from jax_llama_b200.ops import _lib  # noqa: E402
from jax_llama_b200.parallel.comm import init_distributed  # noqa: E402
from jax_llama_b200.parallel.symm import SymmetricHeap  # noqa: E402


def main() -> None:
    rank, world, local = init_distributed("nccl")
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    heap = SymmetricHeap(world, rank, dev, 8 << 20)
    L = _lib.lib()
    L.jl_nvls_selftest.restype = C.c_int
    L.jl_nvls_selftest.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_uint, C.c_void_p]
    res = {"world": world, "backend": heap.backend, "mc": hex(heap.mc_ptr)}
    if not heap.has_multicast:
        if rank == 0:
            print("NVLS_PROBE " + json.dumps(res))
        return
    s = torch.cuda.current_stream().cuda_stream
    n = 4096
    f32 = heap.view(0, (n,), torch.float32)
    f32.copy_(torch.arange(n, device=dev, dtype=torch.float32) * (rank + 1))
    torch.cuda.synchronize(); dist.barrier()
    out = torch.zeros(n, device=dev)
    _lib.check(L.jl_nvls_selftest(heap.mc(0), None, out.data_ptr(), n, 0, 0, s), "ld_reduce")
    torch.cuda.synchronize(); dist.barrier()
    want = torch.arange(n, device=dev, dtype=torch.float32) * (world * (world + 1) // 2)
    res["ld_reduce_f32"] = bool(torch.equal(out, want))
    # packets: rank r broadcasts into slot block r
    heap.alloc(n * 4)                                   # the fp32 block used above
    off = heap.alloc(world * n * 8)
    src = torch.arange(n, device=dev, dtype=torch.float32) + 1000 * rank
    _lib.check(L.jl_nvls_selftest(heap.mc(off + rank * n * 8), src.data_ptr(), None, n, 1, 77, s), "st")
    torch.cuda.synchronize(); dist.barrier()
    pk = heap.view(off, (world, n, 2), torch.int32)
    ok = True
    for r in range(world):
        ok &= bool(torch.equal(pk[r, :, 0].view(torch.float32), torch.arange(n, device=dev, dtype=torch.float32) + 1000 * r))
        ok &= bool((pk[r, :, 1] == 77).all())
    res["st_packets"] = ok
    off2 = heap.alloc(n * 4)
    _lib.check(L.jl_nvls_selftest(heap.mc(off2), torch.ones(n, device=dev).data_ptr(), None, n, 2, 0, s), "red")
    torch.cuda.synchronize(); dist.barrier()
    res["red_f32"] = bool((heap.view(off2, (n,), torch.float32) == world).all())
    off3 = heap.alloc(n * 2)
    heap.view(off3, (n,), torch.bfloat16).fill_(float(rank + 1))
    torch.cuda.synchronize(); dist.barrier()
    outb = torch.zeros(n, device=dev, dtype=torch.bfloat16)
    _lib.check(L.jl_nvls_selftest(heap.mc(off3), None, outb.data_ptr(), n, 3, 0, s), "ld_reduce bf16")
    torch.cuda.synchronize(); dist.barrier()
    res["ld_reduce_bf16"] = bool((outb.float() == world * (world + 1) / 2).all())
    off4 = heap.alloc(256)
    for _ in range(3):
        _lib.check(L.jl_nvls_selftest(heap.mc(off4), None, None, 1, 5, 0, s), "red u32")
    torch.cuda.synchronize(); dist.barrier()
    res["red_u32"] = int(heap.view(off4, (1,), torch.int32).item()) == 3 * world
    # timing: 4096-value in-switch reduction (one launch), averaged
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(5):
        L.jl_nvls_selftest(heap.mc(0), None, out.data_ptr(), n, 0, 0, s)
    torch.cuda.synchronize(); dist.barrier()
    e0.record()
    for _ in range(100):
        L.jl_nvls_selftest(heap.mc(0), None, out.data_ptr(), n, 0, 0, s)
    e1.record()
    torch.cuda.synchronize()
    res["ld_reduce_16KB_launch_us"] = round(e0.elapsed_time(e1) * 10, 2)
    big = 1 << 20
    offb = heap.alloc(big * 4)
    outbig = torch.zeros(big, device=dev)
    for _ in range(3):
        L.jl_nvls_selftest(heap.mc(offb), None, outbig.data_ptr(), big, 0, 0, s)
    torch.cuda.synchronize(); dist.barrier()
    e0.record()
    for _ in range(20):
        L.jl_nvls_selftest(heap.mc(offb), None, outbig.data_ptr(), big, 0, 0, s)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    res["ld_reduce_4MB_us"] = round(ms * 1e3, 1)
    res["ld_reduce_4MB_GBs_out"] = round(big * 4 / (ms / 1e3) / 1e9, 1)
    dist.barrier()
    if rank == 0:
        print("NVLS_PROBE " + json.dumps(res))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
