"""Allocation-free decode attention (`jl_decode_attn_noalloc`) checked through the C ABI.

The entry point exists for callers that cannot allocate device memory on their own — for instance a
serving process whose HBM already belongs to another runtime's arena, or a host that forbids
cudaMalloc/cudaFree inside its token loop because they take the allocator lock and may synchronise the
device.  Its split-KV workspace is a fixed array inside the library, so this test calls the symbol
straight from the shared object with raw pointers (the way such a caller would) and compares the result
with the ordinary entry point, `jl_decode_attn`, that receives a caller-owned workspace.  Two calls are
made back to back with different shapes and inputs: the array is shared between calls and keeps the
first call's contents, and the second call must not be disturbed by them.

The file name sorts last on purpose: under `pytest -x` every other test has already run by the time
this one does, and it must pass both on its own in a fresh process and at the end of a full run.
"""
import ctypes as C
import math

import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = 1.5e-2     # same relative tolerance as the decode-attention tests in test_gpu_kernels.py


def _abi():
    from jax_llama_b200.ops.build import LIB_PATH, build_library

    if not LIB_PATH.exists():
        build_library()
    lib = C.CDLL(str(LIB_PATH))
    vp, i32, f32 = C.c_void_p, C.c_int, C.c_float
    lib.jl_last_error.restype = C.c_char_p
    lib.jl_decode_attn_noalloc_max_floats.restype = C.c_size_t
    lib.jl_decode_attn_noalloc.restype = i32
    lib.jl_decode_attn_noalloc.argtypes = [vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, f32, i32, i32, vp]
    lib.jl_decode_attn.restype = i32
    lib.jl_decode_attn.argtypes = [vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, f32, i32, i32, vp]
    return lib


def _case(seed, B, nh, nkv, hd, length, splits):
    g = torch.Generator().manual_seed(seed)
    T = length + 3
    mk = lambda *s: torch.randn(*s, generator=g).to(torch.bfloat16).cuda()      # noqa: E731
    q, kc, vc = mk(B, nh, hd), mk(B, nkv, T, hd), mk(B, nkv, T, hd)
    pad = torch.tensor([0, 5, 1, 0][:B], dtype=torch.int32).cuda()
    clen = torch.tensor([length - 1], dtype=torch.int32).cuda()
    return dict(q=q, kc=kc, vc=vc, pad=pad, clen=clen, B=B, nh=nh, nkv=nkv, hd=hd, T=T, splits=splits)


def _run_both(lib, c):
    stream = torch.cuda.current_stream().cuda_stream
    scale = 1.0 / math.sqrt(c["hd"])
    want = torch.empty(c["B"], c["nh"] * c["hd"], dtype=torch.bfloat16, device="cuda")
    got = torch.full_like(want, float("nan"))
    scratch = torch.empty(c["B"] * c["nh"] * c["splits"] * (c["hd"] + 2), dtype=torch.float32, device="cuda")
    rc = lib.jl_decode_attn(c["q"].data_ptr(), c["kc"].data_ptr(), c["vc"].data_ptr(), c["clen"].data_ptr(),
                            c["pad"].data_ptr(), scratch.data_ptr(), want.data_ptr(), c["B"], c["nh"], c["nkv"],
                            c["hd"], c["T"], 1, scale, c["splits"], 0, stream)
    assert rc == 0, lib.jl_last_error().decode()
    rc = lib.jl_decode_attn_noalloc(c["q"].data_ptr(), c["kc"].data_ptr(), c["vc"].data_ptr(), c["clen"].data_ptr(),
                                    c["pad"].data_ptr(), got.data_ptr(), c["B"], c["nh"], c["nkv"], c["hd"],
                                    c["T"], 1, scale, c["splits"], 0, stream)
    assert rc == 0, lib.jl_last_error().decode()
    torch.cuda.synchronize()
    return got.float().cpu(), want.float().cpu()


def test_noalloc_matches_caller_scratch_entry_point_twice_in_a_row():
    lib = _abi()
    assert lib.jl_decode_attn_noalloc_max_floats() == 1 << 20
    # call 1: GQA group 4, long context, many splits;  call 2 (right after, smaller and different layout of
    # the shared workspace): MHA, head_dim 64, fewer splits — stale partials from call 1 must not leak in
    for case in (_case(1, 3, 8, 2, 128, 1500, 8), _case(2, 2, 4, 4, 64, 333, 3)):
        got, want = _run_both(lib, case)
        assert torch.isfinite(got).all()
        err = ((got - want).abs().max() / (want.abs().max() + 1e-6)).item()
        assert err < TOL, err


def test_noalloc_rejects_requests_beyond_the_builtin_workspace():
    lib = _abi()
    c = _case(3, 4, 8, 2, 128, 64, 2)
    out = torch.empty(c["B"], c["nh"] * c["hd"], dtype=torch.bfloat16, device="cuda")
    # B*nh*splits*(hd+2) = 64*64*64*130 floats is far beyond 2^20: must be refused before anything is launched
    rc = lib.jl_decode_attn_noalloc(c["q"].data_ptr(), c["kc"].data_ptr(), c["vc"].data_ptr(), c["clen"].data_ptr(),
                                    c["pad"].data_ptr(), out.data_ptr(), 64, 64, 16, 128, c["T"], 1, 0.1, 64, 0,
                                    torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    msg = lib.jl_last_error().decode()
    assert "exceeds the built-in scratch" in msg and "1048576" in msg
    torch.cuda.synchronize()      # and the context is still healthy
