"""ctypes binding of libjllama_b200.so (the C ABI in include/jllama_b200.h).

Loading is lazy and *loud*: on a machine with a CUDA device, a missing or stale library is an error,
never a silent fall back to PyTorch ops.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

from .build import LIB_PATH, build_library

_lib: Optional[C.CDLL] = None

vp, ip, fp = C.c_void_p, C.c_void_p, C.c_void_p   # all pointers travel as integers from tensor.data_ptr()
i32, f32, u64, sz = C.c_int, C.c_float, C.c_ulonglong, C.c_size_t

_SIGNATURES = {
    "jl_last_error": (C.c_char_p, []),
    "jl_version": (i32, []),
    "jl_launch_count": (u64, []),
    "jl_embed": (i32, [ip, vp, vp, i32, i32, i32, vp]),
    "jl_rmsnorm": (i32, [vp, vp, vp, vp, vp, i32, i32, f32, vp]),
    "jl_rope_kv_write": (i32, [vp, vp, vp, vp, fp, fp, ip, ip, ip, i32, i32, i32, i32, i32, i32, i32, vp]),
    "jl_swiglu": (i32, [vp, vp, i32, i32, i32, vp]),
    "jl_gemv": (i32, [vp, vp, vp, vp, vp, i32, i32, i32, f32, i32, i32, i32, vp]),
    "jl_gemv_tma": (i32, [vp, vp, vp, vp, vp, i32, i32, i32, f32, i32, i32, i32, vp]),
    "jl_gemm": (i32, [vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp]),
    "jl_gemm_2cta": (i32, [vp, vp, vp, vp, i32, i32, i32, i32, i32, vp]),
    "jl_gemm_set_workspace": (i32, [vp, sz, vp, i32]),
    "jl_gemm_set_trace": (None, [vp]),
    "jl_decode_attn_scratch_bytes": (sz, [i32, i32, i32, i32]),
    "jl_decode_attn_splits": (i32, [i32, i32, i32, i32]),
    "jl_decode_attn": (i32, [vp, vp, vp, ip, ip, vp, vp, i32, i32, i32, i32, i32, i32, f32, i32, i32, vp]),
    "jl_decode_attn_fused_scratch_bytes": (sz, [i32, i32, i32, i32, i32]),
    "jl_decode_attn_fused_splits": (i32, [i32, i32, i32, i32]),
    "jl_decode_attn_fused": (i32, [vp, vp, vp, fp, fp, ip, ip, vp, vp, i32, i32, i32, i32, i32, i32, f32, i32, vp]),
    "jl_decode_attn_fused_set_variant": (i32, [i32]),
    "jl_decode_mega_scratch_bytes": (sz, [i32, i32, i32, i32, i32, i32]),
    "jl_decode_mega_cand_offset": (sz, [i32, i32, i32, i32, i32, i32]),
    "jl_decode_mega_peer_bytes": (sz, [i32, i32, i32, i32, vp]),
    "jl_nvls_bcast_f32": (i32, [fp, fp, i32, i32, i32, i32, i32, i32, i32, vp, i32, i32, ip, ip, vp]),
    "jl_sample_full": (i32, [fp, i32, i32, i32, f32, i32, f32, u64, ip, ip, vp, vp, i32, i32, i32, i32, ip, ip, ip, ip,
                             i32, i32, i32, ip, ip, ip, ip, vp]),
    "jl_decode_mega": (i32, [vp] * 9 + [ip, fp, i32, i32, i32, vp, vp, fp, fp, ip, ip, vp, vp] + [i32] * 11 +
                       [f32, f32, i32, i32, i32, vp, vp, vp, vp, vp]),
    "jl_mega_select": (i32, [vp, vp] + [i32] * 5 + [ip] * 4 + [i32] * 3 + [ip] * 4 + [vp]),
    "jl_nvls_selftest": (i32, [vp, vp, vp, i32, i32, C.c_uint, vp]),
    "jl_argmax": (i32, [fp, i32, i32, i32, i32, fp, ip, vp]),
    "jl_topk": (i32, [fp, i32, i32, i32, i32, i32, fp, ip, vp]),
    "jl_sample": (i32, [fp, ip, i32, i32, sz, i32, i32, f32, f32, u64, ip, ip, vp]),
    "jl_decode_advance": (i32, [ip, ip, ip, ip, i32, i32, i32, i32, ip, ip, ip, ip, vp]),
}

_OPTIONAL = {
    "jl_flash_prefill_occupancy": (i32, [i32]),
    "jl_decode_mega_stages": (i32, [i32, i32, i32, i32, i32, i32]),
    "jl_flash_prefill": (i32, [vp, vp, vp, ip, ip, vp, i32, i32, i32, i32, i32, i32, i32, f32, vp]),
    "jl_prefill_attn_mma": (i32, [vp, vp, vp, ip, vp, i32, i32, i32, i32, i32, i32, i32, f32, vp]),
}


class NativeLibraryError(RuntimeError):
    pass


def lib() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    path = LIB_PATH
    if os.environ.get("JLLAMA_LIB"):           # A/B of two builds on one box (scripts/sweep_env.sh JLLAMA_LIB "a.so b.so")
        from pathlib import Path
        path = Path(os.environ["JLLAMA_LIB"])
    if not path.exists():
        if os.environ.get("JLLAMA_NO_AUTOBUILD"):
            raise NativeLibraryError(f"{path} is missing: run `python -m jax_llama_b200.ops.build`")
        build_library()
    try:
        handle = C.CDLL(str(path))
    except OSError as e:
        raise NativeLibraryError(f"cannot load {path}: {e}") from e
    for name, (res, args) in {**_SIGNATURES, **_OPTIONAL}.items():
        try:
            fn = getattr(handle, name)
        except AttributeError:
            if name in _OPTIONAL:
                continue
            raise NativeLibraryError(f"{path} is stale: symbol {name} missing; rebuild with "
                                     "`python -m jax_llama_b200.ops.build --force`")
        fn.restype, fn.argtypes = res, args
    _lib = handle
    return handle


def has(symbol: str) -> bool:
    return hasattr(lib(), symbol)


def check(rc: int, what: str) -> None:
    if rc != 0:
        raise RuntimeError(f"{what} failed (code {rc}): {lib().jl_last_error().decode()}")


def launch_count() -> int:
    return int(lib().jl_launch_count())
