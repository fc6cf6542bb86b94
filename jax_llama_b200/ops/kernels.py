"""Python front-ends of the sm_100a kernels.

Every op takes torch tensors.  CUDA bf16 tensors run the hand-written kernel through the C ABI on the
current stream (legal under CUDA-graph capture); CPU tensors run the fp32 PyTorch statement of the same
op, which is also the oracle the GPU tests compare against.  There is no third path: a CUDA tensor with
no native library raises.
"""
from __future__ import annotations

import math
import os
from typing import Optional, Tuple

import torch
import torch.nn.functional as F

from . import _lib

PLAIN, RESIDUAL, SWIGLU = 0, 1, 2
_NUM_SMS = {}
_SMALL_M_PATH = os.environ.get("JLLAMA_SMALL_M", "gemv")    # measured faster in the graph-replayed step than "gemv_tma"


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def num_sms(device: Optional[torch.device] = None) -> int:
    idx = torch.cuda.current_device() if device is None or device.index is None else device.index
    if idx not in _NUM_SMS:
        _NUM_SMS[idx] = torch.cuda.get_device_properties(idx).multi_processor_count
    return _NUM_SMS[idx]


_GEMM_WS: dict = {}


def ensure_gemm_workspace(device: Optional[torch.device] = None, nbytes: int = 32 << 20) -> None:
    """Give the library its split-K scratch for this device (fp32 partials + per-tile tickets).  Small-batch
    GEMMs with few weight-row tiles (N = 4096 at batch 9..128) then spread their K range over all SMs.  Done once,
    outside any stream capture (the engine calls it at construction); without it such GEMMs simply do not split."""
    idx = torch.cuda.current_device() if device is None or device.index is None else device.index
    if idx in _GEMM_WS or torch.cuda.is_current_stream_capturing() or os.environ.get("JLLAMA_NO_SPLITK") == "1":
        return
    with torch.cuda.device(idx):
        ws = torch.empty(nbytes // 4, dtype=torch.float32, device=f"cuda:{idx}")
        tickets = torch.zeros(2048, dtype=torch.int32, device=f"cuda:{idx}")
        _lib.check(_lib.lib().jl_gemm_set_workspace(ws.data_ptr(), ws.numel() * 4, tickets.data_ptr(), tickets.numel()),
                   "jl_gemm_set_workspace")
    _GEMM_WS[idx] = (ws, tickets)


def _p(t: Optional[torch.Tensor]) -> int:
    return 0 if t is None else t.data_ptr()


def _bf16c(t: torch.Tensor, name: str) -> None:
    if t.dtype != torch.bfloat16 or not t.is_contiguous():
        raise ValueError(f"{name} must be a contiguous bf16 tensor (got {t.dtype}, contiguous={t.is_contiguous()})")


# --------------------------------------------------------------------------- embedding
def embed(ids: torch.Tensor, table: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    if not table.is_cuda:
        return table[ids.long().clamp(0, table.shape[0] - 1)]
    _bf16c(table, "table")
    ids32 = ids.to(torch.int32).contiguous().view(-1)
    if out is None:
        out = torch.empty((ids32.numel(), table.shape[1]), dtype=table.dtype, device=table.device)
    _lib.check(_lib.lib().jl_embed(ids32.data_ptr(), table.data_ptr(), out.data_ptr(), ids32.numel(),
                                   table.shape[1], table.shape[0], _stream()), "jl_embed")
    return out


# --------------------------------------------------------------------------- RMSNorm
def rmsnorm(x: torch.Tensor, w: torch.Tensor, eps: float, delta: Optional[torch.Tensor] = None,
            out: Optional[torch.Tensor] = None, update_x: bool = True) -> torch.Tensor:
    """out = RMSNorm(x (+ delta)) · w.  With `delta`, x is updated in place to x + delta (residual)."""
    if not x.is_cuda:
        if delta is not None:
            s = x + delta
            if update_x:
                x.copy_(s)
        else:
            s = x
        xf = s.float()
        return (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps) * w.float()).to(x.dtype)
    _bf16c(x, "x"); _bf16c(w, "w")
    if delta is not None:
        _bf16c(delta, "delta")
    h = x.shape[-1]
    m = x.numel() // h
    if out is None:
        out = torch.empty_like(x)
    _lib.check(_lib.lib().jl_rmsnorm(x.data_ptr(), _p(delta), w.data_ptr(), out.data_ptr(),
                                     x.data_ptr() if (delta is not None and update_x) else 0,
                                     m, h, float(eps), _stream()), "jl_rmsnorm")
    return out


# --------------------------------------------------------------------------- RoPE + KV append
def rope_kv_write(qkv: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, cos: torch.Tensor,
                  sin: torch.Tensor, B: int, S: int, nh: int, nkv: int, hd: int,
                  cache_len: Optional[torch.Tensor] = None, pad_left: Optional[torch.Tensor] = None,
                  positions: Optional[torch.Tensor] = None, q_out: Optional[torch.Tensor] = None) -> None:
    """qkv [B·S, (nh+2nkv)·hd]; caches [B, nkv, T, hd]; cos/sin fp32 [max_pos, hd/2].  In-place on qkv's q
    block (or into q_out) and on the caches.  cache_len: int32 scalar tensor (device) or None (= 0)."""
    T = k_cache.shape[2]
    if not qkv.is_cuda:
        base = int(cache_len.item()) if cache_len is not None else 0
        x = qkv.view(B, S, nh + 2 * nkv, hd)
        slots = base + torch.arange(S)
        if positions is not None:
            pos = positions.view(B, S).long()
        else:
            pl = pad_left.long() if pad_left is not None else torch.zeros(B, dtype=torch.long)
            pos = (slots[None, :] - pl[:, None])
        pos = pos.clamp(0, cos.shape[0] - 1)
        c, s = cos[pos][:, :, None, :], sin[pos][:, :, None, :]          # [B,S,1,hd/2]
        qk = x[:, :, :nh + nkv].float()
        re, im = qk[..., 0::2], qk[..., 1::2]
        rot = torch.stack([re * c - im * s, re * s + im * c], -1).flatten(-2).to(qkv.dtype)
        if q_out is not None:
            q_out.view(B, S, nh, hd).copy_(rot[:, :, :nh])
        else:
            x[:, :, :nh] = rot[:, :, :nh]
        k_cache[:, :, base:base + S] = rot[:, :, nh:].transpose(1, 2)
        v_cache[:, :, base:base + S] = x[:, :, nh + nkv:].transpose(1, 2)
        return
    _bf16c(qkv, "qkv"); _bf16c(k_cache, "k_cache"); _bf16c(v_cache, "v_cache")
    if cos.dtype != torch.float32 or sin.dtype != torch.float32:
        raise ValueError("cos/sin tables must be fp32")
    if positions is not None:
        positions = positions.to(torch.int32).contiguous()
    _lib.check(_lib.lib().jl_rope_kv_write(qkv.data_ptr(), _p(q_out), k_cache.data_ptr(), v_cache.data_ptr(),
                                           cos.data_ptr(), sin.data_ptr(), _p(cache_len), _p(pad_left),
                                           _p(positions), B, S, nh, nkv, hd, T, cos.shape[0], _stream()),
               "jl_rope_kv_write")


# --------------------------------------------------------------------------- SwiGLU
def swiglu(x: torch.Tensor, interleaved: bool = True, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    f = x.shape[-1] // 2
    if not x.is_cuda:
        g, u = (x[..., 0::2], x[..., 1::2]) if interleaved else (x[..., :f], x[..., f:])
        return (F.silu(g.float()) * u.float()).to(x.dtype)
    _bf16c(x, "x")
    m = x.numel() // (2 * f)
    if out is None:
        out = torch.empty((*x.shape[:-1], f), dtype=x.dtype, device=x.device)
    _lib.check(_lib.lib().jl_swiglu(x.data_ptr(), out.data_ptr(), m, f, int(interleaved), _stream()), "jl_swiglu")
    return out


# --------------------------------------------------------------------------- linear layers
def _linear_ref(x, w, norm_w, eps, residual, mode, out_f32):
    xf = x.float()
    if norm_w is not None:
        xf = (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps) * norm_w.float()).to(x.dtype).float()
    y = (xf @ w.float().t())
    if out_f32:
        return y
    y = y.to(x.dtype)
    if mode == RESIDUAL:
        return (y.float() + residual.float()).to(x.dtype)
    if mode == SWIGLU:
        return (F.silu(y[..., 0::2].float()) * y[..., 1::2].float()).to(x.dtype)
    return y


def linear(x: torch.Tensor, w: torch.Tensor, *, norm_w: Optional[torch.Tensor] = None, eps: float = 1e-5,
           residual: Optional[torch.Tensor] = None, mode: int = PLAIN, out_f32: bool = False,
           out: Optional[torch.Tensor] = None, force: Optional[str] = None) -> torch.Tensor:
    """y = f(x) Wᵀ with W [N, K] for the per-op path (the whole-step decode kernel has its own GEMV).  Dispatch
    by token count: M ≤ 4 → register-streaming GEMV with RMSNorm fused (JLLAMA_SMALL_M=gemv_tma selects the
    TMA-ring GEMV, which falls back when the activations do not fit next to its ring), otherwise the tcgen05
    GEMM (swap-AB up to M = 128).  `force` ∈ {"gemv_tma", "gemv", "gemm"} pins the kernel."""
    if not x.is_cuda:
        return _linear_ref(x, w, norm_w, eps, residual, mode, out_f32)
    _bf16c(x, "x"); _bf16c(w, "w")
    n, k = w.shape
    m = x.numel() // k
    odim = n // 2 if mode == SWIGLU else n
    if out is None:
        out = torch.empty((m, odim), dtype=torch.float32 if out_f32 else x.dtype, device=x.device)
    sms = num_sms(x.device)
    if force is None:
        path = ("gemv_tma" if _SMALL_M_PATH == "gemv_tma" else "gemv") if m <= 4 else "gemm"
    else:
        path = force
    L = _lib.lib()
    if path == "gemv_tma":
        rc = L.jl_gemv_tma(x.data_ptr(), w.data_ptr(), _p(norm_w), _p(residual), out.data_ptr(), m, n, k,
                           float(eps), mode, int(out_f32), sms, _stream())
        if rc == 3 and force is None:
            path = "gemv"                      # activations too large for the TMA ring: register-streaming kernel
        else:
            _lib.check(rc, "jl_gemv_tma")
    if path == "gemv":
        if m > 8:
            raise ValueError("gemv path handles at most 8 rows")
        _lib.check(L.jl_gemv(x.data_ptr(), w.data_ptr(), _p(norm_w), _p(residual), out.data_ptr(), m, n, k,
                             float(eps), mode, int(out_f32), sms, _stream()), "jl_gemv")
    elif path == "gemm":
        if norm_w is not None:
            x = rmsnorm(x.view(m, k), norm_w, eps)
        if m <= 128:
            ensure_gemm_workspace(x.device)
        _lib.check(L.jl_gemm(x.data_ptr(), w.data_ptr(), _p(residual), out.data_ptr(), m, n, k, mode,
                             int(out_f32), sms, _stream()), "jl_gemm")
    elif path == "gemm_2cta":
        # the CTA-pair kernel alone (tests): shapes it does not cover are an error here, not a fallback
        if norm_w is not None:
            x = rmsnorm(x.view(m, k), norm_w, eps)
        rc = L.jl_gemm_2cta(x.data_ptr(), w.data_ptr(), _p(residual), out.data_ptr(), m, n, k, mode, sms, _stream())
        if rc == 3:
            raise ValueError(f"jl_gemm_2cta does not cover M={m}, N={n}, K={k} (needs M >= 256, N % 256 == 0, K % 64 == 0)")
        _lib.check(rc, "jl_gemm_2cta")
    elif path != "gemv_tma":
        raise ValueError(f"unknown linear path {path!r}")
    return out


# --------------------------------------------------------------------------- decode attention
def decode_attn_plan(B: int, nh: int, nkv: int, hd: int, max_len: int, device: torch.device
                     ) -> Tuple[int, Optional[torch.Tensor]]:
    """(splits, scratch) for jl_decode_attn at this shape; scratch is reused across layers and steps."""
    L = _lib.lib()
    splits = int(L.jl_decode_attn_splits(B, nkv, max_len, num_sms(device)))
    nbytes = int(L.jl_decode_attn_scratch_bytes(B, nh, hd, splits))
    scratch = torch.empty(max(nbytes, 4) // 4, dtype=torch.float32, device=device) if splits > 1 else None
    return splits, scratch


def decode_attn(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
                cache_len: Optional[torch.Tensor], pad_left: Optional[torch.Tensor], n_new: int = 1,
                splits: int = 1, scratch: Optional[torch.Tensor] = None,
                out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """q [B, nh, hd] (already rotated); caches [B, nkv, T, hd] → [B, nh·hd]."""
    B, nh, hd = q.shape
    nkv, T = k_cache.shape[1], k_cache.shape[2]
    scale = 1.0 / math.sqrt(hd)
    if not q.is_cuda:
        length = (int(cache_len.item()) if cache_len is not None else 0) + n_new
        k = k_cache[:, :, :length].float().repeat_interleave(nh // nkv, dim=1)
        v = v_cache[:, :, :length].float().repeat_interleave(nh // nkv, dim=1)
        s = torch.einsum("bhd,bhtd->bht", q.float(), k) * scale
        if pad_left is not None:
            s = s.masked_fill(torch.arange(length)[None, None, :] < pad_left.long()[:, None, None], float("-inf"))
        return torch.einsum("bht,bhtd->bhd", torch.softmax(s, -1), v).to(q.dtype).reshape(B, nh * hd)
    if q.dtype != torch.bfloat16 or q.stride(2) != 1 or q.stride(1) != hd:
        raise ValueError("q must be bf16 [B, nh, hd] with dense heads (batch pitch may be larger)")
    if out is None:
        out = torch.empty((B, nh * hd), dtype=q.dtype, device=q.device)
    _lib.check(_lib.lib().jl_decode_attn(q.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr(), _p(cache_len),
                                         _p(pad_left), _p(scratch), out.data_ptr(), B, nh, nkv, hd, T, n_new,
                                         scale, splits, q.stride(0) if B > 1 else 0, _stream()), "jl_decode_attn")
    return out


def decode_attn_fused_plan(B: int, nh: int, nkv: int, hd: int, max_len: int, device: torch.device
                           ) -> Tuple[int, torch.Tensor]:
    """(splits, zeroed scratch) for `decode_attn_fused` at this shape."""
    L = _lib.lib()
    splits = int(L.jl_decode_attn_fused_splits(B, nkv, max_len, num_sms(device)))
    nbytes = int(L.jl_decode_attn_fused_scratch_bytes(B, nh, nkv, hd, splits))
    return splits, torch.zeros((nbytes + 3) // 4, dtype=torch.int32, device=device)


def set_decode_attn_variant(variant: int) -> None:
    """Pins the kernel behind `decode_attn_fused`: 0 = by shape (default), 1 = SIMT, 2 = tensor-core (head_dim 128).
    Process-wide; for tests and A/B timing."""
    _lib.check(_lib.lib().jl_decode_attn_fused_set_variant(int(variant)), "jl_decode_attn_fused_set_variant")


def decode_attn_fused(qkv: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, cos: torch.Tensor,
                      sin: torch.Tensor, cache_len: torch.Tensor, pad_left: Optional[torch.Tensor], nh: int,
                      nkv: int, hd: int, splits: int = 1, scratch: Optional[torch.Tensor] = None,
                      out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """RoPE + KV append + GQA decode attention + split merge in one launch.  qkv [B, (nh+2nkv)·hd] is the
    raw fused projection of the new token; the caches are updated in place at slot `cache_len`."""
    B = qkv.shape[0]
    if not qkv.is_cuda:
        work = qkv.clone()
        rope_kv_write(work, k_cache, v_cache, cos, sin, B, 1, nh, nkv, hd, cache_len=cache_len, pad_left=pad_left)
        return decode_attn(work.view(B, nh + 2 * nkv, hd)[:, :nh], k_cache, v_cache, cache_len, pad_left, 1)
    _bf16c(qkv, "qkv"); _bf16c(k_cache, "k_cache"); _bf16c(v_cache, "v_cache")
    if scratch is None:
        splits, scratch = 1, torch.zeros(max(B * nkv, 1), dtype=torch.int32, device=qkv.device)
    if out is None:
        out = torch.empty((B, nh * hd), dtype=qkv.dtype, device=qkv.device)
    _lib.check(_lib.lib().jl_decode_attn_fused(qkv.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr(),
                                               cos.data_ptr(), sin.data_ptr(), cache_len.data_ptr(), _p(pad_left),
                                               scratch.data_ptr(), out.data_ptr(), B, nh, nkv, hd,
                                               k_cache.shape[2], cos.shape[0], 1.0 / math.sqrt(hd), splits,
                                               _stream()), "jl_decode_attn_fused")
    return out


# --------------------------------------------------------------------------- prefill attention
def prefill_attn_ref(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, start: int, S: int,
                     pad_left: Optional[torch.Tensor]) -> torch.Tensor:
    """Dense causal attention of S new tokens (slots start..start+S) against the cache; q [B, S, nh, hd]."""
    B, _, nh, hd = q.shape
    nkv = k_cache.shape[1]
    length = start + S
    k = k_cache[:, :, :length].float().repeat_interleave(nh // nkv, dim=1)
    v = v_cache[:, :, :length].float().repeat_interleave(nh // nkv, dim=1)
    s = torch.einsum("bqhd,bhkd->bhqk", q.float(), k) / math.sqrt(hd)
    qpos = torch.arange(start, length, device=q.device)[:, None]
    kpos = torch.arange(length, device=q.device)[None, :]
    ok = (kpos <= qpos)[None, None]
    if pad_left is not None:
        ok = ok & (kpos[None, None] >= pad_left.long()[:, None, None, None])
    s = s.masked_fill(~ok, torch.finfo(torch.float32).min)
    return torch.einsum("bhqk,bhkd->bqhd", torch.softmax(s, -1), v).to(q.dtype).reshape(B, S, nh * hd)


def prefill_attn(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, start: int, S: int,
                 pad_left: Optional[torch.Tensor], cache_len: Optional[torch.Tensor] = None,
                 out: Optional[torch.Tensor] = None, force: Optional[str] = None) -> torch.Tensor:
    """Causal flash attention of S new tokens per row against the (already appended) cache.

    q [B, S, nh, hd] rotated; caches [B, nkv, T, hd]; keys [pad_left[b], start+S) are visible, causally.
    head_dim 128 → tcgen05 kernel, 64 → warp-level tensor-core kernel; force="mma" pins the latter (64 / 128).
    """
    if not q.is_cuda:
        return prefill_attn_ref(q, k_cache, v_cache, start, S, pad_left)
    if not _lib.has("jl_flash_prefill"):
        raise _lib.NativeLibraryError("jl_flash_prefill is not built into libjllama_b200.so")
    _bf16c(q, "q")
    B, _, nh, hd = q.shape
    nkv, T = k_cache.shape[1], k_cache.shape[2]
    if out is None:
        out = torch.empty((B, S, nh * hd), dtype=q.dtype, device=q.device)
    if force == "mma":
        _lib.check(_lib.lib().jl_prefill_attn_mma(q.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr(), _p(pad_left),
                                                  out.data_ptr(), B, S, nh, nkv, hd, T, start, 1.0 / math.sqrt(hd),
                                                  _stream()), "jl_prefill_attn_mma")
        return out
    _lib.check(_lib.lib().jl_flash_prefill(q.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr(), _p(cache_len),
                                           _p(pad_left), out.data_ptr(), B, S, nh, nkv, hd, T, start,
                                           1.0 / math.sqrt(hd), _stream()), "jl_flash_prefill")
    return out


# --------------------------------------------------------------------------- token selection
def argmax(logits: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    if not logits.is_cuda:
        return logits.float().argmax(-1).to(torch.int32)
    B, V = logits.shape
    if out is None:
        out = torch.empty(B, dtype=torch.int32, device=logits.device)
    _lib.check(_lib.lib().jl_argmax(logits.data_ptr(), B, V, logits.stride(0), 0, 0, out.data_ptr(), _stream()),
               "jl_argmax")
    return out


def topk_candidates(logits: torch.Tensor, k: int, idx_offset: int = 0,
                    out: Optional[Tuple[torch.Tensor, torch.Tensor]] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    B, V = logits.shape
    if not logits.is_cuda:
        vals, idx = torch.topk(logits.float(), min(k, V), dim=-1)
        return vals, (idx + idx_offset).to(torch.int32)
    if out is None:
        out = (torch.empty((B, k), dtype=torch.float32, device=logits.device),
               torch.empty((B, k), dtype=torch.int32, device=logits.device))
    _lib.check(_lib.lib().jl_topk(logits.data_ptr(), B, V, logits.stride(0), idx_offset, k, out[0].data_ptr(),
                                  out[1].data_ptr(), _stream()), "jl_topk")
    return out


def sample_candidates(vals: torch.Tensor, idx: torch.Tensor, *, n_lists: int = 1, list_stride: int = 0,
                      top_k: int = 0, top_p: float = 1.0, temperature: float = 1.0, seed: int = 0,
                      step: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """vals/idx: [n_lists, B, k] (or [B, k]).  CUDA only (the CPU path samples with ops.sampling)."""
    B, k = vals.shape[-2], vals.shape[-1]
    if out is None:
        out = torch.empty(B, dtype=torch.int32, device=vals.device)
    if list_stride == 0:
        list_stride = B * k
    _lib.check(_lib.lib().jl_sample(vals.data_ptr(), idx.data_ptr(), B, n_lists, list_stride, k, int(top_k or 0),
                                    float(top_p), float(temperature), int(seed), _p(step), out.data_ptr(),
                                    _stream()), "jl_sample")
    return out


# --------------------------------------------------------------------------- exact full-vocabulary selection
def sample_full(logits: torch.Tensor, *, temperature: float, top_k: int = 0, top_p: float = 1.0, seed: int = 0,
                step: Optional[torch.Tensor] = None, rows: Optional[int] = None) -> torch.Tensor:
    """Tokens [B] from fp32 logits [B, V] (or [V] read by `rows` independent draws): temperature → top-k → top-p →
    categorical over the whole vocabulary, exact and deterministic (csrc/sample_full.cu).  temperature 0 = arg-max."""
    if not logits.is_cuda or logits.dtype != torch.float32:
        raise ValueError("sample_full needs CUDA fp32 logits")
    if logits.dim() == 1:
        B, V, ld = int(rows or 1), logits.shape[0], 0
    else:
        B, V, ld = logits.shape[0], logits.shape[1], logits.stride(0)
    tok = torch.empty(B, dtype=torch.int32, device=logits.device)
    ticket = torch.zeros(1, dtype=torch.int32, device=logits.device)
    _lib.check(_lib.lib().jl_sample_full(
        logits.data_ptr(), B, V, ld, float(temperature), int(top_k or 0), float(top_p), int(seed), _p(step), tok.data_ptr(),
        None, None, 1, 0, 0, 0, ticket.data_ptr(), None, None, None, 0, 0, 0, None, None, None, None, _stream()),
        "jl_sample_full")
    return tok
