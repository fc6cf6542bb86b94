"""The same-box "NCCL + cuBLAS" yardstick: the model run with library ops only.

BASELINE.md §2: the reference publishes no numbers and cannot run here, so the bar for every hand-written path
is what XLA would have emitted for it — cuBLAS GEMMs (`F.linear`), a library attention (`SDPA`), NCCL
all-reduce after the row-parallel `wo` / `w2` and an all-gather of the vocab-sharded logits — CUDA-graphed so
that launch overhead is not what is being compared.  Like the reference (model.py:194-198) the decode step
attends over the whole static cache under a mask.  This class shares the engine's weights (no copies); it is
a measuring stick, not a product path.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist
import torch.nn.functional as F

from .engine import B200Engine


class TorchBaseline:
    def __init__(self, engine: B200Engine, batch: int, max_len: int):
        self.e = engine
        c, dev = engine.config, engine.device
        self.B, self.T = batch, max_len
        L = len(engine.layers)
        shape = (L, batch, engine.nkv, max_len, engine.hd)
        self.k = torch.zeros(shape, dtype=torch.bfloat16, device=dev)
        self.v = torch.zeros(shape, dtype=torch.bfloat16, device=dev)
        self.len = torch.zeros(1, dtype=torch.long, device=dev)
        self.ids = torch.zeros(batch, dtype=torch.long, device=dev)
        self.arange = torch.arange(max_len, device=dev)
        self.freqs = torch.polar(torch.ones_like(engine.cos), torch.atan2(engine.sin, engine.cos))   # complex64
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self.eps = c.rms_norm_eps

    def _norm(self, x: torch.Tensor, w: torch.Tensor) -> torch.Tensor:
        xf = x.float()
        return (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + self.eps)).to(x.dtype) * w

    def _rope(self, t: torch.Tensor, f: torch.Tensor) -> torch.Tensor:          # t [B, S, h, hd]
        tc = torch.view_as_complex(t.float().reshape(*t.shape[:-1], -1, 2))
        return torch.view_as_real(tc * f[None, :, None, :]).flatten(-2).to(t.dtype)

    def _all_reduce(self, y: torch.Tensor) -> torch.Tensor:
        if self.e.tp > 1:
            dist.all_reduce(y)
        return y

    def forward(self, ids: torch.Tensor, start: Optional[int]) -> torch.Tensor:
        """ids [B, S].  start=None: decode (S = 1) at the device-side position `self.len`; else prefill at `start`."""
        e = self.e
        B, S = ids.shape
        nh, nkv, hd = e.nh, e.nkv, e.hd
        x = F.embedding(ids, e.embedding)
        if start is None:
            pos = self.len                                              # [1] device tensor
            f = self.freqs[pos]
            mask = (self.arange <= pos)[None, None, None, :]            # attend over the whole cache under a mask
        else:
            f = self.freqs[start:start + S]
        for i, l in enumerate(e.layers):
            h = self._norm(x, l.attn_norm)
            qkv = F.linear(h, l.wqkv).view(B, S, nh + 2 * nkv, hd)
            q, k = self._rope(qkv[:, :, :nh], f), self._rope(qkv[:, :, nh:nh + nkv], f)
            v = qkv[:, :, nh + nkv:]
            if start is None:
                self.k[i].index_copy_(2, pos, k.transpose(1, 2))
                self.v[i].index_copy_(2, pos, v.transpose(1, 2))
                a = F.scaled_dot_product_attention(q.transpose(1, 2), self.k[i], self.v[i], attn_mask=mask, enable_gqa=True)
            else:
                self.k[i][:, :, start:start + S] = k.transpose(1, 2)
                self.v[i][:, :, start:start + S] = v.transpose(1, 2)
                a = F.scaled_dot_product_attention(q.transpose(1, 2), self.k[i][:, :, :start + S],
                                                   self.v[i][:, :, :start + S], is_causal=(start == 0), enable_gqa=True)
            a = a.transpose(1, 2).reshape(B, S, nh * hd)
            x = x + self._all_reduce(F.linear(a, l.wo))
            h = self._norm(x, l.ffn_norm)
            gu = F.linear(h, l.w13)
            x = x + self._all_reduce(F.linear(F.silu(gu[..., 0::2]) * gu[..., 1::2], l.w2))
        x = self._norm(x[:, -1], e.final_norm)
        logits = F.linear(x, e.lm_head).float()
        if e.tp > 1:
            parts = [torch.empty_like(logits) for _ in range(e.tp)]
            dist.all_gather(parts, logits)
            logits = torch.cat(parts, -1)
        return logits

    @torch.no_grad()
    def prefill(self, ids: torch.Tensor) -> torch.Tensor:
        logits = self.forward(ids, 0)
        self.len.fill_(ids.shape[1])
        self.ids.copy_(logits.argmax(-1))
        return logits

    def _step(self) -> None:
        logits = self.forward(self.ids[:, None], None)
        self.ids.copy_(logits.argmax(-1))
        self.len.add_(1)

    @torch.no_grad()
    def decode_step(self) -> None:
        """One greedy decode step for every row, replayed from a CUDA graph after the first call."""
        if self.graph is None:
            snap = (self.len.clone(), self.ids.clone())
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                for _ in range(2):
                    self._step()
            torch.cuda.current_stream().wait_stream(side)
            self.len.copy_(snap[0]); self.ids.copy_(snap[1])
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph):
                self._step()
            self.len.copy_(snap[0]); self.ids.copy_(snap[1])
        self.graph.replay()
