"""B200 execution engine: kernel-ready weights, device-resident KV cache, CUDA-graph decode.

What XLA/GSPMD did implicitly for the reference (fused elementwise kernels, cuBLAS GEMMs, a compiled
`lax.while_loop` decode body — SURVEY §3.2/§3.3) is explicit here:

* weights are repacked once into the layouts the kernels want: fused ``wqkv`` ``[(nh+2nkv)·hd, H]``,
  ``w13`` with (w1, w3) rows interleaved so SwiGLU happens in the GEMM epilogue, everything ``[N, K]``;
* a decode step is a fixed sequence of ~7 launches per layer (RMSNorm fused into the GEMV prologue,
  residual / SwiGLU fused into epilogues, RoPE fused with the KV append, split-KV GQA attention, on-device
  sampling and sequence bookkeeping) captured once into a CUDA graph and replayed per token with no
  host synchronisation; kernels are chained with programmatic dependent launch;
* the step's moving parts (``cache_len``, ``cur_len``, RNG step, finished flags) live in device memory so
  the same graph serves every position.

Model math: /root/reference/jax_llama/model.py:201-300 (attention), :337-340 (MLP), :374-398 (block),
:622-741 (backbone + head); decode loop semantics: HF Flax `_sample` as configured by generation.py:28-41.
"""
from __future__ import annotations

import logging
import os
from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional, Tuple

import torch

from ..config import LLaMAConfig
from ..ops import _lib
from ..ops import kernels as K
from .reference import KVCache, ModelOutput, rope_cos_sin

logger = logging.getLogger(__name__)

Tree = Dict[str, Any]


@dataclass
class DeviceKVCache(KVCache):
    """KV cache whose fill level also lives on the device (read by kernels, advanced by kernels)."""

    len_dev: Optional[torch.Tensor] = None      # int32 [1]

    def set_index(self, value: int) -> None:
        self.index = int(value)
        self.len_dev.fill_(int(value))


@dataclass
class LayerWeights:
    attn_norm: torch.Tensor
    wqkv: torch.Tensor
    wo: torch.Tensor
    ffn_norm: torch.Tensor
    w13: torch.Tensor
    w2: torch.Tensor


def interleave_rows(w1: torch.Tensor, w3: torch.Tensor) -> torch.Tensor:
    """[F, H], [F, H] → [2F, H] with row 2j = w1[j], row 2j+1 = w3[j]."""
    f, h = w1.shape
    return torch.stack([w1, w3], dim=1).reshape(2 * f, h).contiguous()


@dataclass
class DecodeState:
    """Static buffers of one captured decode configuration."""

    B: int
    max_len: int
    ids: torch.Tensor
    x: torch.Tensor
    qkv: torch.Tensor
    attn: torch.Tensor
    mlp: torch.Tensor
    logits: torch.Tensor
    tok: torch.Tensor
    finished: torch.Tensor
    sequences: torch.Tensor
    cur_len: torch.Tensor
    step: torch.Tensor
    n_unfinished: torch.Tensor
    splits: int
    scratch: Optional[torch.Tensor]
    graphs: Dict[Tuple, torch.cuda.CUDAGraph] = field(default_factory=dict)


class B200Engine:
    def __init__(self, config: LLaMAConfig, embedding: torch.Tensor, layers: List[LayerWeights],
                 final_norm: torch.Tensor, lm_head: torch.Tensor, *, use_cuda_graph: bool = True,
                 use_megakernel: bool = True,
                 tp: int = 1, rank: int = 0, comm: Any = None):
        """`layers` / `lm_head` hold this rank's tensor-parallel slice (Megatron layout of
        parallel/layout.py: wqkv/w13 cut on output rows, wo/w2 on the reduction dim, lm_head on vocab);
        the embedding table and the norm scales are replicated."""
        self.config = config
        self.tp, self.rank = int(tp), int(rank)
        if self.tp > 1:
            from ..parallel.layout import check_tp_divisibility
            check_tp_divisibility(config, self.tp)
            if comm is None:
                from ..parallel.comm import PeerComm
                comm = PeerComm(self.tp, self.rank, embedding.device)
        self.comm = comm
        self.device = embedding.device
        if self.device.type != "cuda":
            raise ValueError("B200Engine needs CUDA tensors")
        _lib.lib()                                      # fail here, loudly, if the native library is absent
        self.embedding = embedding
        self.layers = layers
        self.final_norm = final_norm
        self.lm_head = lm_head
        self.use_cuda_graph = use_cuda_graph
        self.nh = config.num_attention_heads // self.tp          # local (per-rank) head counts
        self.nkv = config.num_key_value_heads // self.tp
        self.hd = config.head_dim
        self.v_local = lm_head.shape[0]
        cos, sin = rope_cos_sin(self.hd, 2 * config.max_sequence_length, config.rope_theta)
        self.cos, self.sin = cos.to(self.device), sin.to(self.device)
        self._decode_states: Dict[Tuple[int, int], DecodeState] = {}
        self.warned_slow_prefill = False
        self._mega_fit_cache: Dict[int, bool] = {}
        # step counter of the whole-step decode kernel: the base of its packet tags (never reset, see decode_mega.cu)
        self._mega_gen = torch.zeros(1, dtype=torch.int32, device=self.device)
        self._mega_peer: Optional[Tuple] = None
        self.mega_write_logits = False      # True: the decode kernel also stores the fp32 [B, V] logits (tests, forward-style use)
        if self.device.type == "cuda":
            K.ensure_gemm_workspace(self.device)          # split-K scratch of the small-batch GEMMs (before any capture)
        # Prefill at tp > 1 runs over NVLS (parallel/comm.py `NvlsPrefill`): sequence-parallel when the token count is a
        # multiple of 128·tp, otherwise tcgen05 GEMM + in-switch all-reduce kernel.  JLLAMA_SP=0 forces the latter.
        self.sequence_parallel_prefill = os.environ.get("JLLAMA_SP", "1") != "0"
        self.use_megakernel = use_megakernel and self.hd == 128 \
            and config.hidden_size % 256 == 0 and (config.intermediate_size // self.tp) % 256 == 0 \
            and (self.nh * self.hd) % 256 == 0 and (self.nh // self.nkv) in (1, 2, 4, 8)
        self._stack_weights()
        # Largest batch the whole-step kernel takes (it can hold 8 rows).  From batch 5 up the per-op path (tcgen05
        # swap-AB GEMMs + tensor-core decode attention, under TP with the fused exchange / all-reduce / norm kernel) is
        # faster whenever a rank streams a lot of weights per step; the whole-step kernel keeps all 8 rows only where
        # the step is latency-bound (little weight per rank).  Measured, ms/step whole-step vs per-op, 2k context:
        # 8B on one GPU (15 GB) batch 6 5.13 vs 4.32, batch 8 5.23 vs 4.41; 7B batch 8 7.12 vs 5.20; 8B at TP=2 (7.5 GB)
        # batch 6 3.50 vs 3.22, batch 8 3.69 vs 3.31 (batch 5: 3.19 vs 3.36); 70B at TP=8 (17.6 GB) batch 8 9.53 vs
        # 8.23 for batch 16; 8B at TP=8 (1.9 GB) batch 8 1.48 vs 2.17 for batch 16.  JLLAMA_MEGA_MAX_BATCH overrides.
        default_max = 4 if self.weight_bytes_per_step() >= (5 << 30) else 8
        self.mega_max_batch = min(8, int(os.environ.get("JLLAMA_MEGA_MAX_BATCH", str(default_max))))

    def _mega_fits(self, batch: int) -> bool:
        """Whole-step kernel usable for this batch: its staged activations (batch x widest GEMV input) must
        leave room for a weight ring of at least 4 stages in shared memory (8B model: batch <= 4 at TP=1)."""
        if not self.use_megakernel or batch > self.mega_max_batch:
            return False
        if batch not in self._mega_fit_cache:
            c = self.config
            self._mega_fit_cache[batch] = int(_lib.lib().jl_decode_mega_stages(
                batch, c.hidden_size, c.intermediate_size // self.tp, self.nh, self.nkv, self.hd)) >= 4
        return self._mega_fit_cache[batch]

    def _mega_tp_args(self) -> Tuple:
        """(tp, rank, peer table of the partial-sum packets, its multicast address, peer table of the arg-max
        candidates, its multicast address).  The two buffers live in a symmetric heap (parallel/symm.py) sized for
        the largest batch the kernel takes; allocated once, before any graph capture."""
        if self.tp == 1:
            return (1, 0, None, None, None, None)
        if self._mega_peer is None:
            import ctypes as C
            from ..parallel.symm import SymmetricHeap
            L = _lib.lib()
            cand_off = C.c_size_t(0)
            nbytes = int(L.jl_decode_mega_peer_bytes(8, self.config.hidden_size, self.tp, K.num_sms(self.device),
                                                     C.byref(cand_off)))
            logits_off = (nbytes + 1023) // 1024 * 1024          # [8][V] fp32: every rank's shard of the logits, gathered
            nbytes = logits_off + 8 * self.config.vocab_size * 4
            heap = SymmetricHeap(self.tp, self.rank, self.device, nbytes, group=getattr(self.comm, "group", None))
            self._mega_peer = (heap, heap.peer_table(0), heap.mc(0), heap.peer_table(cand_off.value),
                               heap.mc(cand_off.value), heap.local_ptr + cand_off.value)
            self._logits_full = heap.view(logits_off, (8, self.config.vocab_size), torch.float32)
            self._logits_full_mc = heap.mc(logits_off)
            self._bcast_ticket = torch.zeros(1, dtype=torch.int32, device=self.device)
            self._bcast_flag_off = self.comm.flag_block(("logits_bcast",), 8)
        _, p_tab, p_mc, c_tab, c_mc, _ = self._mega_peer
        return (self.tp, self.rank, p_tab, p_mc or None, c_tab, c_mc or None)

    def _stack_weights(self) -> None:
        """Store each op's weights of all layers contiguously ([L·N, K]); the per-layer tensors become views.
        One TMA tensor map per op then describes the whole model (the persistent decode kernel streams
        layer after layer through the same map)."""
        def stack(name: str) -> torch.Tensor:
            parts = [getattr(l, name) for l in self.layers]
            big = torch.empty((len(parts) * parts[0].shape[0],) + tuple(parts[0].shape[1:]),
                              dtype=parts[0].dtype, device=parts[0].device)
            n = parts[0].shape[0]
            for i, (l, t) in enumerate(zip(self.layers, parts)):
                big[i * n:(i + 1) * n].copy_(t)
                setattr(l, name, big[i * n:(i + 1) * n])       # drop the stand-alone copy
            return big

        self.wqkv_all, self.wo_all = stack("wqkv"), stack("wo")
        self.w13_all, self.w2_all = stack("w13"), stack("w2")
        self.attn_norm_all, self.ffn_norm_all = stack("attn_norm"), stack("ffn_norm")

    # ------------------------------------------------------------------ construction
    @classmethod
    def from_tree(cls, config: LLaMAConfig, params: Tree, **kw: Any) -> "B200Engine":
        """Repack a canonical (``[in, out]``) tree that already lives on a CUDA device."""
        t = params["transformer"]
        bf = lambda x: x.to(torch.bfloat16)          # noqa: E731
        layers = []
        for i in range(config.num_hidden_layers):
            l = t["h"][str(i)]
            a, f = l["attention"], l["feed_forward"]
            wqkv = torch.cat([bf(a["wq"]["kernel"]).t(), bf(a["wk"]["kernel"]).t(), bf(a["wv"]["kernel"]).t()], 0)
            layers.append(LayerWeights(
                attn_norm=bf(l["attention_norm"]["kernel"]).contiguous(), wqkv=wqkv.contiguous(),
                wo=bf(a["wo"]["kernel"]).t().contiguous(), ffn_norm=bf(l["ffn_norm"]["kernel"]).contiguous(),
                w13=interleave_rows(bf(f["w1"]["kernel"]).t(), bf(f["w3"]["kernel"]).t()),
                w2=bf(f["w2"]["kernel"]).t().contiguous()))
        emb = bf(t["wte"]["embedding"]).contiguous()
        tp = int(kw.get("tp", 1))
        head = emb if config.tie_word_embeddings else bf(params["lm_head"]["kernel"]).t().contiguous()
        if tp > 1 and emb.shape[0] != config.vocab_size:
            # the layout rules shard the table on vocab; the engine keeps it whole on every rank
            import torch.distributed as dist
            parts = [torch.empty_like(emb) for _ in range(tp)]
            dist.all_gather(parts, emb)
            emb = torch.cat(parts, 0)
        return cls(config, emb, layers, bf(t["ln_f"]["kernel"]).contiguous(), head, **kw)

    @classmethod
    def random_init(cls, config: LLaMAConfig, device: Any = "cuda", seed: int = 0, **kw: Any) -> "B200Engine":
        """Random weights generated directly in kernel layout on the device (benchmarks; no host copy).

        Every rank draws the FULL matrices from the same generator and keeps its Megatron slice (rows of wqkv /
        w13 / lm_head, columns of wo / w2 — parallel/layout.py), so a tensor-parallel engine and a single-GPU
        engine built with the same seed are the same model: `bench.py` uses that for its pre-run `tp_check`."""
        dev = torch.device(device)
        tp, rank = int(kw.get("tp", 1)), int(kw.get("rank", 0))
        g = torch.Generator(device=dev).manual_seed(seed * 1000)
        std = config.initializer_range
        h, F, V = config.hidden_size, config.intermediate_size, config.vocab_size
        NH, NKV, hd = config.num_attention_heads, config.num_key_value_heads, config.head_dim
        f, v, nh, nkv = F // tp, V // tp, NH // tp, NKV // tp

        def full(n: int, k: int) -> torch.Tensor:
            return torch.randn(n, k, generator=g, device=dev, dtype=torch.float32) * std

        def rows(t: torch.Tensor, n: int) -> torch.Tensor:
            return t[rank * n:(rank + 1) * n].to(torch.bfloat16).contiguous()

        def cols(t: torch.Tensor, n: int) -> torch.Tensor:
            return t[:, rank * n:(rank + 1) * n].to(torch.bfloat16).contiguous()

        ones = lambda: torch.ones(h, dtype=torch.bfloat16, device=dev)     # noqa: E731
        layers = []
        for _ in range(config.num_hidden_layers):
            wqkv = full((NH + 2 * NKV) * hd, h)
            q, k_, v_ = wqkv[:NH * hd], wqkv[NH * hd:(NH + NKV) * hd], wqkv[(NH + NKV) * hd:]
            wqkv_r = torch.cat([rows(q, nh * hd), rows(k_, nkv * hd), rows(v_, nkv * hd)], 0)
            del wqkv, q, k_, v_
            wo = cols(full(h, NH * hd), nh * hd)
            w13 = rows(full(2 * F, h), 2 * f)                  # (w1_j, w3_j) rows interleaved: a slice of 2f rows is f pairs
            w2 = cols(full(h, F), f)
            layers.append(LayerWeights(ones(), wqkv_r, wo, ones(), w13, w2))
        ge = torch.Generator(device=dev).manual_seed(seed * 1000 + 999)      # same table on every rank
        emb = (torch.randn(V, h, generator=ge, device=dev, dtype=torch.float32) * std).to(torch.bfloat16)
        head = emb if (config.tie_word_embeddings and tp == 1) else rows(full(V, h), v)
        return cls(config, emb, layers, ones(), head, **kw)

    def to_tree(self) -> Tree:
        """Inverse of `from_tree` (canonical names / ``[in, out]`` layout) for export and tests."""
        c = self.config
        q, kv = self.nh * self.hd, self.nkv * self.hd
        layers = {}
        for i, l in enumerate(self.layers):
            layers[str(i)] = {
                "attention": {"wq": {"kernel": l.wqkv[:q].t().contiguous()},
                              "wk": {"kernel": l.wqkv[q:q + kv].t().contiguous()},
                              "wv": {"kernel": l.wqkv[q + kv:].t().contiguous()},
                              "wo": {"kernel": l.wo.t().contiguous()}},
                "feed_forward": {"w1": {"kernel": l.w13[0::2].t().contiguous()},
                                 "w3": {"kernel": l.w13[1::2].t().contiguous()},
                                 "w2": {"kernel": l.w2.t().contiguous()}},
                "attention_norm": {"kernel": l.attn_norm}, "ffn_norm": {"kernel": l.ffn_norm}}
        tree: Tree = {"transformer": {"wte": {"embedding": self.embedding}, "ln_f": {"kernel": self.final_norm},
                                      "h": layers}}
        if not c.tie_word_embeddings:
            tree["lm_head"] = {"kernel": self.lm_head.t().contiguous()}
        return tree

    # ------------------------------------------------------------------ cache
    def new_cache(self, batch: int, max_length: int) -> DeviceKVCache:
        shape = (self.config.num_hidden_layers, batch, self.nkv, max_length, self.hd)
        return DeviceKVCache(
            k=torch.zeros(shape, dtype=torch.bfloat16, device=self.device),
            v=torch.zeros(shape, dtype=torch.bfloat16, device=self.device), index=0,
            pad_left=torch.zeros(batch, dtype=torch.int32, device=self.device),
            len_dev=torch.zeros(1, dtype=torch.int32, device=self.device))

    @staticmethod
    def pad_left_from_mask(attention_mask: Optional[torch.Tensor], batch: int, seq: int,
                           device: torch.device) -> torch.Tensor:
        """Left-padding count per row.  The fast path supports exactly the masks the generation front-end
        produces (zeros then ones, generation.py:55-60); anything else must use the reference backend."""
        if attention_mask is None:
            return torch.zeros(batch, dtype=torch.int32, device=device)
        m = attention_mask[:, :seq].to(torch.int32)
        pad = (m == 0).to(torch.int32).sum(-1)
        expect = (torch.arange(seq, device=m.device)[None, :] >= pad[:, None]).to(torch.int32)
        if not torch.equal(m, expect):
            raise NotImplementedError("the B200 engine supports left-padding masks only; use the reference "
                                      "backend (CPU tensors or output_attentions=True) for arbitrary masks")
        return pad.to(torch.int32).to(device)

    # ------------------------------------------------------------------ layer stack (eager; any S)
    def _layers(self, x: torch.Tensor, B: int, S: int, cache: DeviceKVCache,
                positions: Optional[torch.Tensor], ws: Optional[DecodeState] = None) -> torch.Tensor:
        """x [B·S, H] → x after all blocks.  Appends S slots at cache.len_dev (does not advance it)."""
        c = self.config
        eps = c.rms_norm_eps
        nh, nkv, hd = self.nh, self.nkv, self.hd
        start = cache.index
        # (see below) only the graph-replayed batched decode of a TP engine takes the fused all-reduce + norm form
        fuse_norm = (self.tp > 1 and S == 1 and positions is None and ws is not None and 4 < B <= 128
                     and os.environ.get("JLLAMA_TP_FUSE_NORM", "1") != "0")
        xn = None
        for i, l in enumerate(self.layers):
            kc, vc = cache.k[i], cache.v[i]
            if xn is None:
                qkv = K.linear(x, l.wqkv, norm_w=l.attn_norm, eps=eps, out=ws.qkv if ws else None)
            else:
                qkv = K.linear(xn, l.wqkv, out=ws.qkv if ws else None)      # normalised by the previous layer's all-reduce
            if S == 1 and positions is None:
                # decode: RoPE + KV append + split-KV attention + merge in one launch
                attn = K.decode_attn_fused(qkv, kc, vc, self.cos, self.sin, cache.len_dev, cache.pad_left, nh, nkv,
                                           hd, ws.splits if ws else 1, ws.scratch if ws else None,
                                           out=ws.attn if ws else None)
            elif S == 1:
                # explicit position_ids: rotate with the caller's positions, then the unfused attention
                K.rope_kv_write(qkv, kc, vc, self.cos, self.sin, B, S, nh, nkv, hd, cache_len=cache.len_dev,
                                pad_left=cache.pad_left, positions=positions)
                attn = K.decode_attn(qkv.view(B, nh + 2 * nkv, hd)[:, :nh], kc, vc, cache.len_dev, cache.pad_left, 1)
            else:
                # prefill: rotated q goes to its own dense [B·S, nh·hd] buffer for the flash kernel
                q_rot = torch.empty((B * S, nh * hd), dtype=qkv.dtype, device=qkv.device)
                K.rope_kv_write(qkv, kc, vc, self.cos, self.sin, B, S, nh, nkv, hd, cache_len=cache.len_dev,
                                pad_left=cache.pad_left, positions=positions, q_out=q_rot)
                attn = self._prefill_attn(q_rot.view(B, S, nh, hd), kc, vc, start, S, cache).view(B * S, nh * hd)
            if fuse_norm:
                # batched decode under TP: GEMM → one kernel that exchanges the "done" flags, pulls the in-switch sum,
                # adds the residual and applies the NEXT RMSNorm (7 launches per layer instead of 11)
                nv = self.comm.nvls_state(B, c.hidden_size)
                K.linear(attn, l.wo, out=nv.next_part())
                xn = nv.allreduce_add_norm(x, l.ffn_norm, eps)
                h = K.linear(xn, l.w13, mode=K.SWIGLU, out=ws.mlp if ws else None)
                K.linear(h, l.w2, out=nv.next_part())
                xn = nv.allreduce_add_norm(x, self.layers[i + 1].attn_norm if i + 1 < len(self.layers) else self.final_norm, eps)
                continue
            x = self._row_parallel(attn, l.wo, x)
            h = K.linear(x, l.w13, norm_w=l.ffn_norm, eps=eps, mode=K.SWIGLU, out=ws.mlp if ws else None)
            x = self._row_parallel(h, l.w2, x)
        if ws is not None:
            ws.xn_final = xn            # rows already normalised with final_norm (fused path), else None
        return x

    def _prefill_sp(self, input_ids: torch.Tensor, cache: DeviceKVCache, B: int, S: int) -> torch.Tensor:
        """Megatron-style sequence-parallel prefill (a capability the reference does not have).

        The residual stream lives sharded by token rows ([M/tp, H] per rank).  Row-parallel GEMMs (wo, w2) keep
        their partial product on the rank that computed it; the owner of a row pulls the sum over ranks with
        `multimem.ld_reduce` (reduced inside the NVSwitch), adds the residual, normalises and stores the row into
        every rank's copy of the gathered activations with `multimem.st` (`NvlsPrefill.rs_reduce`); column-parallel
        GEMMs (wqkv, w1·w3) read that local copy after a one-warp wait for the owners' flags (`NvlsPrefill.ag_gemm`).
        No NCCL call.

        A single prompt whose length allows it is cut into two halves that run the same chain on two streams, the
        second half one step behind the first: while one half's rows cross NVLink (reduce-scatter / all-gather
        kernels, a few CTAs of work) the other half's GEMMs keep the tensor cores busy.  The second half's attention
        reads the keys the first half appended to the same layer's cache (ordered with an event per layer).
        Returns this rank's vocabulary slice of the last-position logits, [B, V/tp] fp32."""
        c = self.config
        eps = c.rms_norm_eps
        nh, nkv, hd = self.nh, self.nkv, self.hd
        M = B * S
        halves = 2 if (B == 1 and M % (2 * 128 * self.tp) == 0 and os.environ.get("JLLAMA_SP_PIPELINE", "1") != "0") else 1
        Mc = M // halves
        start = cache.index
        ids = input_ids.reshape(-1)
        main = torch.cuda.current_stream()
        if halves == 2 and getattr(self, "_sp_side", None) is None:
            self._sp_side = torch.cuda.Stream()
        streams = [main] if halves == 1 else [main, self._sp_side]
        states = [self.comm.nvls_state(Mc, c.hidden_size, tag=h) for h in range(halves)]
        if halves == 2:
            self._sp_side.wait_stream(main)
        Bc, Sc = (B, S) if halves == 1 else (1, Mc)
        last_rows = (torch.arange(B, device=self.device) * S + (S - 1)) if halves == 1 else \
            torch.arange(Mc - 1, Mc, device=self.device)                         # device-side (graph capturable)
        xs, lens = [], []
        for h, (sp, stream) in enumerate(zip(states, streams)):
            with torch.cuda.stream(stream):
                r, mloc = self.rank, sp.mloc
                rows = ids[h * Mc + r * mloc:h * Mc + (r + 1) * mloc]
                xs.append(K.embed(rows, self.embedding))                          # this rank's rows of the stream
                lens.append(cache.len_dev + h * Mc if h else cache.len_dev)       # first cache slot of this half
                sp.rs_reduce(xs[h], self.layers[0].attn_norm, eps, has_partials=False)    # chain start: norm + broadcast
        for i, l in enumerate(self.layers):
            kc, vc = cache.k[i], cache.v[i]
            nxt = self.layers[i + 1].attn_norm if i + 1 < len(self.layers) else self.final_norm
            appended = None
            for h, (sp, stream) in enumerate(zip(states, streams)):
                with torch.cuda.stream(stream):
                    qkv = sp.ag_gemm(l.wqkv, K.PLAIN)                             # [Mc, (nh+2nkv)·hd] local heads
                    q_rot = torch.empty((Mc, nh * hd), dtype=qkv.dtype, device=qkv.device)
                    K.rope_kv_write(qkv, kc, vc, self.cos, self.sin, Bc, Sc, nh, nkv, hd, cache_len=lens[h],
                                    pad_left=cache.pad_left, q_out=q_rot)
                    if halves == 2 and h == 0:
                        appended = torch.cuda.Event()
                        appended.record(stream)
                    elif halves == 2:
                        stream.wait_event(appended)                               # the first half's keys of this layer
                    attn = self._prefill_attn(q_rot.view(Bc, Sc, nh, hd), kc, vc, start + h * Mc, Sc, cache).view(Mc, nh * hd)
                    K.linear(attn, l.wo, out=sp.next_part())                      # partial product stays on this rank
                    sp.rs_reduce(xs[h], l.ffn_norm, eps, has_partials=True)       # in-switch reduce-scatter + norm + all-gather
                    hmid = sp.ag_gemm(l.w13, K.SWIGLU)
                    K.linear(hmid, l.w2, out=sp.next_part())
                    sp.rs_reduce(xs[h], nxt, eps, has_partials=True)
        # last position of every batch row, already final-normed: its owner stored it into every rank's copy
        with torch.cuda.stream(streams[-1]):
            xl = states[-1].read_rows(last_rows).contiguous()
        for sp, stream in zip(states, streams):
            with torch.cuda.stream(stream):
                sp.end_forward()                                                  # device epoch base += collectives issued
        if halves == 2:
            main.wait_stream(self._sp_side)
        return K.linear(xl, self.lm_head, out_f32=True)

    def _prefill_sp_graphed(self, input_ids: torch.Tensor, cache: DeviceKVCache, B: int, S: int) -> torch.Tensor:
        """`_prefill_sp` replayed from a CUDA graph.  At tp > 1 the prefill is a few hundred short kernels per forward
        (GEMMs of 1/tp the width, one-warp flag kernels, two streams): issued eagerly it is bound by the host's launch
        rate, not by the GPU.  One graph per (cache, batch, length, start slot); everything that changes between calls
        (token ids, cache fill level, left-padding counts, flag epochs) is read from device memory."""
        if not self.use_cuda_graph or os.environ.get("JLLAMA_SP_GRAPH", "1") == "0":
            return self._prefill_sp(input_ids, cache, B, S)
        graphs = self.__dict__.setdefault("_sp_graphs", {})
        # the graph bakes the addresses of the cache tensors: all of them are part of the key, so a later cache object
        # that happens to reuse one address can never pick up a graph that points at freed memory
        key = (cache.k.data_ptr(), cache.v.data_ptr(), cache.len_dev.data_ptr(), cache.pad_left.data_ptr(), B, S, cache.index)
        ent = graphs.get(key)
        if ent is None:
            while len(graphs) >= 4:                # each graph owns its activations: keep a handful
                graphs.pop(next(iter(graphs)))
            ids_static = input_ids.clone()
            for _ in range(2):                     # eager warm-up: symmetric buffers, streams, both flag parities
                self._prefill_sp(ids_static, cache, B, S)
            torch.cuda.synchronize(self.device)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                out_static = self._prefill_sp(ids_static, cache, B, S)
            ent = graphs[key] = (g, ids_static, out_static)
        g, ids_static, out_static = ent
        ids_static.copy_(input_ids)
        g.replay()
        return out_static.clone()

    def _row_parallel(self, a: torch.Tensor, w: torch.Tensor, x: torch.Tensor) -> torch.Tensor:
        """x += a · wᵀ summed over tensor-parallel ranks (the reference's GSPMD all-reduce after wo / w2)."""
        if self.tp == 1:
            return K.linear(a, w, residual=x, mode=K.RESIDUAL, out=x)
        m = a.numel() // w.shape[1]
        if m <= 8:
            # decode: GEMV + one-shot all-reduce over NVLink peer memory + residual, one kernel
            return self.comm.gemv_allreduce(a, w, x, out=x)
        # larger M: tcgen05 GEMM into this rank's copy of a symmetric buffer, then one kernel in which every rank
        # pulls the in-switch sum (multimem.ld_reduce) and adds it to its replicated residual stream
        st = self.comm.nvls_state(m, w.shape[0])
        K.linear(a, w, out=st.next_part())
        return st.allreduce_add(x)

    def prefill_paths(self) -> List[str]:
        return ["single"] if self.tp == 1 else ["sequence_parallel", "allreduce"]

    def set_prefill_path(self, name: Optional[str]) -> None:
        """`sequence_parallel` (default when the token count allows), `allreduce`, or None = the default."""
        self.sequence_parallel_prefill = name in (None, "sequence_parallel") and os.environ.get("JLLAMA_SP", "1") != "0"

    def _prefill_attn(self, q: torch.Tensor, kc: torch.Tensor, vc: torch.Tensor, start: int, S: int,
                      cache: DeviceKVCache) -> torch.Tensor:
        if self.hd not in (64, 128):
            # head_dim 128 (every LLaMA-1/2/3) runs the tcgen05 flash kernel, 64 (1B-class models) the warp-level
            # tensor-core kernel; a dense O(S^2) PyTorch statement on the GPU path would be a silent library fallback —
            # other head sizes go through the reference backend (CPU tensors)
            raise NotImplementedError(f"the B200 prefill kernels are built for head_dim 128 and 64, got {self.hd}: run "
                                      "this model through the reference backend (parameters on the CPU)")
        return K.prefill_attn(q, kc, vc, start, S, cache.pad_left, cache.len_dev)

    # ------------------------------------------------------------------ public forward
    @torch.no_grad()
    def forward(self, input_ids: torch.Tensor, attention_mask: Optional[torch.Tensor] = None,
                position_ids: Optional[torch.Tensor] = None, cache: Optional[DeviceKVCache] = None,
                with_head: bool = True, last_only: bool = False, gather_logits: bool = True) -> ModelOutput:
        B, S = input_ids.shape
        c = self.config
        own_cache = cache is None
        if own_cache:
            cache = self.new_cache(B, S)
        if cache.index + S > cache.capacity:
            raise ValueError(f"KV cache overflow: {cache.index}+{S} > capacity {cache.capacity}")
        if cache.index == 0:
            # a cache object is reused across requests (`prepare`): the left-pad counts of the previous request
            # must never leak into a request that comes without a mask
            if attention_mask is None:
                cache.pad_left.zero_()
            else:
                cache.pad_left.copy_(self.pad_left_from_mask(attention_mask, B, S, self.device))
        positions = None
        if position_ids is not None:
            positions = position_ids.to(torch.int32).reshape(-1).contiguous()
        if (self.tp > 1 and S > 1 and last_only and with_head and positions is None and self.hd == 128
                and (B * S) % (128 * self.tp) == 0 and self.sequence_parallel_prefill):
            # sequence-parallel prefill: all-gather→GEMM and GEMM→reduce-scatter fused over NVLink peer memory
            # a cache that lives beyond this call (decode state, caller-owned) is worth a graph; a throw-away one is not
            logits = self._prefill_sp(input_ids, cache, B, S) if own_cache else self._prefill_sp_graphed(input_ids, cache, B, S)
            cache.set_index(cache.index + S)
            out = ModelOutput(past_key_values=None if own_cache else cache)
            if gather_logits:
                import torch.distributed as dist
                parts = [torch.empty_like(logits) for _ in range(self.tp)]
                dist.all_gather(parts, logits.contiguous(), group=self.comm.group)
                logits = torch.cat(parts, dim=-1)
            out.logits = logits.view(B, 1, -1)
            return out
        if self.comm is not None:
            self.comm.begin_sequence()
        x = K.embed(input_ids.reshape(-1), self.embedding)
        x = self._layers(x, B, S, cache, positions)
        if self.comm is not None:
            self.comm.end_sequence()
            if self.tp > 1:
                self.comm.nvls_end_forward()        # in-switch all-reduces of this forward: advance their epoch base
        cache.set_index(cache.index + S)
        out = ModelOutput(past_key_values=None if own_cache else cache)
        if last_only:
            x = x.view(B, S, c.hidden_size)[:, -1].contiguous()
            rows = 1
        else:
            rows = S
        if with_head:
            logits = K.linear(x, self.lm_head, norm_w=self.final_norm, eps=c.rms_norm_eps, out_f32=True)
            if self.tp > 1 and gather_logits:
                # API parity: `.logits` is the full vocabulary.  This gather is off the generate() path,
                # which exchanges only top-k candidates (see _select_and_advance).
                import torch.distributed as dist
                parts = [torch.empty_like(logits) for _ in range(self.tp)]
                dist.all_gather(parts, logits.contiguous(), group=self.comm.group)
                logits = torch.cat(parts, dim=-1)
            out.logits = logits.view(B, rows, -1)
            out.last_hidden_state = None
        else:
            out.last_hidden_state = K.rmsnorm(x, self.final_norm, c.rms_norm_eps).view(B, rows, c.hidden_size)
        return out

    # ------------------------------------------------------------------ decode state + graph
    def _decode_state(self, B: int, max_len: int) -> DecodeState:
        key = (B, max_len)
        if key in self._decode_states:
            return self._decode_states[key]
        c, dev = self.config, self.device
        i32 = dict(dtype=torch.int32, device=dev)
        bf = dict(dtype=torch.bfloat16, device=dev)
        splits, scratch = K.decode_attn_fused_plan(B, self.nh, self.nkv, self.hd, max_len, dev)
        st = DecodeState(
            B=B, max_len=max_len, ids=torch.zeros(B, **i32), x=torch.zeros(B, c.hidden_size, **bf),
            qkv=torch.zeros(B, (self.nh + 2 * self.nkv) * self.hd, **bf),
            attn=torch.zeros(B, self.nh * self.hd, **bf), mlp=torch.zeros(B, c.intermediate_size // self.tp, **bf),
            logits=torch.zeros(B, self.v_local, dtype=torch.float32, device=dev),
            tok=torch.zeros(B, **i32),
            finished=torch.zeros(B, **i32), sequences=torch.zeros(B, max_len, **i32),
            cur_len=torch.zeros(1, **i32), step=torch.zeros(1, **i32), n_unfinished=torch.zeros(1, **i32),
            splits=splits, scratch=scratch)
        self._decode_states[key] = st
        return st

    def _sample_full(self, st: DecodeState, cache: DeviceKVCache, sampling: Tuple, eos: int, pad: int,
                     logits: torch.Tensor, wait_mega: bool) -> None:
        """Exact selection over the whole vocabulary + bookkeeping, one launch (csrc/sample_full.cu).  `logits` is a
        [B, V] fp32 view (any row pitch).  wait_mega: the rows are being filled by the decode kernels of all ranks
        (multicast stores) — the kernel waits for their candidate packets before it reads."""
        do_sample, temperature, top_k, top_p, seed = sampling
        if getattr(st, "ticket", None) is None:
            st.ticket = torch.zeros(1, dtype=torch.int32, device=self.device)
        cand = self._mega_peer[5] if (wait_mega and self.tp > 1) else None
        _lib.check(_lib.lib().jl_sample_full(
            logits.data_ptr(), st.B, self.config.vocab_size, logits.stride(0),
            float(temperature) if do_sample else 0.0, int(top_k or 0), float(top_p), int(seed), st.step.data_ptr(),
            st.tok.data_ptr(), cand, self._mega_gen.data_ptr(), self.tp, self.rank, K.num_sms(self.device),
            len(self.layers), st.ticket.data_ptr(), st.ids.data_ptr(), st.finished.data_ptr(), st.sequences.data_ptr(),
            st.max_len, int(eos), int(pad), st.cur_len.data_ptr(), cache.len_dev.data_ptr(), st.step.data_ptr(),
            st.n_unfinished.data_ptr(), K._stream()), "jl_sample_full")

    def _gather_logits(self, st: DecodeState) -> torch.Tensor:
        """tp > 1: this rank's [B, V/tp] logits shard → every rank's [B, V] buffer over NVLS (multimem.st), then a
        stream-side wait for all ranks' shards.  Returns the local [B, V] view."""
        self._mega_tp_args()                       # the symmetric buffers of the decode kernel exist
        buf, buf_mc = self._logits_full, self._logits_full_mc
        if st.B > buf.shape[0]:
            # batches beyond the decode kernel's 8 rows (per-op path): a larger gathered-logits buffer of its own
            big = self.__dict__.get("_logits_big")
            if big is None or big[0].shape[0] < st.B:
                from ..parallel.symm import SymmetricHeap
                rows = 1 << (st.B - 1).bit_length()
                heap = SymmetricHeap(self.tp, self.rank, self.device, rows * self.config.vocab_size * 4,
                                     group=getattr(self.comm, "group", None))
                big = self._logits_big = (heap.view(0, (rows, self.config.vocab_size), torch.float32), heap.mc(0), heap)
            buf, buf_mc = big[0], big[1]
        L = _lib.lib()
        from ..parallel.comm import _L as comm_lib
        comm_lib()
        cm = self.comm
        # flag value = device epoch counter + position of this collective in the step: the same captured launch stays
        # correct on every CUDA-graph replay (the counter is advanced at the end of the step, `end_sequence`)
        epoch_rel = cm.call_index + 1
        cm.call_index += 1
        _lib.check(L.jl_nvls_bcast_f32(st.logits.data_ptr(), buf_mc, st.B, self.v_local, self.v_local,
                                       self.config.vocab_size, self.rank * self.v_local, self.tp, self.rank, cm.flag_ptrs,
                                       self._bcast_flag_off, epoch_rel, cm.seq.data_ptr(), self._bcast_ticket.data_ptr(),
                                       K._stream()), "jl_nvls_bcast_f32")
        _lib.check(L.jl_wait_flags(cm._own_flags + 4 * self._bcast_flag_off, epoch_rel, cm.seq.data_ptr(), self.tp,
                                   self.rank, K._stream()), "jl_wait_flags")
        return buf[:st.B]

    def _select_and_advance(self, st: DecodeState, cache: DeviceKVCache, sampling: Tuple, eos: int, pad: int) -> None:
        """st.logits (this rank's vocabulary shard) → token (greedy or temperature/top-k/top-p) → sequences / finished /
        counters, on device: one launch of the exact full-vocabulary kernel (csrc/sample_full.cu), at tp > 1 after an
        NVLS gather of the shards so that every rank selects from identical logits."""
        full = st.logits if self.tp == 1 else self._gather_logits(st)
        self._sample_full(st, cache, sampling, eos, pad, full, wait_mega=False)

    def _decode_step(self, st: DecodeState, cache: DeviceKVCache, sampling: Tuple, eos: int, pad: int) -> None:
        """One token for every row: ids → logits → next ids.  All state on device; capturable."""
        c = self.config
        if self.comm is not None:
            self.comm.begin_sequence()
        if self._mega_fits(st.B):
            # whole model in one persistent kernel: embedding → L layers → final norm → LM head (+ running arg-max)
            L = _lib.lib()
            sms = K.num_sms(self.device)
            shape = (st.B, c.hidden_size, c.intermediate_size // self.tp, self.nh, self.nkv, sms)
            if getattr(st, "mega_scratch", None) is None:
                nbytes = int(L.jl_decode_mega_scratch_bytes(*shape))
                st.mega_scratch = torch.zeros((nbytes + 3) // 4, dtype=torch.int32, device=self.device)
                st.mega_cand = st.mega_scratch.data_ptr() + int(L.jl_decode_mega_cand_offset(*shape))
            do_sample = bool(sampling[0])
            tp_args = self._mega_tp_args()
            if do_sample and self.tp > 1:
                # sampling under tensor parallelism: every rank's shard goes into every rank's [B, V] buffer (multimem.st)
                logits_args = (self._logits_full_mc, c.vocab_size, self.rank * self.v_local, 1)
            elif do_sample or self.mega_write_logits:
                logits_args = (st.logits.data_ptr(), self.v_local, 0, 0)
            else:
                logits_args = (None, 0, 0, 0)
            _lib.check(L.jl_decode_mega(
                self.wqkv_all.data_ptr(), self.wo_all.data_ptr(), self.w13_all.data_ptr(), self.w2_all.data_ptr(),
                self.lm_head.data_ptr(), self.attn_norm_all.data_ptr(), self.ffn_norm_all.data_ptr(),
                self.final_norm.data_ptr(), self.embedding.data_ptr(), st.ids.data_ptr(),
                *logits_args,
                cache.k.data_ptr(), cache.v.data_ptr(), self.cos.data_ptr(), self.sin.data_ptr(),
                cache.len_dev.data_ptr(), cache.pad_left.data_ptr(), st.mega_scratch.data_ptr(),
                self._mega_gen.data_ptr(), st.B,
                len(self.layers), c.hidden_size, c.intermediate_size // self.tp, self.v_local, c.vocab_size,
                self.nh, self.nkv, self.hd, cache.capacity, self.cos.shape[0], float(c.rms_norm_eps),
                1.0 / (self.hd ** 0.5), sms, *tp_args, K._stream()), "jl_decode_mega")
            if not do_sample:
                # greedy: the kernel left one candidate per CTA (and rank); merge + bookkeeping in one launch
                cand = st.mega_cand if self.tp == 1 else self._mega_peer[5]
                _lib.check(L.jl_mega_select(
                    cand, self._mega_gen.data_ptr(), self.tp, self.rank, sms, st.B, len(self.layers),
                    st.tok.data_ptr(), st.ids.data_ptr(), st.finished.data_ptr(), st.sequences.data_ptr(), st.max_len,
                    int(eos), int(pad), st.cur_len.data_ptr(), cache.len_dev.data_ptr(), st.step.data_ptr(),
                    st.n_unfinished.data_ptr(), K._stream()), "jl_mega_select")
                if self.comm is not None:
                    self.comm.end_sequence()
                return
            # sampling: exact full-vocabulary kernel on the logits the decode kernel just wrote (second and last launch)
            full = st.logits if self.tp == 1 else self._logits_full[:st.B]
            self._sample_full(st, cache, sampling, eos, pad, full, wait_mega=True)
            if self.comm is not None:
                self.comm.end_sequence()
            return
        else:
            K.embed(st.ids, self.embedding, out=st.x)
            x = self._layers(st.x, st.B, 1, cache, None, ws=st)
            if getattr(st, "xn_final", None) is not None:
                K.linear(st.xn_final, self.lm_head, out_f32=True, out=st.logits)
            else:
                K.linear(x, self.lm_head, norm_w=self.final_norm, eps=c.rms_norm_eps, out_f32=True, out=st.logits)
        self._select_and_advance(st, cache, sampling, eos, pad)
        if self.comm is not None:
            self.comm.end_sequence()
            if self.tp > 1:
                self.comm.nvls_end_forward()

    def _graph_for(self, st: DecodeState, cache: DeviceKVCache, sampling: Tuple, eos: int, pad: int
                   ) -> Optional[torch.cuda.CUDAGraph]:
        if not self.use_cuda_graph:
            return None
        key = (cache.k.data_ptr(), sampling, eos, pad)
        if key in st.graphs:
            return st.graphs[key]
        # warm up outside capture on a side stream (first-use initialisation, cudaFuncSetAttribute, …);
        # the step is re-playable because a snapshot of the mutable state is restored afterwards
        snap = [t.clone() for t in (st.ids, st.finished, st.sequences, st.cur_len, st.step, st.n_unfinished,
                                    cache.len_dev)]
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            self._decode_step(st, cache, sampling, eos, pad)
        torch.cuda.current_stream().wait_stream(side)
        for t, s in zip((st.ids, st.finished, st.sequences, st.cur_len, st.step, st.n_unfinished, cache.len_dev), snap):
            t.copy_(s)
        g = torch.cuda.CUDAGraph()
        n0 = _lib.launch_count()
        with torch.cuda.graph(g):
            self._decode_step(st, cache, sampling, eos, pad)
        st.kernels_per_step = _lib.launch_count() - n0      # kernel nodes in the captured step
        for t, s in zip((st.ids, st.finished, st.sequences, st.cur_len, st.step, st.n_unfinished, cache.len_dev), snap):
            t.copy_(s)
        st.graphs[key] = g
        return g

    def launches_per_step(self, st: DecodeState) -> Optional[int]:
        """Native kernel launches in one decode step (graph nodes when the step is graph-replayed)."""
        return getattr(st, "kernels_per_step", None)

    # ------------------------------------------------------------------ generation
    @torch.no_grad()
    def generate(self, input_ids: torch.Tensor, attention_mask: Optional[torch.Tensor] = None, *,
                 max_length: int, do_sample: bool = False, temperature: float = 1.0,
                 top_k: Optional[int] = 50, top_p: float = 1.0, pad_token_id: int = 0, eos_token_id: int = -1,
                 seed: int = 0, check_every: int = 16) -> torch.Tensor:
        """Prefill + graph-replayed decode.  Returns [B, max_length] int64 (pad-filled), like HF generate."""
        B, S = input_ids.shape
        if max_length > 2 * self.config.max_sequence_length:
            raise ValueError(f"max_length {max_length} exceeds the rotary table (2 x max_sequence_length)")
        from ..utils.stats import CudaTimer, GenerationStats, nvtx_range
        st = self._decode_state(B, max_length)
        stats = GenerationStats(batch=B, prompt_len=S, weight_bytes_per_step=self.weight_bytes_per_step())
        with nvtx_range("prefill"), CudaTimer() as t_prefill:
            cache = self.prepare(st, input_ids, attention_mask, pad_token_id)
        sampling = (bool(do_sample), float(temperature), int(top_k or 0), float(top_p), int(seed))
        # first new token comes from the prefill logits
        self._select_and_advance_after_prefill(st, cache, sampling, eos_token_id, pad_token_id)
        remaining = max_length - S - 1
        graph = self._graph_for(st, cache, sampling, eos_token_id, pad_token_id) if remaining > 0 else None
        done = 0
        windows = []
        while done < remaining:
            n = min(check_every, remaining - done)
            with nvtx_range("decode"), CudaTimer() as t_win:
                for _ in range(n):
                    if graph is not None:
                        graph.replay()
                    else:
                        self._decode_step(st, cache, sampling, eos_token_id, pad_token_id)
            windows.append((t_win, n))
            done += n
            if done < remaining and int(st.n_unfinished.item()) == 0:      # one sync per `check_every` tokens
                break
        cache.index = int(cache.len_dev.item())
        # what this call cost (device time, CUDA events): prefill, ms per token of every window of `check_every` steps
        stats.prefill_ms = t_prefill.elapsed_ms()
        stats.new_tokens = 1 + done
        stats.decode_ms_per_token = [t.elapsed_ms() / n for t, n in windows]
        self.last_stats = stats
        return st.sequences.to(torch.long)

    def weight_bytes_per_step(self) -> int:
        """Bytes of weights one decode step streams on this rank (the HBM-roofline numerator)."""
        return sum(t.numel() * 2 for t in (self.wqkv_all, self.wo_all, self.w13_all, self.w2_all, self.lm_head))

    def prepare(self, st: DecodeState, input_ids: torch.Tensor, attention_mask: Optional[torch.Tensor],
                pad: int) -> DeviceKVCache:
        """Reset the decode state for a new request batch and run the prefill."""
        B, S = input_ids.shape
        cache = getattr(st, "_cache", None)
        if cache is None:
            cache = self.new_cache(B, st.max_len)
            st._cache = cache
        cache.set_index(0)
        st.sequences.fill_(int(pad))
        st.sequences[:, :S] = input_ids.to(torch.int32)
        st.finished.zero_()
        st.cur_len.fill_(S)
        st.step.zero_()
        st.n_unfinished.fill_(B)
        out = self.forward(input_ids, attention_mask=attention_mask, cache=cache, last_only=True, gather_logits=False)
        st.logits.copy_(out.logits.view(B, -1))
        return cache

    def _select_and_advance_after_prefill(self, st: DecodeState, cache: DeviceKVCache, sampling: Tuple,
                                          eos: int, pad: int) -> None:
        # the prefill already advanced cache.len_dev by S; `advance` adds one more for the *next* token's
        # slot only after that token is written, so compensate here
        if self.comm is not None:
            self.comm.begin_sequence()
        self._select_and_advance(st, cache, sampling, eos, pad)
        if self.comm is not None:
            self.comm.end_sequence()          # the candidate exchange used an epoch: advance the counter past it
        cache.len_dev.sub_(1)

    # ------------------------------------------------------------------ streaming (token-at-a-time) API
    @torch.no_grad()
    def begin_stream(self, input_ids: torch.Tensor, attention_mask: Optional[torch.Tensor] = None, *,
                     max_length: int, do_sample: bool = False, temperature: float = 1.0,
                     top_k: Optional[int] = 50, top_p: float = 1.0, pad_token_id: int = 0,
                     eos_token_id: int = -1, seed: int = 0) -> "EngineStream":
        """Prefill and select the first new token; subsequent tokens come from `EngineStream.step`."""
        B, S = input_ids.shape
        st = self._decode_state(B, max_length)
        cache = self.prepare(st, input_ids, attention_mask, pad_token_id)
        sampling = (bool(do_sample), float(temperature), int(top_k or 0), float(top_p), int(seed))
        self._select_and_advance_after_prefill(st, cache, sampling, eos_token_id, pad_token_id)
        graph = self._graph_for(st, cache, sampling, eos_token_id, pad_token_id)
        return EngineStream(self, st, cache, sampling, eos_token_id, pad_token_id, graph, S + 1)


class EngineStream:
    """Handle of one in-flight request batch: one `step()` = one token per row, result on the host."""

    def __init__(self, engine: B200Engine, st: DecodeState, cache: DeviceKVCache, sampling: Tuple, eos: int,
                 pad: int, graph: Optional[torch.cuda.CUDAGraph], produced_upto: int):
        self.engine, self.st, self.cache = engine, st, cache
        self.sampling, self.eos, self.pad, self.graph = sampling, eos, pad, graph
        self.length = produced_upto
        self.host_tok = torch.empty(st.B, dtype=torch.int32).pin_memory()
        self.host_tok.copy_(st.ids)                       # first token (from the prefill)
        self.h2d_bytes = st.B * 4
        self.d2h_bytes = st.B * 4

    def last_tokens(self) -> torch.Tensor:
        return self.host_tok

    def step(self, feed: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Advance every row by one token.  `feed` (pinned int32 [B] on the host) overrides the input
        token of this step — the interactive case where the host owns the sequence; otherwise the token
        selected on the device in the previous step is used.  Returns the new tokens in pinned memory."""
        if self.length >= self.st.max_len:
            raise RuntimeError("stream is at max_length")
        if feed is not None:
            self.st.ids.copy_(feed, non_blocking=True)
        if self.graph is not None:
            self.graph.replay()
        else:
            self.engine._decode_step(self.st, self.cache, self.sampling, self.eos, self.pad)
        self.host_tok.copy_(self.st.ids, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        self.length += 1
        return self.host_tok

    def sequences(self) -> torch.Tensor:
        return self.st.sequences[:, :self.length].to(torch.long)
