"""Timing and observability helpers: CUDA-event timers, NVTX ranges, generation statistics.

The reference has no tracing, timers or metrics; its only observability is bare `print`.
"""
from __future__ import annotations

import contextlib
import json
import logging
import statistics
from dataclasses import asdict, dataclass, field
from typing import Dict, Iterator, List, Optional

import torch

logger = logging.getLogger(__name__)


@contextlib.contextmanager
def nvtx_range(name: str) -> Iterator[None]:
    """NVTX range when CUDA is available (shows up in ncu / nsys timelines); no-op otherwise."""
    use = torch.cuda.is_available()
    if use:
        torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        if use:
            torch.cuda.nvtx.range_pop()


class CudaTimer:
    """Device-side timing with CUDA events on the current stream (never wall clock)."""

    def __init__(self) -> None:
        self._start = torch.cuda.Event(enable_timing=True)
        self._stop = torch.cuda.Event(enable_timing=True)

    def __enter__(self) -> "CudaTimer":
        self._start.record()
        return self

    def __exit__(self, *exc) -> None:
        self._stop.record()

    def elapsed_ms(self) -> float:
        self._stop.synchronize()
        return self._start.elapsed_time(self._stop)


def max_over_ranks(value: float, device: Optional[torch.device] = None) -> float:
    """Multi-GPU numbers are the max over ranks of device-measured times."""
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return value
    t = torch.tensor([value], dtype=torch.float64, device=device or torch.device("cuda", torch.cuda.current_device()))
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


@dataclass
class GenerationStats:
    """What one generate() call cost; dumped as JSON by the benchmark scripts."""

    batch: int = 0
    prompt_len: int = 0
    new_tokens: int = 0
    prefill_ms: float = 0.0
    decode_ms_per_token: List[float] = field(default_factory=list)
    weight_bytes_per_step: int = 0
    hbm_peak_gbs: float = 0.0

    def summary(self) -> Dict[str, float]:
        d = self.decode_ms_per_token
        out = {"batch": self.batch, "prompt_len": self.prompt_len, "new_tokens": self.new_tokens,
               "prefill_ms": self.prefill_ms,
               "prefill_tok_s": self.batch * self.prompt_len / (self.prefill_ms / 1e3) if self.prefill_ms else 0.0}
        if d:
            p50 = statistics.median(d)
            out.update(p50_ms_per_token=p50, p90_ms_per_token=sorted(d)[int(0.9 * (len(d) - 1))],
                       decode_tok_s=self.batch / (p50 / 1e3))
            if self.weight_bytes_per_step and self.hbm_peak_gbs:
                out["hbm_roofline_frac"] = self.weight_bytes_per_step / (p50 / 1e3) / (self.hbm_peak_gbs * 1e9)
        return out

    def to_json(self) -> str:
        return json.dumps({**asdict(self), "summary": self.summary()})
