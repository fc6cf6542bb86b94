from .stats import CudaTimer, GenerationStats, max_over_ranks, nvtx_range

__all__ = ["CudaTimer", "GenerationStats", "max_over_ranks", "nvtx_range"]
