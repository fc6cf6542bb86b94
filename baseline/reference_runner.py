"""Reference arm of bench.py: the UNMODIFIED reference (installed under baseline/_ref) driven through its own public
API and stock code path, for the same metric and config as our arm.

What it runs is the reference's `jax_example.load` flow (/root/reference/jax_example.py:10-31) with random-init
parameters in place of a checkpoint (there are none on the box) and then `LLaMA.generate`
(/root/reference/jax_llama/generation.py:22-45):

    mesh (1, n_devices) over ('dp', 'mp')  →  FlaxLLaMAForCausalLM(config, _do_init=False)
    →  params = model.init_weights(...) placed with get_llama_param_partition_spec(params, fsdp=False)
    →  LLaMA(params, model, tokenizer, mesh).generate(tokens, mask, max_gen_len, temperature=0.0, top_p=1.0)

The reference is a single-controller JAX program: one process drives every visible GPU, so under torchrun only rank 0
runs it and the other ranks return at once.  Its API has no separate decode step: decode throughput is the difference
of two `generate` calls of different lengths on the same prompt (both shapes compiled by a first, untimed call), which
removes the prefill and the fixed costs.  Nothing of this repository's models, kernels or engine is imported here.
"""
from __future__ import annotations

import json
import os
import sys
import time
from typing import Any, Dict

HERE = os.path.dirname(os.path.abspath(__file__))


class _Tok:
    """The three attributes of a tokenizer that `LLaMA.generate` reads."""

    def __init__(self, eos: int):
        self.eos_id, self.bos_id, self.pad_id = eos, 1, -1


MODELS = {      # the shapes of jax_llama_b200.config.preset, restated so that nothing of our package is imported
    "llama3-8b": dict(vocab_size=128256, hidden_size=4096, intermediate_size=14336, num_hidden_layers=32,
                      num_attention_heads=32, num_key_value_heads=8, rms_norm_eps=1e-5, rope_theta=500000.0),
    "llama3-70b": dict(vocab_size=128256, hidden_size=8192, intermediate_size=28672, num_hidden_layers=80,
                       num_attention_heads=64, num_key_value_heads=8, rms_norm_eps=1e-5, rope_theta=500000.0),
    "llama2-7b": dict(vocab_size=32000, hidden_size=4096, intermediate_size=11008, num_hidden_layers=32,
                      num_attention_heads=32, num_key_value_heads=32, rms_norm_eps=1e-5, rope_theta=10000.0),
    "llama2-13b": dict(vocab_size=32000, hidden_size=5120, intermediate_size=13824, num_hidden_layers=40,
                       num_attention_heads=40, num_key_value_heads=40, rms_norm_eps=1e-5, rope_theta=10000.0),
}


def run(args: Any) -> Dict[str, Any]:
    """Returns the JSON line of the reference arm (a measurement, or {"unavailable": why})."""
    ref_dir = os.path.join(HERE, "_ref")
    if ref_dir not in sys.path:
        sys.path.insert(0, ref_dir)
    try:
        import jax
        import jax.numpy as jnp
        import numpy as np
        from jax.experimental import mesh_utils
        from jax.sharding import Mesh, NamedSharding

        from jax_llama import FlaxLLaMAForCausalLM, LLaMA, LLaMAConfig, get_llama_param_partition_spec
    except Exception as e:  # noqa: BLE001
        return {"impl": "reference", "unavailable": f"reference import fails: {type(e).__name__}: {str(e)[:140]}"}
    if int(os.environ.get("RANK", "0")) != 0:
        return {}                                  # single-controller: rank 0 drives every GPU
    try:
        devices = jax.devices()
        n = min(int(args.gpus), len(devices))
        mesh = Mesh(mesh_utils.create_device_mesh((1, n), devices=devices[:n]), axis_names=("dp", "mp"))
        P, B = int(args.prompt_len), int(args.batch)
        steps, warm = int(args.steps), max(3, int(args.warmup))
        cfg = LLaMAConfig(max_sequence_length=P + warm + steps + 8, **MODELS[args.model])
        model = FlaxLLaMAForCausalLM(cfg, _do_init=False, dtype=jnp.bfloat16)
        params = model.init_weights(jax.random.PRNGKey(0), (1, 1))
        params = jax.tree_util.tree_map(lambda x: x.astype(jnp.bfloat16), params)
        spec = get_llama_param_partition_spec(params, fsdp=False)
        params = jax.tree_util.tree_map(lambda p, s: jax.device_put(p, NamedSharding(mesh, s)), params, spec)
        llama = LLaMA(params, model, _Tok(eos=cfg.vocab_size + 7), mesh=mesh)      # eos never fires: fixed-length runs
        rng = np.random.default_rng(1234)
        tokens = jnp.asarray(rng.integers(0, cfg.vocab_size, (B, P)), dtype=jnp.int32)
        mask = jnp.ones((B, P), dtype=jnp.int32)

        def timed(gen_len: int) -> float:
            t0 = time.perf_counter()
            out = llama.generate(tokens, mask, gen_len, temperature=0.0, top_p=1.0)
            jax.block_until_ready(out)
            return time.perf_counter() - t0

        short, long_ = warm, warm + steps
        timed(short), timed(long_)                                   # compile both shapes (untimed)
        t_short = min(timed(short) for _ in range(2))
        t_long = min(timed(long_) for _ in range(2))
        dt = max(t_long - t_short, 1e-9)
        return {"impl": "reference", "metric": "decode_tokens_per_sec", "value": round(B * steps / dt, 2),
                "unit": "tokens/s", "n_gpus": n, "steps": steps, "warmup": warm, "ms_per_step": round(dt / steps * 1e3, 4),
                "higher_is_better": True, "scaling": "strong", "dtype": "bf16",
                "data": "synthetic token ids, random-init weights",
                "config": {"model": args.model, "global_batch": B, "seq_len": P, "parallelism": f"tp{n}",
                           "timing": "generate(P+W+K) minus generate(P+W), wall clock around block_until_ready"},
                "prefill_plus_warmup_s": round(t_short, 4)}
    except Exception as e:  # noqa: BLE001
        return {"impl": "reference", "unavailable": f"reference run fails: {type(e).__name__}: {str(e)[:160]}"}


def main(args: Any) -> None:
    line = run(args)
    if line:
        print(json.dumps(line))
