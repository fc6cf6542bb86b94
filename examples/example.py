#!/usr/bin/env python
"""Command-line text completion — the counterpart of the reference's jax_example.py (same flags).

    python examples/example.py --ckpt_dir Meta-Llama-3-8B --tokenizer_path tokenizer.model --is_llama3 true
    torchrun --nproc-per-node 2 examples/example.py --ckpt_dir llama-2-13b --tokenizer_path tokenizer.model

With torchrun the model is tensor-parallel over the launched ranks (one GPU each); each rank reads only
the Meta shards that overlap its slice.  Without a GPU (or with --device cpu) the fp32 reference backend runs.
"""
import argparse
import logging
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from jax_llama_b200 import (LLaMA, LLaMA2Tokenizer, LLaMA3Tokenizer, LLaMAForCausalLM, Mesh,  # noqa: E402
                            convert_llama_weights, convert_llama_weights_for_rank)
from jax_llama_b200.models.llama import tree_to  # noqa: E402
from jax_llama_b200.parallel.comm import init_distributed  # noqa: E402


def load(ckpt_dir: str, tokenizer_path: str, is_llama3: bool, device: str = "auto", max_seq_len: int = 2048,
         **model_kwargs) -> LLaMA:
    """Checkpoint → ready-to-generate `LLaMA` (mirrors jax_example.load: mesh (1, n_devices) over 'dp','mp')."""
    rank, world, local = init_distributed()
    tokenizer = LLaMA3Tokenizer(tokenizer_path) if is_llama3 else LLaMA2Tokenizer(tokenizer_path)
    use_cuda = torch.cuda.is_available() and device != "cpu"
    mesh = Mesh({"dp": 1, "mp": world})
    if use_cuda:
        dev = torch.device("cuda", local)
        torch.cuda.set_device(dev)
        if world > 1:
            params, config = convert_llama_weights_for_rank(ckpt_dir, tokenizer, world, rank, max_seq_len=max_seq_len)
            model_kwargs.update(tp=world, rank=rank)
        else:
            params, config = convert_llama_weights(ckpt_dir, tokenizer, max_seq_len=max_seq_len, dtype=torch.bfloat16)
        params = tree_to(params, device=dev, dtype=torch.bfloat16)
    else:
        if world > 1:
            raise SystemExit("multi-process runs need one CUDA device per rank")
        params, config = convert_llama_weights(ckpt_dir, tokenizer, max_seq_len=max_seq_len)
    model = LLaMAForCausalLM(config, _do_init=False, **model_kwargs)
    return LLaMA(params, model, tokenizer, mesh=mesh)


def main() -> None:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--ckpt_dir", required=True)
    ap.add_argument("--tokenizer_path", required=True)
    ap.add_argument("--is_llama3", type=lambda s: str(s).lower() in ("1", "true", "yes"), default=False)
    ap.add_argument("--max_gen_len", type=int, default=256)
    ap.add_argument("--temperature", type=float, default=0.8)
    ap.add_argument("--top_p", type=float, default=0.95)
    ap.add_argument("--top_k", type=int, default=50)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--device", default="auto", choices=["auto", "cpu"])
    ap.add_argument("--prompt", action="append", help="may be given several times")
    ap.add_argument("-v", "--verbose", action="store_true")
    a = ap.parse_args()
    logging.basicConfig(level=logging.DEBUG if a.verbose else logging.WARNING)
    generator = load(a.ckpt_dir, a.tokenizer_path, a.is_llama3, device=a.device)
    prompts = a.prompt or ["The capital of Germany is the city of",
                           "Here is my sonnet in the style of Shakespeare about an artificial intelligence:"]
    results = generator.generate_from_str(prompts, max_gen_len=a.max_gen_len, temperature=a.temperature,
                                          top_p=a.top_p, top_k=a.top_k, seed=a.seed)
    if int(os.environ.get("RANK", "0")) == 0:
        for r in results:
            print(r)
            print("\n==================================\n")


if __name__ == "__main__":
    main()
