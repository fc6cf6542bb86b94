"""The command-line entry points a reference user would reach for: parity harness (jax_test.py counterpart),
converter CLI (what convert.sh calls), packaging."""
import os
import subprocess
import sys

import pytest
import torch

from jax_llama_b200 import LLaMAConfig
from jax_llama_b200.checkpoint import load_rank, read_layout
from jax_llama_b200.convert_weights import main as convert_main
from jax_llama_b200.convert_weights import write_meta_checkpoint
from jax_llama_b200.models import reference as ref

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_parity_harness_runs_on_cpu():
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "parity_test.py"), "--trials", "3"], env=env,
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "ALL PASSED" in r.stdout
    for name in ("RMSNorm", "precompute_freqs_cis", "apply_rotary_emb", "Attention", "FeedForward",
                 "TransformerBlock", "Transformer", "ModelLogits", "ModelGenerations"):
        assert f"PASS {name}" in r.stdout, name


def test_convert_cli_writes_rank_files(tmp_path):
    cfg = LLaMAConfig(vocab_size=64, hidden_size=64, intermediate_size=256, num_hidden_layers=2,
                      num_attention_heads=4, num_key_value_heads=2, max_sequence_length=64)
    tree = ref.init_params(cfg, seed=0)
    src, dst = tmp_path / "meta", tmp_path / "cache"
    write_meta_checkpoint(tree, cfg, str(src), mp=2, dtype=torch.float32, multiple_of=256)
    assert convert_main(["--ckpt_dir", str(src), "--out_dir", str(dst), "--tp", "2", "--dtype", "float32"]) == 0
    assert read_layout(str(dst))["tp"] == 2
    for r in range(2):
        shard, c = load_rank(str(dst), r)
        assert c.hidden_size == 64
        wq = tree["transformer"]["h"]["0"]["attention"]["wq"]["kernel"]
        assert torch.equal(shard["transformer"]["h"]["0"]["attention"]["wq"]["kernel"], wq[:, r * 32:(r + 1) * 32])


def test_setup_py_lists_the_package():
    r = subprocess.run([sys.executable, "setup.py", "--name", "--version"], cwd=ROOT, capture_output=True, text=True)
    assert r.returncode == 0 and "jax_llama_b200" in r.stdout


def test_megakernel_fit_plan():
    """Host-side shared-memory plan of the whole-step decode kernel (no GPU needed): which batches stay on it."""
    from jax_llama_b200.ops import _lib
    stages = _lib.lib().jl_decode_mega_stages
    llama3_8b = (4096, 14336, 32, 8, 128)
    assert stages(1, *llama3_8b) >= 8                       # deep weight ring at batch 1
    assert stages(4, *llama3_8b) >= 4                       # single pass over w2 still fits
    assert stages(8, *llama3_8b) >= 4                       # w2 walked in two K passes
    assert stages(9, *llama3_8b) == 0                       # beyond the 8 batch rows of the MMA's first half
    assert stages(1, 4096, 14336, 32, 8, 64) == 0           # head_dim 64 is not built
    assert stages(4, 8192, 3584, 8, 1, 128) >= 4            # LLaMA-3 70B per-rank shape at tp = 8
    assert stages(8, 8192, 3584, 8, 1, 128) >= 4            # fits because staging and attention scratch share a region


def test_bench_reference_arm_contract():
    """`bench.py --impl reference` must exit 0 and print one JSON line: a measurement, or why there is none."""
    import json
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference"], capture_output=True,
                       text=True, timeout=300, env=dict(os.environ, CUDA_VISIBLE_DEVICES=""))
    assert r.returncode == 0, r.stderr[-1000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and ("unavailable" in line or "value" in line)


def test_example_cli_end_to_end_on_cpu(tmp_path, spm_model):
    """The reference's jax_example.py flow: Meta checkpoint directory + tokenizer.model → printed completions."""
    import sentencepiece as spm
    vocab = spm.SentencePieceProcessor(model_file=spm_model).vocab_size()
    cfg = LLaMAConfig(vocab_size=vocab, hidden_size=64, intermediate_size=256, num_hidden_layers=2,
                      num_attention_heads=4, num_key_value_heads=2, max_sequence_length=64)
    write_meta_checkpoint(ref.init_params(cfg, seed=0), cfg, str(tmp_path / "ckpt"), mp=2, dtype=torch.float32,
                          multiple_of=256)
    for entry in ("jax_example.py", os.path.join("examples", "example.py")):
        r = subprocess.run([sys.executable, os.path.join(ROOT, entry), "--ckpt_dir", str(tmp_path / "ckpt"),
                            "--tokenizer_path", spm_model, "--device", "cpu", "--max_gen_len", "4", "--temperature", "0",
                            "--prompt", "the capital of germany is", "--prompt", "hello world"],
                           capture_output=True, text=True, timeout=300, env=dict(os.environ, CUDA_VISIBLE_DEVICES=""))
        assert r.returncode == 0, r.stderr[-1500:]
        assert r.stdout.count("==================================") == 2
        assert "the capital of germany is" in r.stdout.lower()


def test_reference_arm_runner_with_a_stubbed_reference(tmp_path, monkeypatch):
    """The runner drives the reference's own API (mesh → FlaxLLaMAForCausalLM(_do_init=False) → partition spec →
    LLaMA.generate, /root/reference/jax_example.py:10-31).  jax and flax are not installable here, so the API is stubbed:
    the test pins the call sequence and the JSON contract, and that decode time is the difference of two generate calls."""
    import textwrap
    import types

    pkg = tmp_path / "stubs"
    (pkg / "jax" / "experimental").mkdir(parents=True)
    (pkg / "jax_llama").mkdir()
    (pkg / "jax" / "__init__.py").write_text(textwrap.dedent("""
        import types
        def devices(): return ["gpu0", "gpu1"]
        class _R:
            @staticmethod
            def PRNGKey(s): return s
        random = _R()
        class tree_util:
            @staticmethod
            def tree_map(f, *trees):
                t0 = trees[0]
                if isinstance(t0, dict):
                    return {k: tree_util.tree_map(f, *[t[k] for t in trees]) for k in t0}
                return f(*trees)
        def device_put(p, s): return p
        def block_until_ready(x): return x
    """))
    (pkg / "jax" / "numpy.py").write_text("import numpy as _np\nint32 = _np.int32\nbfloat16 = _np.float32\nasarray = _np.asarray\nones = _np.ones\n")
    (pkg / "jax" / "experimental" / "__init__.py").write_text("")
    (pkg / "jax" / "experimental" / "mesh_utils.py").write_text("def create_device_mesh(shape, devices=None): return [list(devices)]\n")
    (pkg / "jax" / "sharding.py").write_text(textwrap.dedent("""
        class Mesh:
            def __init__(self, devs, axis_names): self.devs, self.axis_names = devs, axis_names
        class NamedSharding:
            def __init__(self, mesh, spec): self.mesh, self.spec = mesh, spec
    """))
    (pkg / "jax_llama" / "__init__.py").write_text(textwrap.dedent("""
        import time, numpy as np
        CALLS = []
        class LLaMAConfig:
            def __init__(self, **kw): self.__dict__.update(kw)
        class FlaxLLaMAForCausalLM:
            def __init__(self, config, _do_init=True, dtype=None): CALLS.append(("model", _do_init)); self.config = config
            def init_weights(self, key, shape): CALLS.append(("init", shape)); return {"lm_head": {"kernel": np.zeros((2, 2))}}
        def get_llama_param_partition_spec(params, fsdp=False):
            CALLS.append(("spec", fsdp)); return {"lm_head": {"kernel": ("mp",)}}
        class LLaMA:
            def __init__(self, params, model, tokenizer, mesh=None): self.mesh = mesh; CALLS.append(("llama", len(mesh.devs[0])))
            def generate(self, tokens, mask, max_gen_len, temperature=0.8, top_p=0.95):
                CALLS.append(("generate", max_gen_len, temperature)); time.sleep(0.002 * max_gen_len)
                return np.zeros((tokens.shape[0], tokens.shape[1] + max_gen_len), dtype=np.int32)
    """))
    for name in [m for m in list(sys.modules) if m == "jax" or m.startswith("jax.") or m == "jax_llama"]:
        monkeypatch.delitem(sys.modules, name)
    monkeypatch.syspath_prepend(str(pkg))
    monkeypatch.syspath_prepend(os.path.join(ROOT, "baseline"))
    monkeypatch.setenv("RANK", "0")
    import importlib

    import reference_runner
    importlib.reload(reference_runner)
    # the stub must win over a real install directory of the reference
    monkeypatch.setattr(reference_runner, "HERE", str(tmp_path))
    args = types.SimpleNamespace(gpus=2, steps=20, warmup=3, model="llama3-8b", batch=1, prompt_len=64)
    line = reference_runner.run(args)
    import jax_llama
    assert line["impl"] == "reference" and "unavailable" not in line, line
    assert line["metric"] == "decode_tokens_per_sec" and line["n_gpus"] == 2 and line["config"]["model"] == "llama3-8b"
    assert 250 < line["value"] < 1000            # 20 tokens in ~40 ms of stubbed decode
    kinds = [c[0] for c in jax_llama.CALLS]
    assert kinds[:4] == ["model", "init", "spec", "llama"] and kinds.count("generate") == 6
    assert ("model", False) in jax_llama.CALLS and ("spec", False) in jax_llama.CALLS
    assert all(c[2] == 0.0 for c in jax_llama.CALLS if c[0] == "generate")
    for name in [m for m in list(sys.modules) if m == "jax" or m.startswith("jax.") or m == "jax_llama"]:
        sys.modules.pop(name, None)


def test_c_api_example_compiles_and_links_against_the_library(tmp_path):
    """include/jllama_b200.h is plain C99 and examples/c_api_example.c builds against the in-tree shared object (it was
    run on a B200: 4 launches; here only the compile + link, no GPU needed)."""
    import shutil
    from jax_llama_b200.ops import build
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cuda = "/usr/local/cuda"
    if shutil.which("gcc") is None or not os.path.exists(os.path.join(cuda, "include", "cuda_runtime_api.h")):
        pytest.skip("gcc or the CUDA runtime headers are not available")
    lib = build.build_library()
    hdr = subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-fsyntax-only", "-x", "c",
                          os.path.join(root, "include", "jllama_b200.h")], capture_output=True, text=True)
    assert hdr.returncode == 0, hdr.stderr
    exe = tmp_path / "c_api_example"
    r = subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", f"-I{root}/include", f"-I{cuda}/include",
                        os.path.join(root, "examples", "c_api_example.c"), f"-L{os.path.dirname(lib)}", "-ljllama_b200",
                        f"-L{cuda}/lib64", "-lcudart", "-lm", "-o", str(exe)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert exe.exists()
