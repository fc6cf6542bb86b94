"""Multi-process tensor-parallel checks (launched by torchrun from test_gpu_tp.py, one rank per GPU).

Every rank builds the same tiny model, takes its slice with the partition rules, runs the B200 engine with
the fused peer-memory collectives, and compares against the un-sharded fp32 reference on rank 0."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from jax_llama_b200 import LLaMAConfig, Mesh  # noqa: E402
from jax_llama_b200.models import reference as ref  # noqa: E402
from jax_llama_b200.models.engine import B200Engine  # noqa: E402
from jax_llama_b200.models.llama import tree_to  # noqa: E402
from jax_llama_b200.parallel.comm import PeerComm, init_distributed  # noqa: E402
from jax_llama_b200.parallel.layout import shard_tree  # noqa: E402


def rel(a, b):
    a, b = a.float().cpu(), b.float().cpu()
    return ((a - b).abs().max() / (b.abs().max() + 1e-6)).item()


def main():
    rank, world, local = init_distributed("nccl")
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    cfg = LLaMAConfig(vocab_size=1024, hidden_size=1024, intermediate_size=2048, num_hidden_layers=2,
                      num_attention_heads=8, num_key_value_heads=8 if world > 4 else 4, max_sequence_length=2048,
                      initializer_range=0.05)
    full = tree_to(ref.init_params(cfg, seed=1), dtype=torch.bfloat16)
    full32 = tree_to(full, dtype=torch.float32)
    mine = tree_to(shard_tree(full, Mesh({"dp": 1, "mp": world}), rank), device=dev)
    comm = PeerComm(world, rank, dev)

    # ---- 1. raw fused GEMV + all-reduce + residual against the un-sharded product
    g = torch.Generator().manual_seed(5)
    for M in (1, 3, 8):
        K_, N = 2048, 1024
        x = (torch.randn(M, K_, generator=g)).to(torch.bfloat16)
        w = (torch.randn(N, K_, generator=g) * 0.05).to(torch.bfloat16)
        res = torch.randn(M, N, generator=g).to(torch.bfloat16)
        want = (x.float() @ w.float().t()).to(torch.bfloat16).float() + res.float()
        ks = K_ // world
        comm.begin_sequence()
        got = comm.gemv_allreduce(x[:, rank * ks:(rank + 1) * ks].contiguous().to(dev),
                                  w[:, rank * ks:(rank + 1) * ks].contiguous().to(dev), res.to(dev))
        got2 = comm.gemv_allreduce(x[:, rank * ks:(rank + 1) * ks].contiguous().to(dev),
                                   w[:, rank * ks:(rank + 1) * ks].contiguous().to(dev), res.to(dev))
        comm.end_sequence()
        torch.cuda.synchronize()
        assert rel(got, want) < 2e-2, (M, rel(got, want))
        assert torch.equal(got, got2)
        # every rank must hold bit-identical sums (sampling stays in lock-step without a broadcast)
        ref_t = got.clone()
        dist.broadcast(ref_t, 0)
        assert torch.equal(ref_t, got), "ranks disagree bitwise"

    # ---- 2. engine: TP logits and greedy generation against the fp32 reference
    eng = B200Engine.from_tree(cfg, mine, tp=world, rank=rank, comm=comm)
    ids = torch.randint(3, cfg.vocab_size, (2, 21), generator=torch.Generator().manual_seed(2))
    want = ref.forward(full32, cfg, ids).logits
    got = eng.forward(ids.to(dev)).logits
    assert got.shape == want.shape and rel(got, want) < 3e-2, rel(got, want)
    mask = torch.ones(2, 21, dtype=torch.int32)
    mask[1, :5] = 0
    seqs = eng.generate(ids.to(dev), mask.to(dev), max_length=40, eos_token_id=-7, pad_token_id=0)
    gathered = [torch.empty_like(seqs) for _ in range(world)]
    dist.all_gather(gathered, seqs)
    assert all(torch.equal(gathered[0], t) for t in gathered), "ranks generated different tokens"
    full_mask = torch.ones(2, 40, dtype=torch.int32)
    full_mask[1, :5] = 0
    pos = ref.default_position_ids(full_mask, 2, 40, "cpu")
    logits = ref.forward(full32, cfg, seqs.cpu(), full_mask, pos).logits
    for t in range(21, 40):
        row = logits[:, t - 1]
        chosen = row.gather(-1, seqs[:, t:t + 1].cpu())[:, 0]
        assert ((row.max(-1).values - chosen) < 3e-2 * row.abs().max()).all(), t
    # sampling path: identical tokens on every rank, repeatable per seed
    a = eng.generate(ids.to(dev), mask.to(dev), max_length=32, do_sample=True, temperature=0.9, top_k=20,
                     top_p=0.9, seed=3, eos_token_id=-7)
    b = eng.generate(ids.to(dev), mask.to(dev), max_length=32, do_sample=True, temperature=0.9, top_k=20,
                     top_p=0.9, seed=3, eos_token_id=-7)
    assert torch.equal(a, b), (a[:, 21:].tolist(), b[:, 21:].tolist())
    ga = [torch.empty_like(a) for _ in range(world)]
    dist.all_gather(ga, a)
    assert all(torch.equal(ga[0], t) for t in ga)
    # pure nucleus sampling (top-k disabled): exact over the whole vocabulary, still identical on every rank
    c1 = eng.generate(ids.to(dev), mask.to(dev), max_length=32, do_sample=True, temperature=1.0, top_k=None,
                      top_p=0.95, seed=11, eos_token_id=-7)
    c2 = eng.generate(ids.to(dev), mask.to(dev), max_length=32, do_sample=True, temperature=1.0, top_k=None,
                      top_p=0.95, seed=11, eos_token_id=-7)
    assert torch.equal(c1, c2) and not torch.equal(c1, a)
    gc = [torch.empty_like(c1) for _ in range(world)]
    dist.all_gather(gc, c1)
    assert all(torch.equal(gc[0], t) for t in gc), "ranks sampled different tokens with top_k=None"
    # batch above the whole-step kernel's 8 rows: per-op decode path — tcgen05 GEMMs + in-switch all-reduce kernel inside
    # the captured step, logits of all ranks gathered over NVLS for the exact sampler
    ids9 = torch.randint(3, cfg.vocab_size, (9, 17), generator=torch.Generator().manual_seed(12))
    seq9 = eng.generate(ids9.to(dev), None, max_length=26, eos_token_id=-7, pad_token_id=0)
    g9 = [torch.empty_like(seq9) for _ in range(world)]
    dist.all_gather(g9, seq9)
    assert all(torch.equal(g9[0], t) for t in g9), "ranks generated different tokens at batch 9"
    logits9 = ref.forward(full32, cfg, seq9.cpu()).logits
    for t in range(17, 26):
        row = logits9[:, t - 1]
        chosen = row.gather(-1, seq9[:, t:t + 1].cpu())[:, 0]
        assert ((row.max(-1).values - chosen) < 3e-2 * row.abs().max()).all(), ("batch 9", t)
    s9a = eng.generate(ids9.to(dev), None, max_length=24, do_sample=True, temperature=0.8, top_k=None, top_p=0.9, seed=4,
                       eos_token_id=-7)
    s9b = eng.generate(ids9.to(dev), None, max_length=24, do_sample=True, temperature=0.8, top_k=None, top_p=0.9, seed=4,
                       eos_token_id=-7)
    assert torch.equal(s9a, s9b)
    # batch 6 handed to the per-op path (what a wide model does under TP, engine.mega_max_batch = 4): swap-AB GEMMs at
    # M = 6 + the fused exchange / all-reduce / norm kernel, against the fp32 model
    old_max = eng.mega_max_batch
    eng.mega_max_batch = 4
    seq6 = eng.generate(ids9[:6].to(dev), None, max_length=26, eos_token_id=-7, pad_token_id=0)
    eng.mega_max_batch = old_max
    g6 = [torch.empty_like(seq6) for _ in range(world)]
    dist.all_gather(g6, seq6)
    assert all(torch.equal(g6[0], t) for t in g6), "ranks generated different tokens at batch 6 (per-op path)"
    logits6 = ref.forward(full32, cfg, seq6.cpu()).logits
    for t in range(17, 26):
        row = logits6[:, t - 1]
        chosen = row.gather(-1, seq6[:, t:t + 1].cpu())[:, 0]
        assert ((row.max(-1).values - chosen) < 3e-2 * row.abs().max()).all(), ("batch 6 per-op", t)
    # ---- 3. sequence-parallel prefill (all-gather→GEMM, GEMM→reduce-scatter fused over peer memory)
    S = 64 * world if world > 2 else 128
    ids_sp = torch.randint(3, cfg.vocab_size, (2, S), generator=torch.Generator().manual_seed(9))
    assert (2 * S) % (128 * world) == 0
    want_sp = ref.forward(full32, cfg, ids_sp).logits[:, -1]
    eng.set_prefill_path("sequence_parallel")           # NVLS reduce-scatter + all-gather around the GEMMs (the default)
    for _ in range(2):                                  # twice: epochs, parities and counters must re-arm
        cache = eng.new_cache(2, S + 8)
        got_sp = eng.forward(ids_sp.to(dev), cache=cache, last_only=True).logits[:, 0]
        assert got_sp.shape == want_sp.shape and rel(got_sp, want_sp) < 3e-2, rel(got_sp, want_sp)
    eng.set_prefill_path("allreduce")                   # same prefill through GEMM + in-switch all-reduce kernel
    cache2 = eng.new_cache(2, S + 8)
    got_ar = eng.forward(ids_sp.to(dev), cache=cache2, last_only=True).logits[:, 0]
    assert rel(got_ar, want_sp) < 3e-2, rel(got_ar, want_sp)
    assert rel(got_sp, got_ar) < 3e-2 and rel(cache.k, cache2.k) < 2e-2
    eng.set_prefill_path(None)
    # one long prompt: two halves pipelined on two streams (the second half attends to the first half's keys)
    S1 = 256 * world
    ids_1 = torch.randint(3, cfg.vocab_size, (1, S1), generator=torch.Generator().manual_seed(10))
    want_1 = ref.forward(full32, cfg, ids_1).logits[:, -1]
    for _ in range(2):
        cache_1 = eng.new_cache(1, S1 + 8)
        got_1 = eng.forward(ids_1.to(dev), cache=cache_1, last_only=True).logits[:, 0]
        assert rel(got_1, want_1) < 3e-2, rel(got_1, want_1)
    os.environ["JLLAMA_SP_PIPELINE"] = "0"
    cache_1b = eng.new_cache(1, S1 + 8)
    got_1b = eng.forward(ids_1.to(dev), cache=cache_1b, last_only=True).logits[:, 0]
    del os.environ["JLLAMA_SP_PIPELINE"]
    assert rel(got_1, got_1b) < 1e-2 and rel(cache_1.k, cache_1b.k) < 1e-2
    # a token count that is not a multiple of 128·tp takes the all-reduce path on its own
    ids_odd = ids_sp[:, :S - 3]
    want_odd = ref.forward(full32, cfg, ids_odd).logits[:, -1]
    got_odd = eng.forward(ids_odd.to(dev), last_only=True).logits[:, 0]
    assert rel(got_odd, want_odd) < 3e-2, rel(got_odd, want_odd)
    # ---- 4. flag / epoch protocol stress: ranks enter every collective at random, different times
    # (a device-side sleep of up to ~200 us per rank per call) for many back-to-back calls; payloads change
    # every iteration so that a stale inbox slot, a reused flag or a parity mix-up shows up as a wrong sum.
    import random
    rnd = random.Random(1234 + rank)
    iters = int(os.environ.get("TP_STRESS_ITERS", "150"))
    K_, N = 256 * world, 512
    w_full = (torch.randn(N, K_, generator=torch.Generator().manual_seed(77)) * 0.05).to(torch.bfloat16)
    w_loc = w_full[:, rank * 256:(rank + 1) * 256].contiguous().to(dev)
    zero = torch.zeros(1, N, dtype=torch.bfloat16, device=dev)
    outs, gath = [], []
    for it in range(iters):
        x_full = torch.full((1, K_), 1.0 + (it % 7), dtype=torch.bfloat16)
        x_loc = x_full[:, rank * 256:(rank + 1) * 256].contiguous().to(dev)
        src = torch.full((8,), float(rank * 1000 + it), dtype=torch.float32, device=dev)
        comm.begin_sequence()
        torch.cuda._sleep(rnd.randrange(0, 400_000))
        outs.append(comm.gemv_allreduce(x_loc, w_loc, zero))
        torch.cuda._sleep(rnd.randrange(0, 400_000))
        gath.append(comm.allgather_small(src, torch.empty(world, 8, dtype=torch.float32, device=dev)))
        comm.end_sequence()
    torch.cuda.synchronize()
    base = w_full.float().sum(1)
    for it in range(iters):
        want_it = (base * (1.0 + (it % 7))).to(torch.bfloat16).float()
        assert rel(outs[it][0], want_it) < 2e-2, ("stress all-reduce", it, rel(outs[it][0], want_it))
        exp = torch.tensor([[r * 1000.0 + it] * 8 for r in range(world)], device=dev)
        assert torch.equal(gath[it], exp), ("stress all-gather", it)
    dist.barrier()
    if rank == 0:
        print(f"TP_WORKER_OK world={world}")
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
