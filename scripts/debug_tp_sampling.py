import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from jax_llama_b200 import LLaMAConfig, Mesh
from jax_llama_b200.models import reference as ref
from jax_llama_b200.models.engine import B200Engine
from jax_llama_b200.models.llama import tree_to
from jax_llama_b200.parallel.comm import PeerComm, init_distributed
from jax_llama_b200.parallel.layout import shard_tree
rank, world, local = init_distributed("nccl")
dev = torch.device("cuda", local); torch.cuda.set_device(dev)
cfg = LLaMAConfig(vocab_size=1024, hidden_size=1024, intermediate_size=2048, num_hidden_layers=2, num_attention_heads=8,
                  num_key_value_heads=4, max_sequence_length=256, initializer_range=0.05)
full = tree_to(ref.init_params(cfg, seed=1), dtype=torch.bfloat16)
mine = tree_to(shard_tree(full, Mesh({"dp": 1, "mp": world}), rank), device=dev)
for graph in (False, True):
    eng = B200Engine.from_tree(cfg, mine, tp=world, rank=rank, use_cuda_graph=graph)
    ids = torch.randint(3, cfg.vocab_size, (2, 21), generator=torch.Generator().manual_seed(2)).to(dev)
    outs = [eng.generate(ids, None, max_length=32, do_sample=True, temperature=0.9, top_k=20, top_p=0.9, seed=3, eos_token_id=-7) for _ in range(4)]
    if rank == 0:
        for i, o in enumerate(outs):
            print("graph", graph, "run", i, o[:, 21:].tolist())
    # logits repeatability of single steps
    eng.mega_write_logits = True
    st = eng._decode_state(2, 32)
    L = []
    for rep in range(3):
        cache = eng.prepare(st, ids, None, 0)
        samp = (False, 1.0, 0, 1.0, 0)
        eng._select_and_advance_after_prefill(st, cache, samp, -7, 0)
        ls = []
        for _ in range(4):
            eng._decode_step(st, cache, samp, -7, 0)
            ls.append(st.logits.clone())
        L.append(torch.stack(ls))
    if rank == 0:
        print("graph", graph, "greedy-step logits equal across reps:", [bool(torch.equal(L[0], x)) for x in L[1:]],
              "max diff", [float((L[0] - x).abs().max()) for x in L[1:]])
dist.barrier()
