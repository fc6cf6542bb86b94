import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from jax_llama_b200 import LLaMAConfig
from jax_llama_b200.models import reference as ref
from jax_llama_b200.models.engine import B200Engine
from jax_llama_b200.models.llama import tree_to
B = int(os.environ.get("B", "8"))
cfg = LLaMAConfig(vocab_size=1000, hidden_size=512, intermediate_size=1536, num_hidden_layers=int(os.environ.get("L", "3")),
                  num_attention_heads=4, num_key_value_heads=2, max_sequence_length=256, initializer_range=0.05)
tree = tree_to(ref.init_params(cfg, seed=2), dtype=torch.bfloat16, device="cuda")
engines = [B200Engine.from_tree(cfg, tree, use_cuda_graph=False, use_megakernel=f) for f in (True, False)]
engines[0].mega_write_logits = True
ids = torch.randint(3, cfg.vocab_size, (B, 37), generator=torch.Generator().manual_seed(7)).cuda()
outs = []
for eng in engines:
    st = eng._decode_state(B, 64)
    cache = eng.prepare(st, ids, None, 0)
    samp = (False, 1.0, 0, 1.0, 0)
    eng._select_and_advance_after_prefill(st, cache, samp, -7, 0)
    eng._decode_step(st, cache, samp, -7, 0)
    outs.append((st.logits.clone(), cache.k[:, :, :, 37].clone()))
a, b = outs
for m in range(B):
    e = (a[0][m] - b[0][m]).abs().max() / b[0][m].abs().max()
    ek = [(a[1][l, m] - b[1][l, m]).abs().max().item() / (b[1][l, m].abs().max().item() + 1e-9) for l in range(cfg.num_hidden_layers)]
    print(f"row {m}: logits rel err {e:.4f}  k-cache rel err per layer {[round(x, 4) for x in ek]}")
