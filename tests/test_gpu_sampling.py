"""Exact full-vocabulary sampler (csrc/sample_full.cu) against the fp32 statement of the HF warper pipeline the
reference configures in /root/reference/jax_llama/generation.py:28-41 (temperature → top-k → top-p → categorical)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from jax_llama_b200.ops import kernels as K  # noqa: E402

V = 128256


def kept_set(logits, temperature, top_k, top_p):
    """Exact kept set and probabilities on the CPU in float64, ties ordered by lower index."""
    z = logits.double().cpu() / temperature
    order = torch.sort(z, descending=True, stable=True).indices          # stable: equal values keep index order
    if top_k and top_k < z.numel():
        order = order[:top_k]
    p = torch.softmax(z[order], 0)
    if top_p < 1.0:
        before = torch.cumsum(p, 0) - p
        order = order[before < top_p]
        p = torch.softmax(z[order], 0)
    return order, p


def test_argmax_picks_lowest_index_among_4096_tied_maxima():
    g = torch.Generator().manual_seed(0)
    x = torch.randn(3, V, generator=g)
    tied = torch.randperm(V, generator=g)[:4096]
    x[:, tied] = 9.0
    tok = K.sample_full(x.cuda(), temperature=0.0)
    assert tok.tolist() == [int(tied.min())] * 3


def test_constant_logits_top_k_keeps_the_lowest_indices():
    x = torch.zeros(V).cuda()
    tok = K.sample_full(x, temperature=1.0, top_k=5, seed=1, rows=4096)
    assert set(tok.tolist()) == {0, 1, 2, 3, 4}


@pytest.mark.parametrize("top_k,top_p,temperature,scale", [(0, 0.9, 0.8, 4.0), (100, 1.0, 1.0, 3.0), (1000, 0.95, 0.7, 1.0),
                                                           (0, 0.5, 1.0, 0.3)])
def test_samples_stay_inside_the_exact_nucleus_and_follow_its_distribution(top_k, top_p, temperature, scale):
    """No candidate budget: `top_k=None` is the nucleus over the whole vocabulary (flat logits keep tens of thousands of
    tokens), top_k = 100 / 1000 exceed round 1's 64-candidate limit."""
    g = torch.Generator().manual_seed(2)
    x = torch.randn(V, generator=g) * scale
    order, p = kept_set(x, temperature, top_k, top_p)
    n = 1 << 16
    tok = K.sample_full(x.cuda(), temperature=temperature, top_k=top_k, top_p=top_p, seed=7, rows=n).cpu().long()
    member = torch.zeros(V, dtype=torch.bool)
    member[order] = True
    assert bool(member[tok].all()), "a sampled token lies outside the exact kept set"
    # total variation between the empirical distribution and the exact one; the noise floor for `n` draws over `m`
    # atoms is about sqrt(m / (2 pi n))
    emp = torch.bincount(tok, minlength=V).double() / n
    exact = torch.zeros(V, dtype=torch.float64)
    exact[order] = p
    tv = 0.5 * (emp - exact).abs().sum().item()
    floor = (order.numel() / (6.28 * n)) ** 0.5
    assert tv < 1.5 * floor + 0.01, (tv, floor, order.numel())
    # the most likely token is sampled about as often as it should be
    top = int(order[0])
    assert abs(emp[top].item() - exact[top].item()) < 5 * (exact[top].item() / n) ** 0.5 + 1e-3


def test_same_seed_same_tokens_different_seed_different_tokens():
    x = (torch.randn(8, V, generator=torch.Generator().manual_seed(3)) * 2).cuda()
    step = torch.tensor([5], dtype=torch.int32, device="cuda")
    a = K.sample_full(x, temperature=0.9, top_k=50, top_p=0.9, seed=1, step=step)
    b = K.sample_full(x, temperature=0.9, top_k=50, top_p=0.9, seed=1, step=step)
    c = K.sample_full(x, temperature=0.9, top_k=50, top_p=0.9, seed=2, step=step)
    assert torch.equal(a, b) and not torch.equal(a, c)
