"""Numerics of every sm_100a kernel against a plain PyTorch fp32 statement of the same op."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu

from jax_llama_b200.models import reference as ref      # noqa: E402
from jax_llama_b200.ops import kernels as K             # noqa: E402

DEV = "cuda"


def rel_err(got: torch.Tensor, want: torch.Tensor) -> float:
    got, want = got.float().cpu(), want.float().cpu()
    return ((got - want).abs().max() / (want.abs().max() + 1e-6)).item()


def assert_close(got: torch.Tensor, want: torch.Tensor, tol: float) -> None:
    """Global max error AND, per row and per column, max error relative to that row's / column's own scale: a wrong
    small-magnitude element, a wrong padded-tail row or a wrong last column fails even when the global maximum hides it.
    Then an element-wise allclose with a scale-relative absolute floor (bf16 cancellation near zero)."""
    g, w = got.float().cpu(), want.float().cpu()
    assert g.shape == w.shape
    assert rel_err(g, w) < tol
    g2, w2 = g.reshape(-1, g.shape[-1]), w.reshape(-1, w.shape[-1])
    d = (g2 - w2).abs()
    row = d.max(1).values / (w2.abs().max(1).values + 1e-2 * w2.abs().max())
    col = d.max(0).values / (w2.abs().max(0).values + 1e-2 * w2.abs().max())
    assert row.max().item() < 2.5 * tol, ("row", int(row.argmax()), row.max().item())
    assert col.max().item() < 2.5 * tol, ("column", int(col.argmax()), col.max().item())
    assert torch.allclose(g, w, rtol=4 * tol, atol=tol * float(w.abs().max())), "element-wise allclose"


def rnd(*shape, scale=1.0, seed=0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(torch.bfloat16)


def test_library_loads_and_counts_launches():
    from jax_llama_b200.ops import _lib
    assert _lib.lib().jl_version() == 100
    before = _lib.launch_count()
    K.swiglu(rnd(4, 64).to(DEV))
    assert _lib.launch_count() == before + 1


def test_embed():
    table = rnd(1000, 256)
    ids = torch.tensor([0, 5, 999, 17, -1, 5000])
    got = K.embed(ids.to(DEV), table.to(DEV))
    want = table[ids.clamp(0, 999)]
    assert torch.equal(got.cpu(), want)


@pytest.mark.parametrize("rows", [7, 300])
@pytest.mark.parametrize("H", [256, 2048, 4096, 5120, 8192])
def test_rmsnorm(H, rows):
    x, w, d = rnd(rows, H, seed=1), rnd(H, seed=2) * 0.1 + 1, rnd(rows, H, seed=3)
    want = ref.rms_norm(x.float(), w.float(), 1e-5)
    got = K.rmsnorm(x.to(DEV), w.to(DEV), 1e-5)
    assert rel_err(got, want) < 1e-2
    xd = x.to(DEV).clone()
    got2 = K.rmsnorm(xd, w.to(DEV), 1e-5, delta=d.to(DEV))
    s = (x.float() + d.float()).to(torch.bfloat16)
    assert torch.equal(xd.cpu(), s)                                     # residual stream updated in place
    assert rel_err(got2, ref.rms_norm(s.float(), w.float(), 1e-5)) < 1e-2


@pytest.mark.parametrize("nh,nkv,hd", [(8, 2, 128), (4, 4, 64)])
def test_rope_kv_write(nh, nkv, hd):
    B, S, T = 2, 5, 16
    qkv = rnd(B * S, (nh + 2 * nkv) * hd, seed=4)
    cos, sin = ref.rope_cos_sin(hd, 64, 10000.0)
    pad = torch.tensor([0, 2], dtype=torch.int32)
    clen = torch.tensor([3], dtype=torch.int32)
    kc, vc = torch.zeros(B, nkv, T, hd, dtype=torch.bfloat16), torch.zeros(B, nkv, T, hd, dtype=torch.bfloat16)
    want_qkv = qkv.clone()
    K.rope_kv_write(want_qkv, kc, vc, cos, sin, B, S, nh, nkv, hd, cache_len=clen, pad_left=pad)   # CPU statement
    # the CPU statement itself against the complex-number oracle
    x = qkv.view(B, S, nh + 2 * nkv, hd)
    pos = (3 + torch.arange(S)[None] - pad[:, None].long()).clamp(min=0)
    f = ref.precompute_freqs_cis(hd, 64)[pos]
    oq, ok = ref.apply_rotary_emb(x[:, :, :nh].float(), x[:, :, nh:nh + nkv].float(), f)
    assert rel_err(want_qkv.view(B, S, -1, hd)[:, :, :nh], oq) < 1e-2
    assert rel_err(kc[:, :, 3:8].transpose(1, 2), ok) < 1e-2
    g_qkv, gk, gv = qkv.to(DEV), torch.zeros_like(kc).to(DEV), torch.zeros_like(vc).to(DEV)
    K.rope_kv_write(g_qkv, gk, gv, cos.to(DEV), sin.to(DEV), B, S, nh, nkv, hd, cache_len=clen.to(DEV),
                    pad_left=pad.to(DEV))
    assert rel_err(g_qkv, want_qkv) < 1e-2 and rel_err(gk, kc) < 1e-2 and torch.equal(gv.cpu(), vc)
    assert gk[:, :, :3].abs().sum() == 0 and gk[:, :, 8:].abs().sum() == 0
    # explicit positions + separate q buffer
    q_out = torch.zeros(B * S, nh * hd, dtype=torch.bfloat16, device=DEV)
    g2 = qkv.to(DEV)
    K.rope_kv_write(g2, gk, gv, cos.to(DEV), sin.to(DEV), B, S, nh, nkv, hd, cache_len=clen.to(DEV),
                    positions=pos.reshape(-1).to(DEV), q_out=q_out)
    assert rel_err(q_out.view(B, S, nh, hd), want_qkv.view(B, S, -1, hd)[:, :, :nh]) < 1e-2
    assert torch.equal(g2.cpu(), qkv)                                   # untouched when q_out is given


@pytest.mark.parametrize("interleaved", [True, False])
def test_swiglu(interleaved):
    x = rnd(9, 2 * 768, seed=5)
    assert rel_err(K.swiglu(x.to(DEV), interleaved), K.swiglu(x.float(), interleaved)) < 1e-2


@pytest.mark.parametrize("M", [1, 2, 3, 4, 8])
@pytest.mark.parametrize("N,K_", [(512, 256), (1536, 4096), (4096, 14336), (1000, 520)])
def test_gemv_modes(M, N, K_):
    if M * K_ * 2 > 220 * 1024:
        pytest.skip("activation tile exceeds shared memory; this shape is served by the tcgen05 GEMM")
    x, w = rnd(M, K_, seed=6), rnd(N, K_, scale=0.05, seed=7)
    g = rnd(K_, seed=8) * 0.1 + 1
    res = rnd(M, N, seed=9)
    xd, wd = x.to(DEV), w.to(DEV)
    for norm in (None, g):
        nd = None if norm is None else norm.to(DEV)
        want = K.linear(x, w, norm_w=norm, eps=1e-5)
        assert rel_err(K.linear(xd, wd, norm_w=nd, eps=1e-5, force="gemv"), want) < 1.5e-2
    want = K.linear(x, w, residual=res, mode=K.RESIDUAL)
    assert rel_err(K.linear(xd, wd, residual=res.to(DEV), mode=K.RESIDUAL, force="gemv"), want) < 1.5e-2
    want = K.linear(x, w, norm_w=g, mode=K.SWIGLU)
    assert rel_err(K.linear(xd, wd, norm_w=g.to(DEV), mode=K.SWIGLU, force="gemv"), want) < 2e-2
    want = K.linear(x, w, out_f32=True)
    got = K.linear(xd, wd, out_f32=True, force="gemv")
    assert got.dtype == torch.float32 and rel_err(got, want) < 1e-3


@pytest.mark.parametrize("M", [1, 2, 3, 4])
@pytest.mark.parametrize("N,K_", [(512, 256), (1536, 4096), (4096, 14336), (1000, 520), (28672, 4096), (6144, 4096)])
def test_gemv_tma_modes(M, N, K_):
    x, w = rnd(M, K_, seed=6), rnd(N, K_, scale=0.05, seed=7)
    g = rnd(K_, seed=8) * 0.1 + 1
    res = rnd(M, N, seed=9)
    xd, wd = x.to(DEV), w.to(DEV)
    for norm in (None, g):
        nd = None if norm is None else norm.to(DEV)
        want = K.linear(x, w, norm_w=norm, eps=1e-5)
        assert rel_err(K.linear(xd, wd, norm_w=nd, eps=1e-5, force="gemv_tma"), want) < 1.5e-2
    want = K.linear(x, w, residual=res, mode=K.RESIDUAL)
    r = res.to(DEV).clone()
    assert rel_err(K.linear(xd, wd, residual=r, mode=K.RESIDUAL, out=r, force="gemv_tma"), want) < 1.5e-2
    want = K.linear(x, w, norm_w=g, mode=K.SWIGLU)
    assert rel_err(K.linear(xd, wd, norm_w=g.to(DEV), mode=K.SWIGLU, force="gemv_tma"), want) < 2e-2
    want = K.linear(x, w, out_f32=True)
    got = K.linear(xd, wd, out_f32=True, force="gemv_tma")
    assert got.dtype == torch.float32 and rel_err(got, want) < 1e-3


@pytest.mark.parametrize("M,N,K_", [
    (128, 256, 64), (128, 128, 256), (256, 512, 512),        # single / few tiles, normal orientation... (M>128)
    (300, 1000, 520), (2048, 6144, 4096), (1024, 4096, 14336),
    (16, 256, 256), (9, 512, 4096), (32, 1536, 4096), (64, 4096, 1024), (100, 1000, 520), (128, 28672, 4096)])
def test_gemm_tcgen05(M, N, K_):
    x, w = rnd(M, K_, seed=10), rnd(N, K_, scale=0.05, seed=11)
    want = x.float() @ w.float().t()
    got = K.linear(x.to(DEV), w.to(DEV), force="gemm")
    assert_close(got, want, 1e-2)
    got32 = K.linear(x.to(DEV), w.to(DEV), out_f32=True, force="gemm")
    assert_close(got32, want, 2e-3)


@pytest.mark.parametrize("M,N,K_", [(256, 256, 64), (256, 512, 512), (300, 256, 128), (2048, 6144, 4096),
                                    (1000, 4096, 14336), (4096, 1024, 2048), (8192, 256, 256)])
@pytest.mark.parametrize("mode", ["plain", "residual", "swiglu"])
def test_gemm_cta_pair(M, N, K_, mode, monkeypatch):
    """tcgen05.mma.cta_group::2 kernel (256 x 256 tiles on 2-CTA clusters, TMA-store epilogue) against fp32, including
    a ragged last row block, many tiles per pair and tiles that wrap the ring several times; and bit-equality with the
    single-CTA kernel (same MMA order over K per output element)."""
    if mode != "plain" and K_ > 8192:
        pytest.skip("one big-K case per mode is enough")
    x, w = rnd(M, K_, seed=70), rnd(N, K_, scale=0.05, seed=71)
    xd, wd = x.to(DEV), w.to(DEV)
    if mode == "residual":
        res = rnd(M, N, seed=72)
        want = K.linear(x, w, residual=res, mode=K.RESIDUAL)
        got = K.linear(xd, wd, residual=res.to(DEV), mode=K.RESIDUAL, force="gemm_2cta")
        r2 = res.to(DEV).clone()
        inplace = K.linear(xd, wd, residual=r2, mode=K.RESIDUAL, out=r2, force="gemm_2cta")      # out aliases the residual
        assert torch.equal(got, inplace)
    elif mode == "swiglu":
        want = K.linear(x, w, mode=K.SWIGLU)
        got = K.linear(xd, wd, mode=K.SWIGLU, force="gemm_2cta")
    else:
        want = x.float() @ w.float().t()
        got = K.linear(xd, wd, force="gemm_2cta")
    assert_close(got, want, 2e-2 if mode == "swiglu" else 1.5e-2)
    again = K.linear(xd, wd, force="gemm_2cta") if mode == "plain" else None
    if again is not None:
        assert torch.equal(got, again)
        monkeypatch.setenv("JLLAMA_GEMM_2CTA", "0")


@pytest.mark.parametrize("M,N,K_,mode", [(16, 4096, 14336, "residual"), (64, 4096, 4096, "plain"),
                                         (9, 6144, 4096, "f32"), (100, 2048, 2048, "swiglu")])
def test_gemm_small_batch_split_k(M, N, K_, mode):
    """Few weight-row tiles → the K range is split over the idle SMs; partials are summed in split order by
    the last arriving split (tickets re-arm themselves), so repeated calls agree bit for bit."""
    x, w = rnd(M, K_, seed=60), rnd(N, K_, scale=0.03, seed=61)
    xd, wd = x.to(DEV), w.to(DEV)
    if mode == "residual":
        res = rnd(M, N, seed=62)
        want = K.linear(x, w, residual=res, mode=K.RESIDUAL)
        outs = [K.linear(xd, wd, residual=res.to(DEV), mode=K.RESIDUAL, force="gemm") for _ in range(3)]
    elif mode == "swiglu":
        want = K.linear(x, w, mode=K.SWIGLU)
        outs = [K.linear(xd, wd, mode=K.SWIGLU, force="gemm") for _ in range(3)]
    else:
        want = x.float() @ w.float().t()
        outs = [K.linear(xd, wd, out_f32=(mode == "f32"), force="gemm") for _ in range(3)]
    assert_close(outs[0], want, 2e-2)
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[1], outs[2])


@pytest.mark.parametrize("M", [24, 384])
def test_gemm_epilogues(M):
    N, K_ = 1024, 512
    x, w, res = rnd(M, K_, seed=12), rnd(N, K_, scale=0.05, seed=13), rnd(M, N, seed=14)
    xd, wd = x.to(DEV), w.to(DEV)
    want = K.linear(x, w, residual=res, mode=K.RESIDUAL)
    r = res.to(DEV).clone()
    got = K.linear(xd, wd, residual=r, mode=K.RESIDUAL, out=r, force="gemm")      # in place on the residual
    assert_close(got, want, 1.5e-2)
    want = K.linear(x, w, mode=K.SWIGLU)
    assert_close(K.linear(xd, wd, mode=K.SWIGLU, force="gemm"), want, 2e-2)
    g = rnd(K_, seed=15) * 0.1 + 1
    want = K.linear(x, w, norm_w=g)
    assert_close(K.linear(xd, wd, norm_w=g.to(DEV), force="gemm"), want, 1.5e-2)


@pytest.mark.parametrize("nh,nkv,hd", [(8, 8, 128), (8, 2, 128), (8, 1, 128), (4, 2, 64), (16, 8, 128)])
@pytest.mark.parametrize("B,length,splits", [(1, 1, 1), (2, 77, 1), (2, 700, 4), (3, 2049, 7)])
def test_decode_attention(nh, nkv, hd, B, length, splits):
    T = length + 5
    q = rnd(B, nh, hd, seed=16)
    kc, vc = rnd(B, nkv, T, hd, seed=17), rnd(B, nkv, T, hd, seed=18)
    pad = torch.tensor([0, 3, 0][:B], dtype=torch.int32).clamp(max=length - 1)
    clen = torch.tensor([length - 1], dtype=torch.int32)
    want = K.decode_attn(q.float(), kc.float(), vc.float(), clen, pad)
    scratch = torch.empty(B * nh * splits * (hd + 2), dtype=torch.float32, device=DEV) if splits > 1 else None
    got = K.decode_attn(q.to(DEV), kc.to(DEV), vc.to(DEV), clen.to(DEV), pad.to(DEV), splits=splits, scratch=scratch)
    assert rel_err(got, want) < 1.5e-2
    # q living inside a wider buffer (batch pitch > nh*hd)
    wide = torch.zeros(B, nh + 2 * nkv, hd, dtype=torch.bfloat16)
    wide[:, :nh] = q
    got2 = K.decode_attn(wide.to(DEV)[:, :nh], kc.to(DEV), vc.to(DEV), clen.to(DEV), pad.to(DEV), splits=splits,
                         scratch=scratch)
    assert rel_err(got2, want) < 1.5e-2


@pytest.mark.parametrize("B,S,start,nh,nkv,pads", [
    (1, 128, 0, 2, 2, [0]), (2, 300, 0, 8, 2, [0, 37]), (1, 200, 128, 4, 1, [5]), (1, 2048, 0, 4, 4, [0]),
    (2, 130, 0, 2, 1, [129, 0])])
def test_flash_prefill(B, S, start, nh, nkv, pads):
    hd, T = 128, start + S + 3
    q = rnd(B, S, nh, hd, seed=20)
    kc, vc = rnd(B, nkv, T, hd, seed=21), rnd(B, nkv, T, hd, seed=22)
    pad = torch.tensor(pads, dtype=torch.int32)
    want = K.prefill_attn_ref(q.float(), kc.float(), vc.float(), start, S, pad)
    got = K.prefill_attn(q.to(DEV), kc.to(DEV), vc.to(DEV), start, S, pad.to(DEV))
    # rows that are themselves padding (slot < pad_left) are don't-care
    slots = start + torch.arange(S)
    valid = (slots[None, :] >= pad[:, None].long())
    g, w = got.float().cpu()[valid], want[valid]                      # [tokens, nh, hd]
    assert_close(g.reshape(-1, hd), w.reshape(-1, hd), 2e-2)


@pytest.mark.parametrize("hd", [64, 128])
@pytest.mark.parametrize("B,S,start,nh,nkv,pads", [
    (1, 128, 0, 2, 2, [0]), (2, 300, 0, 8, 2, [0, 37]), (1, 200, 128, 4, 1, [5]), (1, 1000, 0, 4, 4, [0]),
    (2, 130, 0, 2, 1, [129, 0]), (3, 17, 40, 8, 8, [0, 3, 44])])
def test_prefill_attention_warp_mma(hd, B, S, start, nh, nkv, pads):
    """The warp-level tensor-core prefill kernel (the path of 64-wide heads; pinned here at 128 too): causal mask,
    left padding, continuation from a filled cache, ragged row blocks; the cache beyond the visible keys holds NaN —
    nothing outside [pad_left, start + row] may leak into a row."""
    T = start + S + 5
    q = rnd(B, S, nh, hd, seed=23)
    kc, vc = rnd(B, nkv, T, hd, seed=24), rnd(B, nkv, T, hd, seed=25)
    pad = torch.tensor(pads, dtype=torch.int32)
    want = K.prefill_attn_ref(q.float(), kc.float(), vc.float(), start, S, pad)
    kd, vd = kc.to(DEV), vc.to(DEV)
    kd[:, :, start + S:] = float("nan")
    vd[:, :, start + S:] = float("nan")
    got = K.prefill_attn(q.to(DEV), kd, vd, start, S, pad.to(DEV), force="mma")
    slots = start + torch.arange(S)
    valid = (slots[None, :] >= pad[:, None].long())
    g, w = got.float().cpu()[valid], want[valid]
    assert torch.isfinite(g).all()
    assert_close(g.reshape(-1, hd), w.reshape(-1, hd), 2e-2)
    if hd == 64:                                         # the default entry point routes 64-wide heads to the same kernel
        got2 = K.prefill_attn(q.to(DEV), kd, vd, start, S, pad.to(DEV))
        assert torch.equal(got2.cpu()[valid], got.cpu()[valid])


def test_flash_prefill_many_ctas_repeated():
    """Thousands of CTAs, two per SM, launched back to back with cold caches: the barrier-phase discipline
    between the softmax warps and the MMA warp must hold when tile loads are slow (a missing guard showed up
    as an intermittent hang at batch 8)."""
    B, S, nh, nkv, hd = 8, 1024, 16, 4, 128
    q = rnd(B, S, nh, hd, seed=50).to(DEV)
    kc, vc = rnd(B, nkv, S, hd, seed=51).to(DEV), rnd(B, nkv, S, hd, seed=52).to(DEV)
    pad = torch.zeros(B, dtype=torch.int32, device=DEV)
    flush = torch.empty(192 << 20, dtype=torch.uint8, device=DEV)
    first = None
    for it in range(25):
        flush.zero_()
        out = K.prefill_attn(q, kc, vc, 0, S, pad)
        torch.cuda.synchronize()
        if first is None:
            first = out.clone()
            want = K.prefill_attn_ref(q[:1].float().cpu(), kc[:1].float().cpu(), vc[:1].float().cpu(), 0, S, pad[:1].cpu())
            assert ((first[:1].float().cpu() - want).abs().max() / want.abs().max()).item() < 2e-2
        else:
            assert torch.equal(out, first)


@pytest.mark.parametrize("ramp", [0.02, -0.02])
def test_flash_prefill_growing_and_shrinking_scores(ramp):
    """Scores whose row maximum keeps rising along the keys (ramp > 0) force the in-TMEM rescale of O over
    and over; falling scores (ramp < 0) exercise the 'reference maximum stays put' path with tiny tails."""
    B, S, nh, nkv, hd = 1, 640, 2, 1, 128
    q = rnd(B, S, nh, hd, seed=40)
    base = rnd(B, nkv, S, hd, seed=41).float()
    # key t gets a component along a fixed direction that grows with t; queries all have +3 along it
    direction = torch.zeros(hd)
    direction[:8] = 1.0
    kc = (base * 0.3 + direction * (ramp * torch.arange(S).float())[None, None, :, None] * 6).to(torch.bfloat16)
    q = (q.float() * 0.3 + 3.0 * direction).to(torch.bfloat16)
    vc = rnd(B, nkv, S, hd, seed=42)
    pad = torch.tensor([0], dtype=torch.int32)
    want = K.prefill_attn_ref(q.float(), kc.float(), vc.float(), 0, S, pad)
    got = K.prefill_attn(q.to(DEV), kc.to(DEV), vc.to(DEV), 0, S, pad.to(DEV))
    assert torch.isfinite(got.float()).all()
    assert ((got.float().cpu() - want).abs().max() / (want.abs().max() + 1e-6)).item() < 2e-2


@pytest.fixture
def attn_variant(request):
    K.set_decode_attn_variant(request.param)
    yield request.param
    K.set_decode_attn_variant(0)


@pytest.mark.parametrize("attn_variant", [1, 2], indirect=True, ids=["simt", "mma"])
@pytest.mark.parametrize("nh,nkv,hd", [(8, 2, 128), (8, 8, 128), (8, 1, 128), (4, 2, 64), (4, 2, 128)])
@pytest.mark.parametrize("B,filled,max_len", [(1, 0, 64), (2, 77, 256), (1, 2048, 2304), (3, 700, 4096)])
def test_decode_attention_fused(nh, nkv, hd, B, filled, max_len, attn_variant):
    """RoPE + append + attention + merge in one launch == the three separate CPU statements; both kernels behind the
    entry point (SIMT; tensor-core for head_dim 128 — the 64-wide case then stays on the SIMT kernel)."""
    T = max_len
    qkv = rnd(B, (nh + 2 * nkv) * hd, seed=30)
    kc, vc = rnd(B, nkv, T, hd, seed=31), rnd(B, nkv, T, hd, seed=32)
    cos, sin = ref.rope_cos_sin(hd, 2 * max_len, 500000.0)
    pad = torch.tensor([0, 3, 1][:B], dtype=torch.int32).clamp(max=max(filled - 1, 0))
    clen = torch.tensor([filled], dtype=torch.int32)
    kc_ref, vc_ref = kc.float(), vc.float()
    want = K.decode_attn_fused(qkv.float(), kc_ref, vc_ref, cos, sin, clen, pad, nh, nkv, hd)
    splits, scratch = K.decode_attn_fused_plan(B, nh, nkv, hd, max_len, torch.device(DEV))
    gk, gv = kc.to(DEV), vc.to(DEV)
    for _ in range(2):                                   # twice: the ticket counters must re-arm themselves
        got = K.decode_attn_fused(qkv.to(DEV), gk, gv, cos.to(DEV), sin.to(DEV), clen.to(DEV), pad.to(DEV), nh, nkv,
                                  hd, splits, scratch)
        assert rel_err(got, want) < 1.5e-2
    assert rel_err(gk[:, :, filled], kc_ref[:, :, filled]) < 1e-2 and rel_err(gv[:, :, filled], vc_ref[:, :, filled]) < 1e-2
    assert torch.equal(gk[:, :, :filled].cpu(), kc[:, :, :filled]) and (scratch[-B * nkv:] == 0).all()


@pytest.mark.parametrize("nh,nkv", [(32, 8), (8, 1), (8, 8)])
def test_decode_attention_fused_batched_shape_picks_tensor_core_kernel(nh, nkv):
    """Batch 16-64 shapes (batch x kv heads >= 16 → the tensor-core kernel by default): rows with different left
    padding, a context that is not a multiple of the 128-key chunk, per-head / per-column error bounds."""
    B, hd, filled, T = 64 // nkv + 8, 128, 1111, 1280
    qkv = rnd(B, (nh + 2 * nkv) * hd, seed=33)
    kc, vc = rnd(B, nkv, T, hd, seed=34), rnd(B, nkv, T, hd, seed=35)
    cos, sin = ref.rope_cos_sin(hd, 2 * T, 500000.0)
    pad = (torch.arange(B, dtype=torch.int32) * 37) % 600
    clen = torch.tensor([filled], dtype=torch.int32)
    want = K.decode_attn_fused(qkv.float(), kc.float(), vc.float(), cos, sin, clen, pad, nh, nkv, hd)
    splits, scratch = K.decode_attn_fused_plan(B, nh, nkv, hd, T, torch.device(DEV))
    outs = []
    for variant in (0, 1):
        K.set_decode_attn_variant(variant)
        gk, gv = kc.to(DEV), vc.to(DEV)
        got = K.decode_attn_fused(qkv.to(DEV), gk, gv, cos.to(DEV), sin.to(DEV), clen.to(DEV), pad.to(DEV), nh, nkv,
                                  hd, splits, scratch)
        K.set_decode_attn_variant(0)
        assert_close(got.view(B * nh, hd), want.view(B * nh, hd), 1.5e-2)
        outs.append((got.float().cpu(), gk[:, :, filled].cpu(), gv[:, :, filled].cpu()))
    # the two kernels write the same cache slot bit for bit and agree on the output within bf16 noise
    assert torch.equal(outs[0][1], outs[1][1]) and torch.equal(outs[0][2], outs[1][2])
    assert rel_err(outs[0][0], outs[1][0]) < 1e-2


def test_decode_attention_plan():
    splits, scratch = K.decode_attn_plan(1, 32, 8, 128, 2304, torch.device(DEV))
    assert 1 <= splits <= 64 and (scratch is None) == (splits == 1)


def test_argmax_and_topk():
    g = torch.Generator().manual_seed(19)
    logits = torch.randn(3, 128256, generator=g)
    logits[1, 77] = logits[1, 99] = 50.0                      # tie → lowest index
    d = logits.to(DEV)
    assert K.argmax(d).cpu().tolist() == [int(logits[0].argmax()), 77, int(logits[2].argmax())]
    for k in (1, 50, 64):
        v, i = K.topk_candidates(d, k, idx_offset=1000)
        wv, wi = torch.topk(logits, k, dim=-1)
        assert torch.equal(v.cpu(), wv)
        assert torch.equal(torch.sort(i.cpu() - 1000, -1).values[0], torch.sort(wi, -1).values[0].to(torch.int32))
    v, i = K.topk_candidates(d[:, :40].contiguous(), 50)     # fewer entries than k → padded with -inf / -1
    assert (i[:, 40:] == -1).all() and torch.isinf(v[:, 40:]).all()


def test_topk_with_thousands_of_ties_is_exact_and_repeatable():
    """More elements tie with the selection bound than the survivor list holds (constant row; 4096 tied maxima; a
    plateau below a few distinct maxima): the k largest are still exact, ties go to the lower index, and two runs
    agree (the survivor list used to keep whichever tied elements arrived first)."""
    V, k = 128256, 50
    g = torch.Generator().manual_seed(23)
    rows = torch.randn(4, V, generator=g)
    rows[0] = 0.0                                                        # constant row
    tied = torch.randperm(V, generator=g)[:4096]
    rows[1, tied] = 9.0                                                  # 4096 tied maxima
    rows[2] = -1.0
    rows[2, [5, 70000, 128255]] = torch.tensor([3.0, 2.0, 1.0])          # 3 distinct maxima over a plateau
    rows[3, :3000] = float("-inf")                                       # ordinary row with a masked prefix
    d = rows.to(DEV)
    v1, i1 = K.topk_candidates(d, k)
    v2, i2 = K.topk_candidates(d, k)
    assert torch.equal(v1, v2) and torch.equal(i1, i2)
    v, i = v1.cpu(), i1.cpu().long()
    assert (v[0] == 0).all() and i[0].tolist() == list(range(k))
    assert (v[1] == 9).all() and i[1].tolist() == sorted(tied.tolist())[:k]
    assert v[2, :3].tolist() == [3.0, 2.0, 1.0] and i[2, :3].tolist() == [5, 70000, 128255]
    assert (v[2, 3:] == -1).all() and i[2, 3:].tolist() == [j for j in range(k) if j != 5][:k - 3]
    wv, wi = torch.topk(rows[3], k)
    assert torch.equal(v[3], wv) and torch.equal(i[3], wi)


def test_sampling_distribution_and_greedy():
    p = torch.tensor([0.5, 0.25, 0.15, 0.07, 0.03])
    B = 4096
    logits = torch.log(p)[None].repeat(B, 1).contiguous()
    v, i = K.topk_candidates(logits.to(DEV), 5)
    step = torch.zeros(1, dtype=torch.int32, device=DEV)
    tok = K.sample_candidates(v, i, top_k=5, top_p=1.0, temperature=1.0, seed=1, step=step)
    freq = torch.bincount(tok.cpu().long(), minlength=5).float() / B
    assert torch.allclose(freq, p, atol=0.03)
    tok = K.sample_candidates(v, i, top_k=5, top_p=0.7, temperature=1.0, seed=2, step=step)   # keeps {0,1}
    freq = torch.bincount(tok.cpu().long(), minlength=5).float() / B
    assert freq[2:].sum() == 0 and abs(freq[0] - 2 / 3) < 0.03
    tok = K.sample_candidates(v, i, top_k=3, top_p=1.0, temperature=0.5, seed=3, step=step)
    freq = torch.bincount(tok.cpu().long(), minlength=5).float() / B
    want = torch.tensor([0.25, 0.0625, 0.0225]) / 0.335
    assert freq[3:].sum() == 0 and torch.allclose(freq[:3], want, atol=0.03)
    a = K.sample_candidates(v, i, top_k=5, seed=7, step=step)
    b = K.sample_candidates(v, i, top_k=5, seed=7, step=step)
    assert torch.equal(a, b)                                                  # counter-based RNG: repeatable
    step.fill_(1)
    assert not torch.equal(a, K.sample_candidates(v, i, top_k=5, seed=7, step=step))
    assert (K.sample_candidates(v, i, top_k=5, temperature=0.0) == 0).all()   # greedy
    # cross-rank merge: two vocab shards' candidate lists give the same answer as the full row
    full = torch.randn(2, 1000, generator=torch.Generator().manual_seed(5))
    va, ia = K.topk_candidates(full[:, :500].contiguous().to(DEV), 8, idx_offset=0)
    vb, ib = K.topk_candidates(full[:, 500:].contiguous().to(DEV), 8, idx_offset=500)
    vals, idx = torch.stack([va, vb]), torch.stack([ia, ib])
    tok = K.sample_candidates(vals, idx, n_lists=2, top_k=8, temperature=0.0)
    assert tok.cpu().tolist() == full.argmax(-1).tolist()
