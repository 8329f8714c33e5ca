"""ALST (Ulysses SP for HF-style models, tiled compute) and FPDT (chunked attention / FFN / logits-loss) on CPU."""
from types import SimpleNamespace

import pytest
import torch

from conftest import run_dist


# ---------------------------------------------------------------------------------------------- FPDT
def test_fpdt_chunked_attention_matches_full_causal():
    from shuffle_exchange_b200.ops import functional as F
    from shuffle_exchange_b200.sequence.fpdt_layer import fpdt_chunked_attention
    torch.manual_seed(0)
    B, H, Hk, D, n, Lc = 2, 4, 2, 16, 4, 8
    q = torch.randn(B, n * Lc, H, D, requires_grad=True)
    k = torch.randn(B, n * Lc, Hk, D, requires_grad=True)
    v = torch.randn(B, n * Lc, Hk, D, requires_grad=True)
    ref = F.attention_ref(q, k, v, True)
    out = torch.cat(fpdt_chunked_attention(list(q.split(Lc, 1)), list(k.split(Lc, 1)), list(v.split(Lc, 1)),
                                           D ** -0.5), 1)
    assert torch.allclose(out, ref, atol=1e-5)
    g = torch.randn_like(ref)
    for a, b in zip(torch.autograd.grad(out, (q, k, v), g), torch.autograd.grad(ref, (q, k, v), g)):
        assert torch.allclose(a, b, atol=1e-5)


def test_fpdt_ffn_and_logits_loss():
    from shuffle_exchange_b200.sequence.fpdt_layer import FPDT_FFN, FPDT_LogitsLoss, bias_gelu
    torch.manual_seed(0)
    w1 = (torch.randn(128, 64) * 0.1).requires_grad_(True)
    b1 = (torch.randn(128) * 0.1).requires_grad_(True)
    w2 = (torch.randn(64, 128) * 0.1).requires_grad_(True)
    b2 = (torch.randn(64) * 0.1).requires_grad_(True)
    x = torch.randn(32, 2, 64, requires_grad=True)
    y, rb = FPDT_FFN.apply(x, w1, b1, w2, b2, True, 8)
    ref = bias_gelu(x @ w1.t() + b1) @ w2.t() + b2
    assert rb is None and torch.allclose(y, ref, atol=1e-5)
    g = torch.randn_like(y)
    for a, b in zip(torch.autograd.grad(y, (x, w1, b1, w2, b2), g), torch.autograd.grad(ref, (x, w1, b1, w2, b2), g)):
        assert torch.allclose(a, b, atol=2e-5)
    y2, rb2 = FPDT_FFN.apply(x, w1, b1, w2, b2, False, 16)
    assert torch.equal(rb2, b2) and torch.allclose(y2 + b2, ref, atol=1e-5)
    V = 50
    lw = (torch.randn(V, 64) * 0.1).requires_grad_(True)
    labels = torch.randint(0, V, (2, 32))
    loss = FPDT_LogitsLoss.apply(x, labels, lw, 0, 1, None, 4)
    rl = torch.nn.functional.cross_entropy((x @ lw.t()).reshape(-1, V), labels.t().reshape(-1),
                                           reduction="none").view(32, 2)
    assert torch.allclose(loss, rl, atol=1e-5)
    gl = torch.randn_like(loss)
    for a, b in zip(torch.autograd.grad(loss, (x, lw), gl), torch.autograd.grad(rl, (x, lw), gl)):
        assert torch.allclose(a, b, atol=1e-5)


def test_fpdt_input_construct_is_load_balanced_permutation():
    from shuffle_exchange_b200.sequence.fpdt_layer import FPDT_InputConstruct
    args = SimpleNamespace(ds_sequence_parallel_fpdt_chunk_size=16)
    tok = torch.arange(64)[None].repeat(2, 1)
    seen = []
    for r in range(2):
        t, l, lm, am, pos = FPDT_InputConstruct(tok, tok, torch.ones(2, 64), None, tok, args, 2, r).generate()
        assert t.shape == (2, 32) and torch.equal(t, l) and torch.equal(t, pos) and lm.shape == (2, 64)
        seen.append(t[0])
        # chunks of 8 tokens, every other chunk
        assert t[0, :8].tolist() == list(range(8 * r, 8 * r + 8))
        assert t[0, 8:16].tolist() == list(range(16 + 8 * r, 24 + 8 * r))
    assert sorted(torch.cat(seen).tolist()) == list(range(64))


def _fpdt_attention_rank(rank, world):
    import torch.distributed as dist
    from shuffle_exchange_b200.ops import functional as F
    from shuffle_exchange_b200.sequence.fpdt_layer import FPDT_Attention, FPDT_InputConstruct
    torch.manual_seed(0)
    cfg = SimpleNamespace(kv_channels=16, num_attention_heads=4, num_key_value_heads=2, hidden_size=64,
                          attention_dropout=0.0)
    S, b = 64, 2
    W1 = (torch.randn(64 + 2 * 32, 64) * 0.1).requires_grad_(True)
    b1 = torch.zeros(128, requires_grad=True)
    W2 = (torch.randn(64, 64) * 0.1).requires_grad_(True)
    b2 = torch.zeros(64, requires_grad=True)
    x_full = torch.randn(S, b, 64)
    # reference: plain causal attention over the whole sequence
    xr = x_full.clone().requires_grad_(True)
    qkv = xr @ W1.t() + b1
    q = qkv[..., :64].reshape(S, b, 4, 16).permute(1, 0, 2, 3)
    k = qkv[..., 64:96].reshape(S, b, 2, 16).permute(1, 0, 2, 3)
    v = qkv[..., 96:].reshape(S, b, 2, 16).permute(1, 0, 2, 3)
    ref = F.attention_ref(q, k, v, True).flatten(2).permute(1, 0, 2) @ W2.t()
    # FPDT: load-balanced shard of the sequence on this rank, global chunk = 16 tokens
    args = SimpleNamespace(ds_sequence_parallel_fpdt_chunk_size=16)
    idx = torch.arange(S)[None]
    mine = FPDT_InputConstruct(idx, None, None, None, None, args, world, rank).generate()[0][0]
    xl = x_full[mine].clone().requires_grad_(True)
    att = FPDT_Attention(cfg, W1, b1, W2, b2, dist.group.WORLD, chunk_size=16, enable_offloading=False)
    y, bias = att(xl, None, None, None)
    assert torch.equal(bias, b2)
    assert torch.allclose(y, ref[mine].detach(), atol=1e-5), (y - ref[mine]).abs().max()
    g = torch.randn(S, b, 64)
    W1.grad = W2.grad = None
    y.backward(g[mine])
    gw1 = W1.grad.clone()
    dist.all_reduce(gw1)
    W1.grad = None
    ref.backward(g)
    assert torch.allclose(xl.grad, xr.grad[mine], atol=1e-5)
    assert torch.allclose(gw1, W1.grad, atol=1e-4)


def test_fpdt_attention_two_ranks():
    run_dist(_fpdt_attention_rank, world=2)


# ---------------------------------------------------------------------------------------------- ALST tiled compute
def test_sequence_tiled_compute_and_tiled_mlp():
    from shuffle_exchange_b200.runtime.sequence_parallel import TiledMLP, sequence_tiled_compute
    torch.manual_seed(0)
    mlp = torch.nn.Sequential(torch.nn.Linear(16, 64), torch.nn.GELU(), torch.nn.Linear(64, 16))
    x = torch.randn(2, 10, 16, requires_grad=True)
    ref = mlp(x)
    g = torch.randn_like(ref)
    ref.backward(g)
    gx_ref, gw_ref = x.grad.clone(), mlp[0].weight.grad.clone()
    x.grad = None
    mlp.zero_grad()
    out = TiledMLP.apply(lambda m, t: m(t), mlp, x, 3, list(mlp.parameters()))
    assert torch.allclose(out, ref, atol=1e-6)
    out.backward(g)
    assert torch.allclose(x.grad, gx_ref, atol=1e-6) and torch.allclose(mlp[0].weight.grad, gw_ref, atol=1e-5)
    # generic version, un-reduced and mean-reduced
    x.grad = None
    mlp.zero_grad()
    out = sequence_tiled_compute(lambda hidden, scale: mlp(hidden) * scale, 10, 4, {"hidden": x}, {"scale": 2.0},
                                 "hidden", list(mlp.parameters()), output_unshard_dimension=1, output_reduction=None)
    assert torch.allclose(out, ref * 2, atol=1e-6)
    out.backward(g)
    assert torch.allclose(x.grad, gx_ref * 2, atol=1e-5)
    x.grad = None
    mlp.zero_grad()
    m = sequence_tiled_compute(lambda hidden: mlp(hidden), 10, 5, {"hidden": x}, {}, "hidden",
                               list(mlp.parameters()), output_reduction="mean")
    m.backward()
    xr = x.detach().clone().requires_grad_(True)
    mlp(xr).mean().backward()
    assert torch.allclose(m, mlp(xr).mean(), atol=1e-6) and torch.allclose(x.grad, xr.grad, atol=1e-6)


def test_tiled_loss_is_sft_aware():
    from shuffle_exchange_b200.runtime.sequence_parallel import TiledFusedLogitsLoss

    def loss_fn(logits, labels, vocab_size, shift_labels):
        return torch.nn.functional.cross_entropy(logits.reshape(-1, vocab_size).float(), shift_labels.reshape(-1),
                                                 ignore_index=-100)
    torch.manual_seed(0)
    logits = torch.randn(1, 12, 20, requires_grad=True)
    labels = torch.randint(0, 20, (1, 12))
    labels[:, :5] = -100           # the first tile (4 tokens) is fully masked
    ref = loss_fn(logits, None, 20, labels)
    out = TiledFusedLogitsLoss.apply(loss_fn, logits, 20, labels, 3)
    assert torch.allclose(out, ref, atol=1e-6)
    (g,) = torch.autograd.grad(out, logits)
    (gr,) = torch.autograd.grad(ref, logits)
    assert torch.allclose(g, gr, atol=1e-6)
    allmasked = torch.full((1, 12), -100)
    z = TiledFusedLogitsLoss.apply(loss_fn, logits, 20, allmasked, 3)
    assert float(z) == 0.0


# ---------------------------------------------------------------------------------------------- ALST attention + loader
def _ulysses_hf_rank(rank, world):
    import torch.distributed as dist
    from shuffle_exchange_b200.ops import functional as F
    from shuffle_exchange_b200.runtime.sequence_parallel import (UlyssesSPAttentionHF, UlyssesSPDataLoaderAdapter,
                                                                 sxb_flash_core_attention)
    torch.manual_seed(0)
    b, H, Hk, S, D = 2, 4, 1, 16, 8            # kv heads < sp: replicated
    q = torch.randn(b, H, S, D)
    k = torch.randn(b, Hk, S, D)
    v = torch.randn(b, Hk, S, D)
    ref = F.attention_ref(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2), True)       # [b, S, H, D]
    sl = slice(rank * S // world, (rank + 1) * S // world)
    att = UlyssesSPAttentionHF(sxb_flash_core_attention, S // world, S, b, H, D, Hk, 2, dist.group.WORLD)
    ql = q[:, :, sl].clone().requires_grad_(True)
    ctx, w = att(None, ql, k[:, :, sl].contiguous(), v[:, :, sl].contiguous(), None)
    assert w is None and ctx.shape == (b, S // world, H, D)
    assert torch.allclose(ctx, ref[:, sl], atol=1e-5)
    ctx.sum().backward()
    assert ql.grad is not None and ql.grad.abs().sum() > 0
    with pytest.raises(ValueError):
        UlyssesSPAttentionHF(sxb_flash_core_attention, 8, 16, b, 3, D, 1, 2, dist.group.WORLD)
    # dataloader adapter: each rank feeds a different sample; both see shards of rank 0's sample, then rank 1's
    sample = {"input_ids": torch.arange(8)[None] + 100 * rank, "labels": torch.arange(8)[None] + 100 * rank,
              "position_ids": torch.arange(8)[None], "meta": f"r{rank}"}
    dl = UlyssesSPDataLoaderAdapter([sample], rank, dist.group.WORLD, world, torch.device("cpu"))
    assert len(dl) == 2
    first, second = next(dl), next(dl)
    half = slice(rank * 4, rank * 4 + 4)
    assert first["input_ids"].tolist() == (torch.arange(8)[None])[:, half].tolist() and first["meta"] == "r0"
    assert second["input_ids"].tolist() == (torch.arange(8)[None] + 100)[:, half].tolist() and second["meta"] == "r1"
    exp = torch.cat([torch.arange(1, 8), torch.tensor([-100])])[None]
    assert first["shift_labels"].tolist() == exp[:, half].tolist() and "labels" not in first
    with pytest.raises(StopIteration):
        next(dl)


def test_ulysses_hf_attention_and_dataloader_two_ranks():
    run_dist(_ulysses_hf_rank, world=2)
