"""Ulysses sequence parallelism on gloo: sharded attention == full attention; SP cross entropy."""
import torch

from conftest import run_dist


def _attn(q, k, v):     # [s, b, h, d] full-sequence causal attention
    s, b, h, d = q.shape
    hk = k.shape[2]
    qf = q.permute(1, 2, 0, 3).float()
    kf = k.permute(1, 2, 0, 3).float().repeat_interleave(h // hk, 1)
    vf = v.permute(1, 2, 0, 3).float().repeat_interleave(h // hk, 1)
    sc = (qf @ kf.transpose(-1, -2)) / d ** 0.5
    sc = sc.masked_fill(~torch.ones(s, s, dtype=torch.bool).tril(), float("-inf"))
    return (sc.softmax(-1) @ vf).permute(2, 0, 1, 3)


def _ulysses_worker(rank, world, hq, hk):
    from shuffle_exchange_b200.sequence.layer import DistributedAttention
    torch.manual_seed(0)
    S, B, D = 16, 2, 8
    q = torch.randn(S, B, hq, D, requires_grad=True)
    k = torch.randn(S, B, hk, D, requires_grad=True)
    v = torch.randn(S, B, hk, D, requires_grad=True)
    ref = _attn(q, k, v)
    g = torch.randn_like(ref)
    ref.backward(g)
    sl = slice(rank * S // world, (rank + 1) * S // world)
    ql, kl, vl = (t.detach()[sl].clone().requires_grad_() for t in (q, k, v))
    da = DistributedAttention(_attn, torch.distributed.group.WORLD, scatter_idx=2, gather_idx=0)
    out = da(ql, kl, vl)
    assert torch.allclose(out, ref[sl].detach(), atol=1e-5), (out - ref[sl]).abs().max()
    out.backward(g[sl])
    assert torch.allclose(ql.grad, q.grad[sl], atol=1e-5)
    assert torch.allclose(kl.grad, k.grad[sl], atol=1e-5)
    assert torch.allclose(vl.grad, v.grad[sl], atol=1e-5)


def test_ulysses_equal_heads():
    run_dist(_ulysses_worker, 2, (4, 4))


def test_ulysses_gqa_heads():
    run_dist(_ulysses_worker, 2, (4, 2))


def test_ulysses_uneven_heads():
    run_dist(_ulysses_worker, 2, (3, 3))


def _ce_worker(rank, world):
    from shuffle_exchange_b200.sequence.cross_entropy import vocab_sequence_parallel_cross_entropy
    torch.manual_seed(0)
    S, B, V = 8, 2, 11
    logits = torch.randn(S, B, V, requires_grad=True)
    tgt = torch.randint(0, V, (S, B))
    ref = torch.nn.functional.cross_entropy(logits.view(-1, V), tgt.view(-1), reduction="none").view(S, B)
    ref.sum().backward()
    sl = slice(rank * S // world, (rank + 1) * S // world)
    ll = logits.detach()[sl].clone().requires_grad_()
    loss = vocab_sequence_parallel_cross_entropy(ll, tgt[sl], torch.distributed.group.WORLD)
    assert loss.shape == (S, B) and torch.allclose(loss, ref.detach(), atol=1e-5)
    loss.sum().backward()
    assert torch.allclose(ll.grad, logits.grad[sl], atol=1e-5)


def test_sp_cross_entropy():
    run_dist(_ce_worker, 2)
