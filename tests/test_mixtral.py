"""Dropless MoE / Mixtral: single-process numerics against a dense per-token reference, and expert parallelism over 2
gloo ranks against the same model run without EP."""
import torch

from conftest import run_dist


def _dense_moe_reference(moe, x):
    """Every token through its top-k experts, computed token by token in fp32."""
    T, H = x.shape
    probs = torch.softmax((x.float() @ moe.wg.float().t()), dim=-1)
    top_p, top_e = torch.topk(probs, moe.k, dim=-1)
    top_p = top_p / top_p.sum(-1, keepdim=True)
    out = torch.zeros(T, H)
    for t in range(T):
        for j in range(moe.k):
            e = moe.experts[int(top_e[t, j])]
            gu = x[t].float() @ e.w_gate_up.float().t()
            g, u = gu.chunk(2)
            out[t] += top_p[t, j] * ((torch.nn.functional.silu(g) * u) @ e.w_down.float().t())
    return out


def test_dropless_moe_matches_dense_reference():
    from shuffle_exchange_b200.moe.dropless import DroplessMoE
    torch.manual_seed(0)
    moe = DroplessMoE(32, 64, num_experts=4, k=2)
    moe.reset_parameters(0.2)
    x = torch.randn(19, 32, requires_grad=True)
    y, l_aux, counts = moe(x)
    ref = _dense_moe_reference(moe, x.detach())
    assert torch.allclose(y, ref, atol=1e-4, rtol=1e-4)
    assert int(counts.sum()) == 19 * 2 and l_aux.item() > 0
    (y.sum() + l_aux).backward()
    assert x.grad is not None and moe.wg.grad is not None and moe.wg.grad.abs().sum() > 0
    assert all(p.grad is not None for p in moe.experts[int(counts.argmax())].parameters())


def test_mixtral_tiny_trains():
    from shuffle_exchange_b200.models.mixtral import MixtralForCausalLM, mixtral_config
    torch.manual_seed(0)
    cfg = mixtral_config("mixtral-tiny", activation_checkpointing=False)
    m = MixtralForCausalLM(cfg)
    opt = torch.optim.AdamW(m.parameters(), lr=3e-3)
    ids = torch.randint(0, cfg.vocab_size, (2, 32))
    losses = []
    for _ in range(8):
        loss = m(ids, labels=ids)
        opt.zero_grad()
        loss.backward()
        opt.step()
        losses.append(loss.item())
    assert losses[-1] < losses[0] - 0.5, losses


def _ep_worker(rank, world):
    import torch.distributed as td
    from shuffle_exchange_b200.moe.dropless import DroplessMoE
    from shuffle_exchange_b200.utils import groups
    torch.manual_seed(0)
    full = DroplessMoE(16, 32, num_experts=4, k=2)           # identical on both ranks
    full.reset_parameters(0.3)
    ep = DroplessMoE(16, 32, num_experts=4, ep_size=2, k=2)
    ep.wg.data.copy_(full.wg.data)
    for j in range(2):                                        # rank r owns experts 2r, 2r+1
        ep.experts[j].load_state_dict(full.experts[rank * 2 + j].state_dict())
    ep.set_deepspeed_parallelism()
    g = torch.Generator().manual_seed(10 + rank)
    x = torch.randn(13 + rank, 16, generator=g, requires_grad=True)       # different token counts per rank
    y, _, _ = ep(x)
    ref, _, _ = full(x.detach())
    assert torch.allclose(y, ref, atol=1e-5), (y - ref).abs().max()
    y.sum().backward()
    assert x.grad is not None and all(p.grad is not None for p in ep.experts.parameters())


def test_expert_parallel_matches_local():
    run_dist(_ep_worker, 2)
