"""The reference-shaped capacity-factor gates (``moe/sharded_moe.py``: ``top1gating`` / ``top2gating`` / ``topkgating``,
``TopKGate``) against their definitions, and ``MoE`` (capacity path, residual variant) against a dense evaluation."""
import math

import pytest
import torch
import torch.nn.functional as F

from shuffle_exchange_b200.moe import sharded_moe as sm


def _check_slots(dispatch_mask):
    assert (dispatch_mask.sum(dim=0) <= 1).all()                      # one token per (expert, slot)


def test_top1_gating_capacity_order_and_aux_loss():
    torch.manual_seed(0)
    S, E = 32, 4
    logits = torch.randn(S, E)
    logits[:, 0] += 1.5                                               # expert 0 is oversubscribed
    l_aux, cw, dm, counts = sm.top1gating(logits, capacity_factor=1.0, min_capacity=2, use_rts=False)
    gates = F.softmax(logits, 1)
    choice = gates.argmax(1)
    cap = math.ceil(S / E * 1.0)
    assert cw.shape == (S, E, cap) and counts.tolist() == torch.bincount(choice, minlength=E).tolist()
    _check_slots(dm)
    assert (dm.sum(dim=(1, 2)) <= 1).all()                            # top-1: a token sits in at most one slot
    for e in range(E):                                                # first come, first served up to the capacity
        mine = (choice == e).nonzero().flatten().tolist()
        kept, dropped = mine[:cap], mine[cap:]
        for pos, t in enumerate(kept):
            assert dm[t, e, pos] and abs(float(cw[t, e, pos]) - float(gates[t, e])) < 1e-6
        assert not dm[dropped].any()
    assert counts[0] > cap                                            # something really was dropped
    want_aux = (gates.mean(0) * F.one_hot(choice, E).float().mean(0)).sum() * E
    assert abs(float(l_aux) - float(want_aux)) < 1e-6
    # random token selection keeps the same NUMBER of tokens per expert, chosen at random
    _, _, dm_rts, _ = sm.top1gating(logits, 1.0, 2, use_rts=True)
    assert dm_rts.sum(dim=(0, 2)).tolist() == dm.sum(dim=(0, 2)).tolist()
    # no token dropping: capacity grows to the largest expert load
    _, cw_nd, dm_nd, _ = sm.top1gating(logits, 1.0, 2, drop_tokens=False, use_rts=False)
    assert cw_nd.shape[2] == int(counts.max()) and dm_nd.sum() == S
    used = torch.ones(S)
    used[::2] = 0                                                     # padded tokens take no slot
    _, _, dm_u, c_u = sm.top1gating(logits, 4.0, 2, used_token=used, use_rts=False)
    assert not dm_u[::2].any() and int(c_u.sum()) == S // 2


def test_top2_and_topk_gating_weights():
    torch.manual_seed(1)
    S, E = 24, 6
    logits = torch.randn(S, E)
    gates = F.softmax(logits, 1)
    l_aux, cw, dm, counts = sm.top2gating(logits, capacity_factor=4.0, min_capacity=1, top2_2nd_expert_sampling=False)
    _check_slots(dm)
    top2 = gates.topk(2, dim=1)
    for t in range(S):
        e1, e2 = top2.indices[t].tolist()
        w = cw[t].sum(dim=1)                                          # per expert
        g1, g2 = float(gates[t, e1]), float(gates[t, e2])
        assert abs(float(w[e1]) - g1 / (g1 + g2)) < 1e-5 and abs(float(w[e2]) - g2 / (g1 + g2)) < 1e-5
        assert int(dm[t].sum()) == 2
    assert int(counts.sum()) == 2 * S and float(l_aux) > 0
    # tight capacity drops second choices first and never exceeds the buffers
    _, cw_t, dm_t, _ = sm.top2gating(logits, capacity_factor=0.5, min_capacity=1, top2_2nd_expert_sampling=False)
    _check_slots(dm_t)
    assert dm_t.sum() < 2 * S and cw_t.shape[2] == math.ceil(2 * S / E * 0.5)
    # general k
    l3, cw3, dm3, c3 = sm.topkgating(logits, 3, capacity_factor=8.0, min_capacity=1)
    _check_slots(dm3)
    assert (dm3.sum(dim=(1, 2)) == 3).all() and int(c3.sum()) == 3 * S
    top3 = gates.topk(3, dim=1)
    t = 5
    w = cw3[t].sum(dim=1)
    norm = top3.values[t].sum()
    for e, g in zip(top3.indices[t].tolist(), top3.values[t].tolist()):
        assert abs(float(w[e]) - g / float(norm)) < 1e-5
    _, cw_nd, dm_nd, c_nd = sm.topkgating(logits, 2, capacity_factor=0.1, min_capacity=1, drop_tokens=False)
    assert dm_nd.sum() == 2 * S and cw_nd.shape[2] == int(c_nd.max())


class _Expert(torch.nn.Module):
    def __init__(self, d):
        super().__init__()
        self.fc = torch.nn.Linear(d, d)

    def forward(self, x):
        return torch.tanh(self.fc(x))


def _moe_worker(rank, world):
    from shuffle_exchange_b200.moe.layer import MoE
    from shuffle_exchange_b200.utils import groups
    torch.manual_seed(0)
    d, E = 8, 4
    moe = MoE(hidden_size=d, expert=_Expert(d), num_experts=E, ep_size=1, k=2, capacity_factor=8.0, min_capacity=1,
              use_rts=False, top2_2nd_expert_sampling=False)
    moe.set_deepspeed_parallelism()
    x = torch.randn(3, 5, d, requires_grad=True)
    y, l_aux, counts = moe(x)
    experts = moe.deepspeed_moe.experts.deepspeed_experts
    gates = F.softmax(F.linear(x.reshape(-1, d).float(), moe.deepspeed_moe.gate.wg.weight.float()), 1)
    top = gates.topk(2, dim=1)
    want = torch.zeros(15, d)
    for t in range(15):
        norm = top.values[t].sum()
        for e, g in zip(top.indices[t].tolist(), top.values[t].tolist()):
            want[t] += (g / norm) * experts[e](x.reshape(-1, d)[t])
    assert torch.allclose(y.reshape(-1, d), want, atol=1e-5), float((y.reshape(-1, d) - want).abs().max())
    assert int(counts.sum()) == 30 and float(l_aux) > 0
    (y.sum() + l_aux).backward()
    assert x.grad is not None and all(p.grad is not None for p in experts[0].parameters()) and \
        moe.deepspeed_moe.gate.wg.weight.grad is not None
    # residual MoE: a dense MLP next to the experts, mixed by a learned 2-way coefficient
    torch.manual_seed(1)
    res = MoE(hidden_size=d, expert=_Expert(d), num_experts=E, ep_size=1, k=1, capacity_factor=8.0, min_capacity=1,
              use_residual=True, use_rts=False)
    res.set_deepspeed_parallelism()
    x2 = torch.randn(2, 4, d)
    out, _, _ = res(x2)
    moe_part, _, _ = res.deepspeed_moe(x2) if not isinstance(res.deepspeed_moe(x2), torch.Tensor) else (res.deepspeed_moe(x2), 0, 0)
    coef = F.softmax(res.coefficient(x2), dim=-1)
    want2 = moe_part * coef[..., 0:1] + res.mlp(x2) * coef[..., 1:]
    assert torch.allclose(out, want2, atol=1e-5)


def test_moe_layer_capacity_path_matches_dense_evaluation():
    from conftest import run_dist
    run_dist(_moe_worker, 1)


def _moe_ep_worker(rank, world):
    """Expert parallelism of the capacity path: 4 experts over 2 ranks, each rank routes ITS tokens; the all-to-alls
    carry them to the owning rank and back.  Every rank's output equals the dense evaluation with all four experts."""
    from shuffle_exchange_b200.moe.layer import MoE
    d, E = 8, 4
    torch.manual_seed(0)
    all_experts = [_Expert(d) for _ in range(E)]                       # same on both ranks (same seed)
    gate_w = torch.randn(E, d) * 0.5
    moe = MoE(hidden_size=d, expert=_Expert(d), num_experts=E, ep_size=2, k=1, capacity_factor=16.0, min_capacity=1,
              use_rts=False)
    moe.set_deepspeed_parallelism()
    with torch.no_grad():
        moe.deepspeed_moe.gate.wg.weight.copy_(gate_w)
        for local, ex in enumerate(moe.deepspeed_moe.experts.deepspeed_experts):
            ex.load_state_dict(all_experts[rank * 2 + local].state_dict())   # rank r owns experts 2r, 2r+1
    g = torch.Generator().manual_seed(10 + rank)
    x = torch.randn(2, 6, d, generator=g, requires_grad=True)
    y, l_aux, counts = moe(x)
    flat = x.detach().reshape(-1, d)
    gates = F.softmax(F.linear(flat, gate_w), 1)
    want = torch.stack([gates[t].max() * all_experts[int(gates[t].argmax())](flat[t]) for t in range(flat.shape[0])])
    assert torch.allclose(y.reshape(-1, d), want, atol=1e-5), float((y.reshape(-1, d) - want).abs().max())
    y.sum().backward()
    assert x.grad is not None and x.grad.abs().sum() > 0
    # an expert's gradient contains the tokens the OTHER rank sent to it
    g0 = moe.deepspeed_moe.experts.deepspeed_experts[0].fc.weight.grad
    assert g0 is not None


def test_moe_capacity_path_with_expert_parallelism_two_ranks():
    from conftest import run_dist
    run_dist(_moe_ep_worker, 2)
