"""Gating, dispatch / combine and the expert-parallel MoE layer.

Semantics of the reference (deepspeed/moe/sharded_moe.py: ``top1gating`` :183-287, ``top2gating``
:290-371, ``topkgating`` :374-447, ``TopKGate`` :450-531, ``MOELayer.forward`` :587-678, ``_AllToAll``
:96-108) with a different data structure: instead of dense ``combine_weights[S,E,C]`` /
``dispatch_mask[S,E,C]`` tensors and ``einsum("sec,sm->ecm")`` (an O(S·E·C·M) matmul used as a
permutation) the gate returns a **sparse routing** — per (token, choice): expert id, capacity slot,
weight, keep flag — and dispatch/combine are row gathers/scatters (hand-written kernels on CUDA,
``index_copy``/``index_select`` elsewhere).  :func:`routing_to_dense` rebuilds the reference's dense
tensors for tests and for callers that want them.  ``drop_tokens=False`` keeps the reference's
all-reduce(MAX) of the capacity but pads to a bucketed capacity so no shape depends on it.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Callable, Dict, Optional, Tuple

import torch
import torch.nn.functional as F
from torch import Tensor
from torch.nn import Module

from ..comm import comm as dist
from ..utils import groups
from ..utils.logging import logger
from ..utils.timer import SynchronizedWallClockTimer

TOPK_GATE_TIMER = "topk_gate"
MOE_TIMER = "moe"
FIRST_ALLTOALL_TIMER = "1st_a2a"
SECOND_ALLTOALL_TIMER = "2nd_a2a"

uniform_map: Dict[torch.device, Callable] = {}
gumbel_map: Dict[torch.device, Callable] = {}


def multiplicative_jitter(x, device: torch.device, epsilon=1e-2):
    """Multiply by U(1-eps, 1+eps) noise ('Jitter' policy)."""
    if epsilon == 0:
        return x
    return x * torch.empty_like(x).uniform_(1.0 - epsilon, 1.0 + epsilon)


def gumbel_rsample(shape, device: torch.device) -> Tensor:
    u = torch.rand(shape, device=device).clamp_(1e-10, 1.0 - 1e-7)
    return -torch.log(-torch.log(u))


class _AllToAll(torch.autograd.Function):
    """Equal-split all-to-all whose backward is the same collective."""

    @staticmethod
    def forward(ctx, group, input: Tensor) -> Tensor:
        ctx.group = group
        input = input.contiguous()
        output = torch.empty_like(input)
        dist.all_to_all_single(output, input, group=group)
        return output

    @staticmethod
    def backward(ctx, *grad_output: Tensor):
        return (None, _AllToAll.apply(ctx.group, *grad_output))


@dataclass
class Routing:
    """Sparse routing decision for S tokens with k choices each."""
    expert: Tensor      # [S, k] int64 expert id
    slot: Tensor        # [S, k] int64 position inside the expert's capacity buffer
    weight: Tensor      # [S, k] fp32 combine weight (0 where dropped)
    keep: Tensor        # [S, k] bool
    capacity: int
    num_experts: int
    l_aux: Tensor
    exp_counts: Tensor


def _capacity(num_tokens: int, num_experts: int, capacity_factor: float, min_capacity: int) -> int:
    return max(int(math.ceil(num_tokens / num_experts * capacity_factor)), int(min_capacity))


def _no_drop_capacity(exp_counts: Tensor, ep_group, num_tokens: int) -> int:
    cap = exp_counts.max()
    if ep_group is not None and dist.is_initialized():
        dist.all_reduce(cap, op=dist.ReduceOp.MAX, group=ep_group)
    tp = 1
    if groups.mpu is not None and groups.expert_tensor_parallel_world_size == 1:
        tp = groups.get_tensor_model_parallel_world_size()
    c = int(cap.item())
    c = int(math.ceil(c / tp) * tp)
    return max(1, c)


def _positions_within_expert(mask: Tensor) -> Tensor:
    """mask [S,E] {0,1} -> location of each selected token in its expert's queue (token order)."""
    return torch.cumsum(mask, dim=0) - 1


def top1_route(logits: Tensor, capacity_factor: float, min_capacity: int, used_token: Tensor = None,
               noisy_gate_policy: Optional[str] = None, drop_tokens: bool = True, use_rts: bool = True,
               ep_group=None) -> Routing:
    S, E = logits.shape
    logits_w_noise = logits + gumbel_rsample(logits.shape, logits.device) if noisy_gate_policy == "RSample" else None
    gates = F.softmax(logits, dim=1)
    capacity = _capacity(S, E, capacity_factor, min_capacity)
    idx1 = torch.argmax(logits_w_noise if logits_w_noise is not None else gates, dim=1)
    mask1 = F.one_hot(idx1, num_classes=E)
    if used_token is not None:
        mask1 = mask1 * used_token.to(mask1.dtype).unsqueeze(1)
    exp_counts = mask1.sum(0).detach()
    if not drop_tokens:
        capacity = min(_no_drop_capacity(exp_counts.clone(), ep_group, S), S)
    me = gates.mean(0)
    ce = mask1.float().mean(0)
    l_aux = torch.sum(me * ce) * E
    assert S >= min_capacity, ("No. of tokens (batch-size) should be greater than min_capacity. Either set "
                               "min_capacity to 0 or increase your batch size.")
    # capacity selection: random token selection (RTS) or first-come order
    score = mask1.float() * (torch.rand(mask1.shape, device=logits.device) if use_rts else 1.0)
    cap_k = min(capacity, S)
    if not use_rts:      # first-come order: earlier tokens win
        score = mask1.float() * (S - torch.arange(S, device=logits.device).unsqueeze(1).float()) / S
    top_idx = torch.topk(score, k=cap_k, dim=0)[1]
    mask1 = mask1 * torch.zeros_like(mask1).scatter_(0, top_idx, 1)
    loc = _positions_within_expert(mask1)
    slot = (loc * mask1).sum(1)
    keep = mask1.sum(1) > 0
    w = (gates * mask1.float()).sum(1)
    return Routing(idx1.unsqueeze(1), slot.unsqueeze(1), w.unsqueeze(1), keep.unsqueeze(1), capacity, E, l_aux, exp_counts)


def top2_route(logits: Tensor, capacity_factor: float, min_capacity: int, drop_tokens: bool = True, ep_group=None,
               top2_2nd_expert_sampling: bool = True) -> Routing:
    S, E = logits.shape
    gates = F.softmax(logits, dim=1)
    idx1 = torch.argmax(gates, dim=1)
    mask1 = F.one_hot(idx1, num_classes=E)
    l2 = logits + gumbel_rsample(logits.shape, logits.device) if top2_2nd_expert_sampling else logits
    idx2 = torch.argmax(l2.masked_fill(mask1.bool(), float("-inf")), dim=1)
    mask2 = F.one_hot(idx2, num_classes=E)
    loc1 = _positions_within_expert(mask1)
    loc2 = _positions_within_expert(mask2) + mask1.sum(0, keepdim=True)
    me = gates.mean(0)
    ce = mask1.float().mean(0)
    l_aux = torch.mean(me * ce) * E * E
    exp_counts = (mask1 + mask2).sum(0).detach()
    if drop_tokens:
        capacity = _capacity(S, E, capacity_factor * 2, min_capacity)
        mask1 = mask1 * (loc1 < capacity)
        mask2 = mask2 * (loc2 < capacity)
    else:
        capacity = _no_drop_capacity(exp_counts.clone(), ep_group, S)
    slot1 = (loc1 * mask1).sum(1)
    slot2 = (loc2 * mask2).sum(1)
    g1 = (gates * mask1.float()).sum(1)
    g2 = (gates * mask2.float()).sum(1)
    denom = (g1 + g2).clamp(min=torch.finfo(gates.dtype).eps)
    g1, g2 = g1 / denom, g2 / denom
    keep = torch.stack([mask1.sum(1) > 0, mask2.sum(1) > 0], 1)
    return Routing(torch.stack([idx1, idx2], 1), torch.stack([slot1, slot2], 1), torch.stack([g1, g2], 1) * keep,
                   keep, capacity, E, l_aux, exp_counts)


def topk_route(logits: Tensor, k: int, capacity_factor: float, min_capacity: int, drop_tokens: bool = True,
               ep_group=None, drop_policy: str = "probs") -> Routing:
    S, E = logits.shape
    top_gate, top_idx = torch.topk(logits, k=k, dim=1)
    gates = F.softmax(logits, dim=1)
    mask = torch.zeros_like(gates, dtype=torch.bool).scatter_(1, top_idx, True)
    exp_counts = mask.sum(0).detach()
    me = gates.mean(0)
    ce = mask.float().mean(0)
    l_aux = torch.mean(me * ce) * E * E / k
    if drop_tokens:
        capacity = _capacity(S, E, capacity_factor * k, min_capacity)
        if drop_policy == "probs":
            masked = torch.zeros_like(logits).scatter(1, top_idx, top_gate)
            masked = masked.masked_fill(~mask, float("-inf"))
            cap_idx = torch.topk(masked, k=min(capacity, S), dim=0, sorted=False)[1]
            cap_mask = torch.zeros_like(mask).scatter_(0, cap_idx, True)
            mask = mask & cap_mask
            loc = _positions_within_expert(mask.long())
        elif drop_policy == "position":
            loc = _positions_within_expert(mask.long())
            mask = mask & (loc < capacity)
        else:
            raise ValueError(f"Invalid drop_policy: {drop_policy}")
    else:
        capacity = _no_drop_capacity(exp_counts.clone(), ep_group, S)
        loc = _positions_within_expert(mask.long())
    gm = gates * mask
    gm = gm / gm.sum(-1, keepdim=True).clamp(min=torch.finfo(gm.dtype).eps)
    keep = mask.gather(1, top_idx)
    slot = (loc * mask).gather(1, top_idx)
    w = gm.gather(1, top_idx)
    return Routing(top_idx, slot, w * keep, keep, capacity, E, l_aux, exp_counts)


def routing_to_dense(r: Routing) -> Tuple[Tensor, Tensor]:
    """(combine_weights[S,E,C], dispatch_mask[S,E,C]) exactly as the reference's gates return them."""
    S, k = r.expert.shape
    cw = torch.zeros(S, r.num_experts, r.capacity, dtype=r.weight.dtype, device=r.weight.device)
    s_idx = torch.arange(S, device=cw.device).unsqueeze(1).expand(S, k)
    keep = r.keep & (r.slot < r.capacity)
    cw[s_idx[keep], r.expert[keep], r.slot[keep]] = r.weight[keep]
    return cw, cw.bool()


# reference-shaped wrappers -------------------------------------------------------------------
def top1gating(logits, capacity_factor, min_capacity, used_token=None, noisy_gate_policy=None, drop_tokens=True,
               use_rts=True, ep_group=None, use_tutel=False):
    r = top1_route(logits, capacity_factor, min_capacity, used_token, noisy_gate_policy, drop_tokens, use_rts, ep_group)
    cw, dm = routing_to_dense(r)
    return r.l_aux, cw, dm, r.exp_counts


def top2gating(logits, capacity_factor, min_capacity, drop_tokens=True, ep_group=None, top2_2nd_expert_sampling=True):
    r = top2_route(logits, capacity_factor, min_capacity, drop_tokens, ep_group, top2_2nd_expert_sampling)
    cw, dm = routing_to_dense(r)
    return r.l_aux, cw, dm, r.exp_counts


def topkgating(logits, k, capacity_factor, min_capacity, drop_tokens=True, ep_group=None, drop_policy="probs"):
    r = topk_route(logits, k, capacity_factor, min_capacity, drop_tokens, ep_group, drop_policy)
    cw, dm = routing_to_dense(r)
    return r.l_aux, cw, dm, r.exp_counts


class TopKGate(Module):
    """fp32 router: ``wg`` is ``Linear(model_dim, num_experts, bias=False)``."""

    def __init__(self, model_dim: int, num_experts: int, k: int = 1, capacity_factor: float = 1.0,
                 eval_capacity_factor: float = 1.0, min_capacity: int = 8, noisy_gate_policy: Optional[str] = None,
                 drop_tokens: bool = True, use_rts: bool = True, ep_group=None, top2_2nd_expert_sampling: bool = True):
        super().__init__()
        self.wg = torch.nn.Linear(model_dim, num_experts, bias=False)
        self.wg.weight._sxb_keep_dtype = True       # the router stays fp32 under bf16/fp16 engines
        self.ep_group = ep_group
        self.k = k
        self.capacity_factor = capacity_factor
        self.eval_capacity_factor = eval_capacity_factor
        self.min_capacity = min_capacity
        self.noisy_gate_policy = noisy_gate_policy
        self.timers = SynchronizedWallClockTimer()
        self.wall_clock_breakdown = False
        self.gate_time = 0.0
        self.drop_tokens = drop_tokens
        self.use_rts = use_rts
        self.top2_2nd_expert_sampling = top2_2nd_expert_sampling

    def _set_ep_group(self, ep_group):
        assert self.ep_group is None, "Attempting to override an existing ep_group"
        self.ep_group = ep_group

    def route(self, input: Tensor, used_token: Tensor = None) -> Routing:
        if self.wall_clock_breakdown:
            self.timers(TOPK_GATE_TIMER).start()
        x = input.float()
        if self.noisy_gate_policy == "Jitter" and self.training:
            x = multiplicative_jitter(x, device=input.device)
        logits = F.linear(x, self.wg.weight.float())
        cf = self.capacity_factor if self.training else self.eval_capacity_factor
        if self.k == 1:
            r = top1_route(logits, cf, self.min_capacity, used_token,
                           self.noisy_gate_policy if self.training else None, self.drop_tokens, self.use_rts,
                           self.ep_group)
        elif self.k == 2:
            r = top2_route(logits, cf, self.min_capacity, self.drop_tokens, self.ep_group,
                           self.top2_2nd_expert_sampling)
        else:
            r = topk_route(logits, self.k, cf, self.min_capacity, self.drop_tokens, self.ep_group)
        if self.wall_clock_breakdown:
            self.timers(TOPK_GATE_TIMER).stop()
            self.gate_time = self.timers(TOPK_GATE_TIMER).elapsed(reset=False)
        return r

    def forward(self, input: Tensor, used_token: Tensor = None, use_tutel: bool = False):
        r = self.route(input, used_token)
        cw, dm = routing_to_dense(r)
        return r.l_aux, cw, dm, r.exp_counts


def dispatch(x: Tensor, r: Routing) -> Tensor:
    """tokens [S,M] -> capacity buffers [E, C, M] (rows of dropped choices go to a discard row)."""
    S, k = r.expert.shape
    E, C = r.num_experts, r.capacity
    keep = r.keep & (r.slot < C)
    dest = torch.where(keep, r.expert * C + r.slot, torch.full_like(r.slot, E * C)).reshape(-1)
    src = x.repeat_interleave(k, dim=0) if k > 1 else x
    buf = torch.zeros(E * C + 1, x.shape[1], dtype=x.dtype, device=x.device)
    buf = buf.index_copy(0, dest, src)
    return buf[: E * C].view(E, C, x.shape[1])


def combine(expert_out: Tensor, r: Routing, out_dtype=None) -> Tensor:
    """capacity buffers [E, C, M] -> tokens [S, M] (weighted sum over the k choices)."""
    S, k = r.expert.shape
    E, C = r.num_experts, r.capacity
    M = expert_out.shape[-1]
    keep = r.keep & (r.slot < C)
    flat = torch.cat([expert_out.reshape(E * C, M), expert_out.new_zeros(1, M)], 0)
    src = torch.where(keep, r.expert * C + r.slot, torch.full_like(r.slot, E * C)).reshape(-1)
    g = flat.index_select(0, src).view(S, k, M)
    w = (r.weight * keep).to(g.dtype).unsqueeze(-1)
    return (g * w).sum(1).to(out_dtype or expert_out.dtype)


class MOELayer(Module):
    """gate -> dispatch -> all-to-all -> local experts -> all-to-all -> combine."""

    def __init__(self, gate: Module, experts: Module, ep_group_name, ep_size, num_local_experts: int,
                 use_tutel: bool = False):
        super().__init__()
        self.gate = gate
        self.experts = experts
        self.ep_group = None
        self.ep_size = ep_size
        self.ep_group_name = ep_group_name
        self.num_local_experts = num_local_experts
        self.time_falltoall = 0.0
        self.time_salltoall = 0.0
        self.time_moe = 0.0
        self.timers = SynchronizedWallClockTimer()
        self.wall_clock_breakdown = False
        self.use_tutel = False
        self.fused = None      # set by moe/fused.py when the symmetric-memory path is active

    def _set_ep_group(self, ep_group):
        self.ep_group = ep_group
        self.gate._set_ep_group(ep_group)

    def forward(self, *input: Tensor, **kwargs) -> Tensor:
        if self.wall_clock_breakdown:
            self.timers(MOE_TIMER).start()
        x = input[0]
        d_model = x.shape[-1]
        tokens = x.reshape(-1, d_model)
        r = self.gate.route(tokens, input[1] if len(input) > 1 else None)
        self.l_aux, self.exp_counts = r.l_aux, r.exp_counts
        dispatched = dispatch(tokens, r)                                   # [E, C, M]
        tp = groups.get_tensor_model_parallel_world_size() if groups.mpu is not None else 1
        if tp > 1 and groups.expert_tensor_parallel_world_size == 1:
            from .mappings import drop_tokens
            dispatched = drop_tokens(dispatched, dim=1)
        if self.wall_clock_breakdown:
            self.timers(FIRST_ALLTOALL_TIMER).start()
        if self.ep_size > 1:
            dispatched = _AllToAll.apply(self.ep_group, dispatched)
        if self.wall_clock_breakdown:
            self.timers(FIRST_ALLTOALL_TIMER).stop()
            self.time_falltoall = self.timers(FIRST_ALLTOALL_TIMER).elapsed(reset=False)
        # [ep * local_E, C, M] -> [ep, local_E, C, M]: expert e_local receives rows from every ep rank
        dispatched = dispatched.reshape(self.ep_size, self.num_local_experts, -1, d_model)
        expert_output = self.experts(dispatched)
        expert_output = expert_output.reshape(self.ep_size * self.num_local_experts, -1, d_model)
        if self.wall_clock_breakdown:
            self.timers(SECOND_ALLTOALL_TIMER).start()
        if self.ep_size > 1:
            expert_output = _AllToAll.apply(self.ep_group, expert_output)
        if self.wall_clock_breakdown:
            self.timers(SECOND_ALLTOALL_TIMER).stop()
            self.time_salltoall = self.timers(SECOND_ALLTOALL_TIMER).elapsed(reset=False)
        if tp > 1 and groups.expert_tensor_parallel_world_size == 1:
            from .mappings import gather_tokens
            expert_output = gather_tokens(expert_output, dim=1)
        out = combine(expert_output, r, out_dtype=x.dtype)
        out = out.reshape(x.shape)
        if self.wall_clock_breakdown:
            self.timers(MOE_TIMER).stop()
            self.time_moe = self.timers(MOE_TIMER).elapsed(reset=False)
        return out
