"""Batched gradient reductions used by ZeRO-3: coalesced reduce-scatter and the quantised hierarchical
all-to-all of ZeRO++ (qgZ), with and without LoCo error feedback.

Reference: ``deepspeed/runtime/comm/coalesced_collectives.py`` — ``reduce_scatter_coalesced`` (:159-218),
``all_to_all_quant_reduce`` (:30-79), ``all_to_all_loco_quant_reduce`` (:82-156).

Differences that matter on B200: the interleave is ONE strided copy per tensor into a ``[world, padded_total]``
staging buffer (the reference concatenates ``world x n_tensors`` slices), pre-division is folded into that copy, and
the quantise / dequant-reduce-requant / dequantise steps are the kernels of ``csrc/quantize.cu``.  When the ZeRO-3
optimizer runs with fused collectives this module is bypassed altogether (``runtime/zero/fused_collectives.py``:
reduce-scatter + Adam in one NVLink kernel); it remains the NCCL path and the multi-node path.
"""
from __future__ import annotations

import math
from typing import Dict, List

import torch
from torch import Tensor

from ... import comm as dist
from ...ops import quantizer as Q
from ...utils.logging import logger


@torch.no_grad()
def reduce_scatter_coalesced(tensors: List[Tensor], group=None) -> List[Tensor]:
    """Average every tensor over ``group`` and return, per tensor, this rank's ``ceil(numel/world)`` chunk (the
    last rank's chunk may be shorter)."""
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    chunks = [math.ceil(t.numel() / world) for t in tensors]
    if len(tensors) == 1 and tensors[0].numel() % world == 0:
        flat = tensors[0].view(-1)
        flat.div_(world)
    else:
        total = sum(chunks)
        staging = torch.empty(world, total, dtype=tensors[0].dtype, device=tensors[0].device)
        off = 0
        for t, c in zip(tensors, chunks):
            src = t.reshape(-1)
            full_rows = src.numel() // c
            dst = staging[:, off: off + c]
            if full_rows:
                torch.div(src[: full_rows * c].view(full_rows, c), world, out=dst[:full_rows])
            rem = src.numel() - full_rows * c
            if rem:
                torch.div(src[full_rows * c:], world, out=dst[full_rows, :rem])
                dst[full_rows, rem:].zero_()
            if full_rows + (1 if rem else 0) < world:
                dst[full_rows + (1 if rem else 0):].zero_()
            off += c
        flat = staging.view(-1)
    mine = flat.view(world, -1)[rank]
    dist.reduce_scatter_tensor(mine, flat, group=group)
    out, off = [], 0
    for t, c in zip(tensors, chunks):
        valid = max(0, min(c, t.numel() - rank * c))
        out.append(mine.narrow(0, off, valid))
        off += c
    return out


def _local_world(groups=None) -> int:
    """Ranks per node: what the group builder used (``utils.groups._get_local_all_to_all_group``), else the number of
    visible devices."""
    if groups is not None and "devices_per_node" in groups:
        return int(groups["devices_per_node"])
    return max(1, torch.cuda.device_count()) if torch.cuda.is_available() else 1


@torch.no_grad()
def all_to_all_quant_reduce(tensors: List[Tensor], groups: Dict[str, object]) -> List[Tensor]:
    """qgZ: int4 swizzle-quant -> intra-node all-to-all -> dequant-reduce-requant -> inter-node all-to-all ->
    dequant + mean over nodes.  ``groups`` holds ``local_<node>`` and ``global_<local rank>`` process groups
    (``utils.groups._get_local_all_to_all_group``).  1-D tensors and sizes not divisible by ``2 * world`` take the
    plain reduce-scatter."""
    local = _local_world(groups)
    world = dist.get_world_size()
    nodes = max(1, world // local)
    rank = dist.get_rank()
    node_idx, local_idx = rank // local, rank % local
    out: List[Tensor] = [None] * len(tensors)
    for i, t in enumerate(tensors):
        if t.dim() == 1 or t.numel() % (2 * world) != 0:
            if t.dim() != 1:
                logger.warning(f"qgZ falls back to reduce_scatter because tensor size = {t.numel()} is not divisible "
                               f"by (2 * global_world_size) = {2 * world}.")
            out[i] = reduce_scatter_coalesced([t])[0]
            continue
        intra_groups = max(t.shape[0], t.shape[1], world)
        while t.numel() % intra_groups or (t.numel() // intra_groups) % 8:
            intra_groups //= 2
        inter_groups = intra_groups // local
        q, s = Q.swizzle_quant(t, intra_groups, 4, True, 1, nodes, local)
        q_local, s_local = torch.empty_like(q), torch.empty_like(s)
        dist.all_to_all_single(q_local, q, group=groups[f"local_{node_idx}"])
        dist.all_to_all_single(s_local, s, group=groups[f"local_{node_idx}"])
        q_red, s_red = Q.quantized_reduction(q_local, s_local, intra_groups, inter_groups, 4, True, local)
        if nodes > 1:
            q_glob, s_glob = torch.empty_like(q_red), torch.empty_like(s_red)
            dist.all_to_all_single(q_glob, q_red, group=groups[f"global_{local_idx}"])
            dist.all_to_all_single(s_glob, s_red, group=groups[f"global_{local_idx}"])
        else:
            q_glob, s_glob = q_red, s_red
        full = Q.dequantize(q_glob, s_glob, s_glob.numel(), 4, True, t.dtype).view(-1)
        assert full.numel() % nodes == 0
        out[i] = (full.view(nodes, -1).float().sum(0) / (nodes * local)).to(t.dtype)
    return out


@torch.no_grad()
def all_to_all_loco_quant_reduce(params: List[Tensor], groups: Dict[str, object], loco_param: dict = None) -> List[Tensor]:
    """qgZ with LoCo error feedback (reference :82-156): the quantisation error of both hops is kept in per-parameter
    buffers (``p.intra_ef_buf`` / ``p.inter_ef_buf``), damped by ``err_beta`` and re-injected at the next step; the
    buffers are reset every ``reset_T`` steps."""
    loco_param = loco_param or {"err_beta": 0.8, "reset_T": 1024}
    beta, reset_T = loco_param["err_beta"], loco_param["reset_T"]
    local = _local_world(groups)
    world = dist.get_world_size()
    nodes = max(1, world // local)
    rank = dist.get_rank()
    node_idx, local_idx = rank // local, rank % local
    out: List[Tensor] = [None] * len(params)
    for i, p in enumerate(params):
        t = p.grad
        if t.dim() == 1 or t.numel() % (2 * world) != 0:
            out[i] = reduce_scatter_coalesced([t])[0]
            continue
        step = getattr(p, "_loco_step", 0)
        if not hasattr(p, "intra_ef_buf") or step % reset_T == 0:
            p.intra_ef_buf = torch.zeros_like(t, dtype=torch.float32)
            p.inter_ef_buf = None
        p._loco_step = step + 1
        intra_groups = max(t.shape[0], t.shape[1], world)
        while t.numel() % intra_groups or (t.numel() // intra_groups) % 8:
            intra_groups //= 2
        inter_groups = intra_groups // local
        comp = t.float() + p.intra_ef_buf
        q, s = Q.swizzle_quant(comp, intra_groups, 4, True, 1, nodes, local)
        # error of hop 1 in the ORIGINAL element order: undo the group swizzle on the dequantised tensor
        deq = Q.dequantize(q, s, intra_groups, 4, True, torch.float32)
        perm = Q.quantizer._swizzle_perm(intra_groups, nodes, local, t.device)
        p.intra_ef_buf = beta * (comp.view(intra_groups, -1) - deq[perm]).view_as(t)
        q_local, s_local = torch.empty_like(q), torch.empty_like(s)
        dist.all_to_all_single(q_local, q, group=groups[f"local_{node_idx}"])
        dist.all_to_all_single(s_local, s, group=groups[f"local_{node_idx}"])
        summed = Q.dequantize(q_local, s_local, intra_groups, 4, True, torch.float32).view(local, inter_groups, -1).sum(0)
        if p.inter_ef_buf is None:
            p.inter_ef_buf = torch.zeros_like(summed)
        summed = summed + p.inter_ef_buf
        q_red, s_red = Q.quantize(summed, inter_groups, 4, True)
        p.inter_ef_buf = beta * (summed - Q.dequantize(q_red, s_red, inter_groups, 4, True, torch.float32))
        if nodes > 1:
            q_glob, s_glob = torch.empty_like(q_red), torch.empty_like(s_red)
            dist.all_to_all_single(q_glob, q_red, group=groups[f"global_{local_idx}"])
            dist.all_to_all_single(s_glob, s_red, group=groups[f"global_{local_idx}"])
        else:
            q_glob, s_glob = q_red, s_red
        full = Q.dequantize(q_glob, s_glob, s_glob.numel(), 4, True, torch.float32).view(-1)
        out[i] = (full.view(nodes, -1).sum(0) / (nodes * local)).to(t.dtype)
    return out
