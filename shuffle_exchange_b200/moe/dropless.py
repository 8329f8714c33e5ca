"""Dropless (capacity-free) expert-parallel MoE block — the Mixtral-style layer.

The reference's ``MOELayer`` (``moe/sharded_moe.py:587-678``) pads every expert to a fixed capacity, dispatches with a
dense ``[S, E, C]`` one-hot einsum and moves ``[E, C, M]`` through ``all_to_all_single``.  This block keeps the same
place in the model and the same return convention ``(output, l_aux, exp_counts)`` but:

* routing is ``sxb_top_k_gating`` (softmax over all experts, top-k, renormalised) and dispatch/combine are the row
  movers ``sxb_moe_scatter`` / ``sxb_moe_gather`` — O(T·k·H) bytes instead of O(S·E·C·M) FLOPs;
* no token is dropped and nothing is padded: ranks exchange the per-expert row counts first, then exactly the
  routed rows (``all_to_all_single`` with split sizes, or the symmetric-memory row kernel ``sxb_symm_a2a_rows``);
* each local expert runs its SwiGLU as two tcgen05 GEMMs over its contiguous row range.

It subclasses :class:`MoE` so that engine discovery, expert process groups and ZeRO-1/2 expert parameter groups
(``param.allreduce = False``, ``param.group_name``) work unchanged.
"""
from __future__ import annotations

from typing import List, Optional

import torch
import torch.nn as nn

from .. import comm as dist
from ..ops import functional as F
from ..ops import moe_ops
from ..utils import groups
from ..utils.logging import log_dist
from .layer import MoE


class _AllToAllV(torch.autograd.Function):
    """``all_to_all_single`` of rows with per-rank split sizes; backward runs the transposed exchange."""

    @staticmethod
    def forward(ctx, group, x, send_splits: List[int], recv_splits: List[int]):
        ctx.group, ctx.send_splits, ctx.recv_splits = group, send_splits, recv_splits
        out = x.new_empty((sum(recv_splits),) + tuple(x.shape[1:]))
        dist.all_to_all_single(out, x.contiguous(), output_split_sizes=recv_splits, input_split_sizes=send_splits,
                               group=group)
        return out

    @staticmethod
    def backward(ctx, g):
        out = g.new_empty((sum(ctx.send_splits),) + tuple(g.shape[1:]))
        dist.all_to_all_single(out, g.contiguous(), output_split_sizes=ctx.send_splits,
                               input_split_sizes=ctx.recv_splits, group=ctx.group)
        return None, out, None, None


class _RouterLogits(torch.autograd.Function):
    """fp32 router logits from bf16 activations / gate weights: tcgen05 GEMM with an fp32 epilogue, fp32 gradients in."""

    @staticmethod
    def forward(ctx, x, wg):
        ctx.save_for_backward(x, wg)
        return F.gemm(x, wg, F.GEMM_NT, out_dtype=torch.float32)

    @staticmethod
    def backward(ctx, dl):
        x, wg = ctx.saved_tensors
        E = wg.shape[0]
        pad = (-E) % 8                       # the GEMM wants 16-byte rows: pad the expert dimension with zeros
        dlb = dl.to(torch.bfloat16)
        wgp = wg
        if pad:
            dlb = torch.nn.functional.pad(dlb, (0, pad))
            wgp = torch.nn.functional.pad(wg, (0, 0, 0, pad))
        dx = F.gemm(dlb, wgp, F.GEMM_NN) if ctx.needs_input_grad[0] else None
        dw = F.gemm(dlb, x, F.GEMM_TN, out_dtype=wg.dtype)[:E] if ctx.needs_input_grad[1] else None
        return dx, dw


class SwiGLUExpert(nn.Module):
    def __init__(self, hidden: int, intermediate: int, device=None, dtype=None):
        super().__init__()
        self.w_gate_up = nn.Parameter(torch.empty(2 * intermediate, hidden, device=device, dtype=dtype))
        self.w_down = nn.Parameter(torch.empty(hidden, intermediate, device=device, dtype=dtype))

    def reset_parameters(self, std: float):
        for w in (self.w_gate_up, self.w_down):
            if w.device.type != "meta":
                nn.init.normal_(w, 0.0, std)

    def forward(self, x):
        return F.linear(F.swiglu(F.linear(x, self.w_gate_up)), self.w_down)


class DroplessMoE(MoE):
    def __init__(self, hidden_size: int, intermediate_size: int, num_experts: int, ep_size: int = 1, k: int = 2,
                 aux_loss_coef: float = 0.0, device=None, dtype=None, fused_exchange: Optional[bool] = None,
                 fp8_experts: bool = False, ep_capacity_factor: float = 2.0):
        nn.Module.__init__(self)
        assert num_experts % ep_size == 0, f"Number of experts ({num_experts}) should be divisible by expert parallel size ({ep_size})"
        self.hidden_size, self.num_experts, self.ep_size, self.k = hidden_size, num_experts, ep_size, k
        self.num_local_experts = num_experts // ep_size
        self.expert_group_name = f"ep_size_{ep_size}"
        self.use_residual = False
        self.enable_expert_tensor_parallelism = False
        self.aux_loss_coef = aux_loss_coef
        # fused_exchange: None = use the symmetric-memory dispatch/combine + grouped GEMM whenever it is available
        # (CUDA bf16; NCCL group with symmetric memory, or a single rank); False = NCCL all_to_all_single path
        self.fused_exchange = fused_exchange
        self.fp8_experts = fp8_experts
        self.ep_capacity_factor = ep_capacity_factor
        self._fused_ep = None
        self.wg = nn.Parameter(torch.empty(num_experts, hidden_size, device=device, dtype=dtype))
        self.experts = nn.ModuleList([SwiGLUExpert(hidden_size, intermediate_size, device, dtype)
                                      for _ in range(self.num_local_experts)])
        for p in self.experts.parameters():
            p.allreduce = False
            p.group_name = self.expert_group_name
        self.ep_group = None
        self.l_aux = torch.zeros((), device=device if device is not None and str(device) != "meta" else None)
        self.exp_counts = None
        log_dist(f"Creating dropless MoE layer with num_experts: {num_experts} | num_local_experts: "
                 f"{self.num_local_experts} | expert_parallel_size: {ep_size} | top-{k}", [0])

    def reset_parameters(self, std: float):
        if self.wg.device.type != "meta":
            nn.init.normal_(self.wg, 0.0, std)
        for e in self.experts:
            e.reset_parameters(std)

    def _create_process_groups(self, use_data_before_expert_parallel_: bool = False) -> None:
        if not dist.is_initialized():
            return
        if self.expert_group_name not in groups._get_expert_parallel_group_dict():
            groups._create_expert_and_data_parallel(self.ep_size, use_data_before_expert_parallel_)
        self.ep_group = groups._get_expert_parallel_group(self.expert_group_name)

    # --------------------------------------------------------------------------------------------
    def _run_experts(self, rows: torch.Tensor, ends: List[int]) -> torch.Tensor:
        out = torch.empty_like(rows)
        pieces, lo = [], 0
        for e, hi in enumerate(ends):
            if hi > lo:
                pieces.append(self.experts[e](rows[lo:hi]))
            lo = hi
        if len(pieces) == 1:
            return pieces[0]                   # one local expert (EP = number of experts): no concatenation copy
        return torch.cat(pieces) if pieces else out

    # ------------------------------------------------------------------------------ fused path
    def _fused_available(self, x) -> bool:
        if self.fused_exchange is False or not x.is_cuda or x.dtype != torch.bfloat16:
            return False
        import os
        if os.environ.get("SXB_MOE_FUSED", "1") == "0":
            return False
        ep = self.ep_size if (self.ep_group is not None and self.ep_size > 1) else 1
        if ep == 1:
            return True
        from ..comm import symm
        return symm.available(self.ep_group)

    def _stacked(self, name: str):
        """[L, N, K] weights of the local experts; for one local expert the parameter itself (main-grad protocol
        stays attached), otherwise a stacked copy that autograd un-stacks."""
        ws = [getattr(e, name) for e in self.experts]
        return ws[0] if len(ws) == 1 else torch.stack(ws)

    def _forward_fused(self, x, scores, assign, offsets, counts):
        """K4/K5: device-side count exchange, fused permute+dispatch, grouped expert GEMMs (optionally fp8 forward),
        fused combine+un-permute — no host synchronisation anywhere (moe/fused_ep.py)."""
        from . import fused_ep
        from ..ops.gemm_ext import grouped_linear
        ep = self.ep_size if (self.ep_group is not None and self.ep_size > 1) else 1
        fe = fused_ep.FusedEP.get(self.ep_group if ep > 1 else None, self.hidden_size, self.num_experts, x.shape[0],
                                  self.k, self.ep_capacity_factor)
        self._fused_ep = fe
        plan = fused_ep.make_plan(fe, counts, assign, offsets)
        rows = fused_ep.dispatch(x, fe, plan)
        h = grouped_linear(rows, self._stacked("w_gate_up"), plan.group_offsets, self.fp8_experts)
        a = F.swiglu(h)
        c = grouped_linear(a, self._stacked("w_down"), plan.group_offsets, self.fp8_experts, out=fe.ret_rows)
        return fused_ep.combine(c, scores.to(torch.float32), fe, plan)

    def check_overflow(self):
        if self._fused_ep is not None:
            self._fused_ep.check_overflow()

    def forward(self, hidden_states: torch.Tensor, used_token: Optional[torch.Tensor] = None):
        shape = hidden_states.shape
        x = hidden_states.reshape(-1, shape[-1])
        T = x.shape[0]
        # router: [T, H] x [E, H]^T; fp32 logits out of the tcgen05 GEMM on CUDA (the reference gates in fp32,
        # moe/sharded_moe.py:500-531)
        if x.is_cuda and x.dtype == torch.bfloat16 and self.wg.dtype == torch.bfloat16 and shape[-1] % 8 == 0:
            logits = _RouterLogits.apply(x.contiguous(), self.wg)
        else:
            logits = torch.nn.functional.linear(x, self.wg)
        scores, assign, offsets, counts = moe_ops.top_k_gating(logits.detach(), self.k, renormalize=True)
        # the scores must carry gradient to the router: recompute them differentiably from the logits
        probs = torch.softmax(logits.float(), dim=-1)
        picked = probs.gather(1, assign.long())
        scores = picked / picked.sum(dim=-1, keepdim=True)
        self.exp_counts = counts
        if self.aux_loss_coef > 0.0 or self.training:
            frac = counts.float() / max(1, T * self.k)
            self.l_aux = self.num_experts * torch.sum(frac * probs.mean(dim=0))
        if self._fused_available(x):
            y = self._forward_fused(x, scores, assign, offsets, counts)
            return y.reshape(shape), self.l_aux, self.exp_counts
        moe_in, cumsum, slots = moe_ops.moe_scatter(x, assign, offsets, counts)
        ep = self.ep_size if (self.ep_group is not None and self.ep_size > 1) else 1
        if ep == 1:
            out_rows = self._run_experts(moe_in, cumsum.tolist())
        else:
            L = self.num_local_experts
            send_counts = counts.view(ep, L)
            recv_counts = torch.empty_like(send_counts)
            dist.all_to_all_single(recv_counts, send_counts.contiguous(), group=self.ep_group)
            both = torch.stack([send_counts, recv_counts]).cpu()          # ONE host sync per layer
            send_splits, recv_splits = both[0].sum(1).tolist(), both[1].sum(1).tolist()
            recv = _AllToAllV.apply(self.ep_group, moe_in, send_splits, recv_splits)
            rc = both[1]                                                   # [src, local expert]
            if L == 1:
                out_local = self._run_experts(recv, [recv.shape[0]])
            else:
                # received order is [src][expert]; experts want [expert][src]
                starts = torch.cumsum(rc.reshape(-1), 0) - rc.reshape(-1)
                idx = torch.cat([torch.arange(int(starts[s * L + e]), int(starts[s * L + e] + rc[s, e]))
                                 for e in range(L) for s in range(ep)]).to(recv.device)
                grouped = recv.index_select(0, idx)
                ends = torch.cumsum(rc.sum(0), 0).tolist()
                out_grouped = self._run_experts(grouped, ends)
                out_local = torch.empty_like(out_grouped).index_copy(0, idx, out_grouped)
            out_rows = _AllToAllV.apply(self.ep_group, out_local, recv_splits, send_splits)
        y = moe_ops.moe_gather(out_rows, scores.to(torch.float32), slots)
        return y.reshape(shape), self.l_aux, self.exp_counts
