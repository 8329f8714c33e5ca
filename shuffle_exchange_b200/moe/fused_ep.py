"""Expert-parallel token exchange over symmetric memory (K4 / K5): fused permute + dispatch, fused combine +
weighted un-permute, device-side count exchange — no NCCL, no host synchronisation.

Data flow of one MoE layer on rank ``r`` (``T`` local tokens, top-``k``, ``E = ep * L`` experts)::

    gating kernel                         -> assign[T,k], offsets[T,k] (arrival order inside the expert), counts[E]
    plan     (1 CTA, csrc/moe_ep.cu)      -> counts of every rank exchanged by peer stores; send_base[E] and the
                                             256-aligned group offsets of my local experts, all on the device
    push                                  -> x[t] stored straight into the owning rank's receive buffer at its final,
                                             expert-contiguous row (peer stores over NVLink)
    grouped GEMMs (ops/gemm_ext.py)       -> gate|up, SwiGLU, down over the local experts' rows; the last GEMM
                                             writes into the symmetric return buffer
    pull                                  -> y[t] = sum_j score[t,j] * (row in the owner's return buffer) (peer loads)

Backward is the same two movers with the roles swapped: the combine's backward PUSHES ``score * dy`` rows (and forms
``d score = <dy, expert output>`` from the rows it kept in forward), the dispatch's backward PULLS the rows of
``d(expert input)`` and sums them per token.

The reference's equivalent is the dense ``[S, E, C]`` one-hot einsum + ``all_to_all_single`` (+ a host sync when
tokens may not be dropped): ``moe/sharded_moe.py:587-678, 214-225``.

Capacity: receive / return buffers are allocated once for ``capacity_factor * T * k`` rows (plus alignment padding).
Routing that exceeds it on some rank is reported by the plan kernel and raised at the next :meth:`check_overflow`
(once per step, one 4-byte read) — tokens are never dropped silently.
"""
from __future__ import annotations

import ctypes
from typing import Dict, Optional

import torch

from ..comm import symm
from ..native import lib as _n
from ..ops.gemm_ext import GROUP_ALIGN_M

_n.register({
    "sxb_moe_ep_plan": "ppii" "pp" "iii" "iii" "p",
    "sxb_moe_ep_push": "ppii" "pppp" "ppp" "iiiii" "i" "i" "pi" "p",
    "sxb_moe_ep_pull": "ppii" "ppp" "ppp" "iiiii" "i" "i" "pi" "p",
    "sxb_moe_ep_hang_info": "p",
})

CH_PLAN, CH_LANDED, CH_READY, CH_PULLED = 8, 9, 10, 11


class EPPlan:
    """Device-side routing tables of one layer call (kept for its backward)."""

    def __init__(self, tables, assign, offsets, T, K, group_offsets):
        self.tables, self.assign, self.offsets, self.T, self.K = tables, assign, offsets, T, K
        self.group_offsets = group_offsets          # int32 [L+1] view into ``tables``


class _LocalBuffer:
    """Stand-in for :class:`symm.SymmBuffer` on a single rank: the same pointer-table interface over plain memory, so
    the single-GPU layer runs the very same kernels (world = 1: every "peer" access is local)."""

    def __init__(self, numel, dtype, device):
        self.tensor = torch.zeros(numel, dtype=dtype, device=device)
        self.pad = torch.zeros(32 * 16, dtype=torch.int32, device=device)
        self._bufs = (ctypes.c_void_p * 1)(self.tensor.data_ptr())
        self._sigs = (ctypes.c_void_p * 1)(self.pad.data_ptr())

    @property
    def bufs(self):
        return ctypes.cast(self._bufs, ctypes.c_void_p)

    @property
    def sigs(self):
        return ctypes.cast(self._sigs, ctypes.c_void_p)


class _LocalGroup:
    rank, world = 0, 1

    def __init__(self, device):
        self.counter = torch.zeros(16, dtype=torch.int32, device=device)

    def _ctr(self, i):
        return self.counter.data_ptr() + 4 * i


class FusedEP:
    """Symmetric buffers + epochs of one expert-parallel group; shared by every MoE layer of the model (layers run
    one after the other, and every use of a buffer is fenced by the plan handshake)."""

    _instances: Dict[tuple, "FusedEP"] = {}

    def __init__(self, group, hidden: int, num_experts: int, max_tokens: int, top_k: int,
                 capacity_factor: float = 1.5, dtype=torch.bfloat16):
        self.group = group
        local = group is None
        dev = torch.device("cuda", torch.cuda.current_device())
        self.sg = _LocalGroup(dev) if local else symm.SymmGroup.get(group)
        self.rank, self.world = self.sg.rank, self.sg.world
        self.E, self.L = num_experts, num_experts // self.world
        self.capacity_factor = 1.0 if group is None else capacity_factor
        self.H, self.K = hidden, top_k
        rows = int(max_tokens * top_k * (1.0 if local else capacity_factor)) + self.L * GROUP_ALIGN_M
        self.cap_rows = (rows + GROUP_ALIGN_M - 1) // GROUP_ALIGN_M * GROUP_ALIGN_M
        mk = (lambda n, dt: _LocalBuffer(n, dt, dev)) if local else (lambda n, dt: symm.SymmBuffer(n, dt, group))
        self.recv = mk(self.cap_rows * hidden, dtype)
        self.ret = mk(self.cap_rows * hidden, dtype)
        self.counts = mk(self.world * num_experts, torch.int32)
        self.recv_rows = self.recv.tensor.view(self.cap_rows, hidden)
        self.ret_rows = self.ret.tensor.view(self.cap_rows, hidden)
        self.epoch = 0
        self.pull_epoch = 0
        self._overflow_sources = []

    @classmethod
    def get(cls, group, hidden, num_experts, max_tokens, top_k, capacity_factor=1.5) -> "FusedEP":
        key = (id(group) if group is not None else 0, hidden, num_experts, top_k)
        inst = cls._instances.get(key)
        world = 1 if group is None else torch.distributed.get_world_size(group)
        cf = 1.0 if group is None else capacity_factor
        # the same row count the constructor reserves: routed rows plus one alignment pad per local expert
        need = int(max_tokens * top_k * cf) + (num_experts // world) * GROUP_ALIGN_M
        if inst is None or inst.cap_rows < need:
            inst = cls._instances[key] = FusedEP(group, hidden, num_experts, max_tokens, top_k, capacity_factor)
        return inst

    # ---------------------------------------------------------------------------------- steps
    def plan(self, counts: Optional[torch.Tensor], tables: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Handshake (+ count exchange when ``counts`` is given).  Returns the tables tensor."""
        self.epoch += 1
        exchange = counts is not None
        if tables is None:
            tables = torch.zeros(self.E + 2 * self.L + 2, dtype=torch.int32, device=self.recv.tensor.device)
        # signals live on the return buffer's pad
        _n.call("sxb_moe_ep_plan", self.counts.bufs, self.ret.sigs, self.rank, self.world, _n.ptr(counts),
                tables.data_ptr(), self.E, self.L, self.cap_rows, self.epoch, self.pull_epoch, int(exchange),
                _n.stream_ptr())
        if exchange:
            self._overflow_sources.append(tables)
        return tables

    def push(self, x, plan: EPPlan, weights=None, pulled=None, d_scores=None):
        """Rows of ``x`` -> peers' receive buffers (after :meth:`plan` of this epoch); then wait until every peer's rows
        for ME have landed."""
        _n.call("sxb_moe_ep_push", self.recv.bufs, self.ret.sigs, self.rank, self.world, x.data_ptr(),
                _n.ptr(weights), _n.ptr(pulled), _n.ptr(d_scores), plan.assign.data_ptr(), plan.offsets.data_ptr(),
                plan.tables.data_ptr(), plan.T, plan.K, self.H, self.E, self.L, self.cap_rows, self.epoch,
                self.sg._ctr(3), 0, _n.stream_ptr())
        _n.call("sxb_symm_wait", self.ret.bufs, self.ret.sigs, self.rank, self.world, CH_LANDED, self.epoch,
                _n.stream_ptr())

    def pull(self, out, plan: EPPlan, weights=None, keep=None):
        """Announce that my return buffer holds this epoch's rows, wait for every peer's, then gather-reduce."""
        _n.call("sxb_symm_signal", self.ret.bufs, self.ret.sigs, self.rank, self.world, CH_READY, self.epoch,
                _n.stream_ptr())
        _n.call("sxb_symm_wait", self.ret.bufs, self.ret.sigs, self.rank, self.world, CH_READY, self.epoch,
                _n.stream_ptr())
        _n.call("sxb_moe_ep_pull", self.ret.bufs, self.ret.sigs, self.rank, self.world, out.data_ptr(),
                _n.ptr(weights), _n.ptr(keep), plan.assign.data_ptr(), plan.offsets.data_ptr(),
                plan.tables.data_ptr(), plan.T, plan.K, self.H, self.E, self.L, self.cap_rows, self.epoch,
                self.sg._ctr(4), 0, _n.stream_ptr())
        self.pull_epoch = self.epoch

    def check_overflow(self):
        """Raise if any plan since the last call found a receive buffer too small (one small D2H read)."""
        if not self._overflow_sources:
            return
        flags = torch.stack([t[self.E + 2 * self.L + 1] for t in self._overflow_sources]).max().reshape(1)
        self._overflow_sources = []
        if self.group is not None:
            # every rank must reach the same verdict, or the ranks that did not overflow wait forever for the one that raised
            torch.distributed.all_reduce(flags, op=torch.distributed.ReduceOp.MAX, group=self.group)
        need = int(flags.item())
        if need != 0:
            raise RuntimeError(
                f"expert-parallel receive buffer overflow: routing sent {need} rows (with alignment) to one rank, the "
                f"buffers hold {self.cap_rows}; raise the MoE layer's ep_capacity_factor to at least "
                f"{need / max(1, self.cap_rows - self.L * GROUP_ALIGN_M) * self.capacity_factor:.2f}")


class _Dispatch(torch.autograd.Function):
    """x [T, H] -> this rank's expert input rows [cap_rows, H] (expert-contiguous, groups aligned to 256 rows,
    padding rows zero)."""

    @staticmethod
    def forward(ctx, x, ep: FusedEP, plan: EPPlan):
        ep.push(x.contiguous(), plan)
        ctx.ep, ctx.plan = ep, plan
        return ep.recv_rows.clone()          # saved by the expert GEMM for its wgrad: must outlive the buffer

    @staticmethod
    def backward(ctx, d_rows):
        ep, plan = ctx.ep, ctx.plan
        ep.plan(None, plan.tables)           # handshake: return buffers of all ranks are free
        ep.ret_rows.copy_(d_rows)
        dx = torch.empty((plan.T, ep.H), dtype=d_rows.dtype, device=d_rows.device)
        ep.pull(dx, plan)
        return dx, None, None


class _Combine(torch.autograd.Function):
    """expert output rows (living in the symmetric return buffer) + routing weights -> y [T, H]."""

    @staticmethod
    def forward(ctx, rows, scores, ep: FusedEP, plan: EPPlan):
        if rows.data_ptr() != ep.ret_rows.data_ptr():
            ep.ret_rows.copy_(rows)
        y = torch.empty((plan.T, ep.H), dtype=rows.dtype, device=rows.device)
        keep = torch.empty((plan.T * plan.K, ep.H), dtype=rows.dtype, device=rows.device)
        scores = scores.contiguous()
        ep.pull(y, plan, weights=scores, keep=keep)
        ctx.ep, ctx.plan = ep, plan
        ctx.save_for_backward(keep, scores)
        return y

    @staticmethod
    def backward(ctx, dy):
        ep, plan = ctx.ep, ctx.plan
        keep, scores = ctx.saved_tensors
        ep.plan(None, plan.tables)           # handshake: receive buffers of all ranks are free
        d_scores = torch.empty_like(scores)
        ep.push(dy.contiguous(), plan, weights=scores, pulled=keep, d_scores=d_scores)
        # gradient w.r.t. the expert output rows: what the peers pushed into my receive buffer.  A view is enough: the
        # expert backward consumes it before this rank's next push, which the next handshake orders across ranks.
        return ep.recv_rows, d_scores, None, None


def dispatch(x, ep: FusedEP, plan: EPPlan):
    return _Dispatch.apply(x, ep, plan)


def combine(rows, scores, ep: FusedEP, plan: EPPlan):
    return _Combine.apply(rows, scores, ep, plan)


def make_plan(ep: FusedEP, counts, assign, offsets) -> EPPlan:
    tables = ep.plan(counts.contiguous())
    T, K = assign.shape
    return EPPlan(tables, assign.contiguous(), offsets.contiguous(), T, K, tables[ep.E: ep.E + ep.L + 1])


@torch.no_grad()
def grouped_experts_inference(x, scores, assign, offsets, counts, w1, w2, activation, top_k: int):
    """Inference MoE body without a host synchronisation: expert-contiguous rows through the plan / push kernels (one
    rank: every "peer" access is local), ONE grouped GEMM per projection over the 256-aligned groups, weighted
    un-permute through the pull kernel.  ``w1`` ``[E, N1, H]``, ``w2`` ``[E, H, F]`` (stacked expert weights),
    ``activation``: callable on the first projection's output (e.g. SwiGLU).  Replaces the per-expert loop over
    ``cumsum.tolist()`` (reference: ``inference/v2/kernels/cutlass_ops/moe_gemm``)."""
    from ..ops.gemm_ext import grouped_gemm_nt
    T, H = x.shape
    E = w1.shape[0]
    ep = FusedEP.get(None, H, E, T, top_k, 1.0)
    plan = make_plan(ep, counts, assign, offsets)
    ep.push(x.contiguous(), plan)
    h = activation(grouped_gemm_nt(ep.recv_rows, w1, plan.group_offsets))
    grouped_gemm_nt(h, w2, plan.group_offsets, out=ep.ret_rows)
    y = torch.empty((T, H), dtype=x.dtype, device=x.device)
    ep.pull(y, plan, weights=scores.to(torch.float32).contiguous())
    return y


def grouped_inference_ok(x, w1, w2, *biases) -> bool:
    return (x.is_cuda and x.dtype == torch.bfloat16 and w1.dtype == torch.bfloat16 and w2.dtype == torch.bfloat16
            and w1.dim() == 3 and w2.dim() == 3 and w1.is_contiguous() and w2.is_contiguous()
            and all(b is None for b in biases) and x.shape[1] % 64 == 0 and w1.shape[1] % 64 == 0)
