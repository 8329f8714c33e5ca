"""Ulysses sequence parallelism with the all-to-alls fused into the producing kernels (K6 / K7).

One attention block (QKV projection, RoPE, causal flash attention, output projection) over a sequence that is sharded
across ``P`` ranks (``S_local`` tokens each, batch 1):

    forward   x[S_local, H] @ Wqkv^T     GEMM whose EPILOGUE stores every head's columns straight into the rank that
                                         owns the head, at the token's global row  (K6: csrc/gemm_ext.cu,
                                         ``sxb_gemm_scatter`` — peer stores over NVLink, no all-to-all pass)
              RoPE (in place), flash attention over the FULL sequence on ``H/P`` heads, whose EPILOGUE stores every
                                         output row into the rank that owns that piece of the sequence (K7:
                                         ``sxb_attn_fwd_scatter``); a local copy of O stays for the backward
              ctx[S_local, H] @ Wo^T     plain GEMM
    backward  dy @ Wo                    GEMM with the same column-scatter epilogue (d ctx to the head owners)
              flash attention backward   dQ / dK / dV rows scattered back to the sequence owners by the epilogues
              inverse RoPE, wgrad / dgrad of the QKV projection: local

The reference issues three ``all_to_all_single`` for q, k, v, one for the context and their mirrors in backward around
library attention (``sequence/layer.py:331-440``).  Cross-rank ordering here: a barrier before peers' receive buffers
are overwritten and a signal + wait after the stores, all one-warp kernels on the compute stream
(``comm/symm.py``).  The receive buffers are symmetric memory shared by all layers.

``UlyssesPeers`` also has an *emulated* form (``world`` logical ranks on one GPU, run one after the other — none of the
kernels waits on another rank, only the stream-ordered sync kernels do, and those are no-ops here), which is what the
single-GPU tests use to check the multi-rank indexing.
"""
from __future__ import annotations

import ctypes
import math
from typing import List, Optional

import torch

from ..native import lib as _n
from ..ops import functional as F

_n.register({
    "sxb_gemm_scatter": "pp" "iiii" "ll" "pp" "i" "ll" "f" "i" "p",
    "sxb_attn_fwd_scatter": "ppp" "p" "ii" "pl" "p" "iiiii" "llll" "i" "f" "p",
    "sxb_attn_bwd_scatter": "ppppp" "pp" "pp" "l" "ii" "iiiii" "llllllll" "i" "f" "p",
})

CH_FREE, CH_LANDED = 12, 13


def _ptrs(tensors, offset_elems=0):
    arr = (ctypes.c_void_p * len(tensors))(*[t.data_ptr() + 2 * offset_elems for t in tensors])
    return arr, ctypes.cast(arr, ctypes.c_void_p)


class UlyssesPeers:
    """Receive buffers of one sequence-parallel group.

    ``group`` = a process group (symmetric memory, real ranks) or ``None`` with ``emulate_world`` logical ranks whose
    buffers all live on this GPU."""

    def __init__(self, group, S_local: int, Hq: int, Hk: int, D: int, emulate_world: int = 1,
                 dtype=torch.bfloat16, device=None):
        self.group = group
        if group is not None:
            from ..comm import symm
            self.sg = symm.SymmGroup.get(group)
            self.world, self.rank = self.sg.world, self.sg.rank
        else:
            self.sg, self.world, self.rank = None, emulate_world, None
        P = self.world
        assert Hq % P == 0 and Hk % P == 0, f"fused Ulysses needs head counts divisible by sp={P} (got {Hq}/{Hk})"
        self.S_local, self.S = S_local, S_local * P
        self.Hq, self.Hk, self.D = Hq, Hk, D
        self.hq, self.hk = Hq // P, Hk // P
        self.C_loc = (self.hq + 2 * self.hk) * D            # columns of the gathered qkv on one rank
        self.C_all = (Hq + 2 * Hk) * D
        dev = device or torch.device("cuda", torch.cuda.current_device())
        sizes = {"qkv": self.S * self.C_loc, "ctx": S_local * Hq * D, "dctx": self.S * self.hq * D,
                 "dqkv": S_local * self.C_all}
        self.bufs = {}
        if group is not None:
            from ..comm import symm
            for k, n in sizes.items():
                b = symm.SymmBuffer(n, dtype, group)
                self.bufs[k] = (b, [b.peer_tensor(r) for r in range(P)])
            self._sync_buf = self.bufs["qkv"][0]
        else:
            for k, n in sizes.items():
                self.bufs[k] = (None, [torch.zeros(n, dtype=dtype, device=dev) for _ in range(P)])
        self.epoch = 0
        # scatter tables: 64-column block -> (owner, destination column)
        self.table_qkv = self._table([(Hq, self.hq, 0), (Hk, self.hk, self.hq * D), (Hk, self.hk, (self.hq + self.hk) * D)], dev)
        self.table_q = self._table([(Hq, self.hq, 0)], dev)

    def _table(self, sections, dev):
        rows = []
        for heads, per, base in sections:
            for h in range(heads):
                for half in range(self.D // 64):
                    rows += [h // per, base + (h % per) * self.D + half * 64]
        return torch.tensor(rows, dtype=torch.int32, device=dev)

    def peers(self, name) -> List[torch.Tensor]:
        return self.bufs[name][1]

    # cross-rank ordering (no-ops in the emulated form: logical ranks run one after the other)
    def sync_free(self):
        """Every rank finished consuming what it received last time: its buffers may be overwritten."""
        if self.sg is None:
            return
        self.sg.barrier(self._sync_buf, CH_FREE)

    def sync_landed(self):
        if self.sg is None:
            return
        self.epoch += 1
        b = self._sync_buf
        _n.call("sxb_symm_signal", b.bufs, b.sigs, self.rank, self.world, CH_LANDED, self.epoch, _n.stream_ptr())
        _n.call("sxb_symm_wait", b.bufs, b.sigs, self.rank, self.world, CH_LANDED, self.epoch, _n.stream_ptr())


def gemm_scatter(a, b, layout, table, peers: List[torch.Tensor], dst_ld: int, row0: int):
    """``a op(b)`` with the column-scatter epilogue; result rows land at ``row0 + r`` of the owners' buffers."""
    if layout == F.GEMM_NT:
        M, K = a.shape
        N = b.shape[0]
    else:
        M, K = a.shape
        N = b.shape[1]
    keep, pp = _ptrs(peers)
    _n.call("sxb_gemm_scatter", a.data_ptr(), b.data_ptr(), M, N, K, layout, a.stride(0), b.stride(0),
            table.data_ptr(), pp, len(peers), int(dst_ld), int(row0), 1.0, 0, _n.stream_ptr())


class _Block(torch.autograd.Function):
    """Real-rank path: all phases in one autograd node."""

    @staticmethod
    def forward(ctx, x, wqkv, wo, cos, sin, up: UlyssesPeers, causal: bool):
        st = _Stage(up, up.rank, causal)
        st.fwd1(x, wqkv)
        up.sync_landed()
        up.sync_free()
        st.fwd2(cos, sin)
        up.sync_landed()
        out = st.fwd3(wo)
        F._note_use(wqkv)
        F._note_use(wo)
        ctx.st = st
        ctx.save_for_backward(x, wqkv, wo, cos, sin)
        return out

    @staticmethod
    def backward(ctx, dy):
        x, wqkv, wo, cos, sin = ctx.saved_tensors
        st, up = ctx.st, ctx.st.up
        dwo = st.bwd1(dy.contiguous(), wo)
        up.sync_landed()
        up.sync_free()
        st.bwd2()
        up.sync_landed()
        dx, dwqkv = st.bwd3(x, wqkv, cos, sin)
        return dx, dwqkv, dwo, None, None, None, None


class _Stage:
    """The block cut at its cross-rank synchronisation points, for one (logical) rank."""

    def __init__(self, up: UlyssesPeers, rank: int, causal: bool):
        self.up, self.rank, self.causal = up, rank, causal
        self.scale = 1.0 / math.sqrt(up.D)

    # ------------------------------------------------------------------------------- forward
    def fwd1(self, x, wqkv):
        up = self.up
        self.x = x.contiguous()
        up.sync_free()
        gemm_scatter(self.x, wqkv, F.GEMM_NT, up.table_qkv, up.peers("qkv"), up.C_loc, self.rank * up.S_local)

    def fwd2(self, cos, sin):
        up, r = self.up, self.rank
        S, D = up.S, up.D
        qkv = up.peers("qkv")[r].view(S, up.C_loc).clone()
        _n.call("sxb_rope", qkv.data_ptr(), None, cos.data_ptr(), sin.data_ptr(), S, up.hq + up.hk, D, up.C_loc, 0,
                _n.dtype_code(qkv.dtype), _n.stream_ptr())
        self.qkv = qkv
        self.o_keep = torch.empty((S, up.hq * D), dtype=qkv.dtype, device=qkv.device)
        self.lse = torch.empty((1, up.hq, S), dtype=torch.float32, device=qkv.device)
        q, k, v = self._qkv_views()
        kk, pp = _ptrs(up.peers("ctx"), r * up.hq * D)
        _n.call("sxb_attn_fwd_scatter", q.data_ptr(), k.data_ptr(), v.data_ptr(), pp, up.world, up.S_local,
                self.o_keep.data_ptr(), self.o_keep.stride(0), self.lse.data_ptr(), 1, S, up.hq, up.hk, D,
                qkv.stride(0), qkv.stride(0), qkv.stride(0), up.Hq * D, int(self.causal), float(self.scale),
                _n.stream_ptr())

    def _qkv_views(self):
        up, D = self.up, self.up.D
        return (self.qkv[:, : up.hq * D], self.qkv[:, up.hq * D: (up.hq + up.hk) * D],
                self.qkv[:, (up.hq + up.hk) * D:])

    def fwd3(self, wo):
        up = self.up
        self.ctx_local = up.peers("ctx")[self.rank].view(up.S_local, up.Hq * up.D).clone()
        return F.gemm(self.ctx_local, wo, F.GEMM_NT)

    # ------------------------------------------------------------------------------ backward
    def bwd1(self, dy, wo):
        up = self.up
        dwo = F._deliver_wgrad(wo, dy, self.ctx_local)
        up.sync_free()
        gemm_scatter(dy, wo, F.GEMM_NN, up.table_q, up.peers("dctx"), up.hq * up.D, self.rank * up.S_local)
        return dwo

    def bwd2(self):
        up, r, D = self.up, self.rank, self.up.D
        S = up.S
        dctx = up.peers("dctx")[r].view(S, up.hq * D)
        q, k, v = self._qkv_views()
        delta = torch.empty((1, up.hq, S), dtype=torch.float32, device=dctx.device)
        kq, pq = _ptrs(up.peers("dqkv"), r * up.hq * D)
        kk, pk = _ptrs(up.peers("dqkv"), up.Hq * D + r * up.hk * D)
        _n.call("sxb_attn_bwd_scatter", q.data_ptr(), k.data_ptr(), v.data_ptr(), self.o_keep.data_ptr(),
                dctx.data_ptr(), self.lse.data_ptr(), delta.data_ptr(), pq, pk, up.Hk * D, up.world, up.S_local,
                1, S, up.hq, up.hk, D, self.qkv.stride(0), self.qkv.stride(0), self.qkv.stride(0),
                self.o_keep.stride(0), dctx.stride(0), up.C_all, up.C_all, up.C_all, int(self.causal),
                float(self.scale), _n.stream_ptr())

    def bwd3(self, x, wqkv, cos, sin):
        up, r = self.up, self.rank
        dqkv = up.peers("dqkv")[r].view(up.S_local, up.C_all)
        pos = torch.arange(r * up.S_local, (r + 1) * up.S_local, device=dqkv.device, dtype=torch.long)
        _n.call("sxb_rope", dqkv.data_ptr(), pos.data_ptr(), cos.data_ptr(), sin.data_ptr(), up.S_local,
                up.Hq + up.Hk, up.D, up.C_all, 1, _n.dtype_code(dqkv.dtype), _n.stream_ptr())
        dw = F._deliver_wgrad(wqkv, dqkv, self.x)
        return F.gemm(dqkv, wqkv, F.GEMM_NN), dw


def ulysses_attention(x, wqkv, wo, cos, sin, up: UlyssesPeers, causal: bool = True):
    """Attention block of THIS rank (real sequence-parallel group)."""
    return _Block.apply(x, wqkv, wo, cos, sin, up, causal)


def ulysses_attention_emulated(xs: List[torch.Tensor], wqkv, wo, cos, sin, up: UlyssesPeers, causal: bool = True,
                               dys: Optional[List[torch.Tensor]] = None):
    """All logical ranks of an emulated group, stage by stage (each stage for every rank before the next stage — the
    order the cross-rank syncs would impose).  Returns ``(outs, dxs, dwqkv, dwo)`` (gradients only with ``dys``)."""
    P = up.world
    stages = [_Stage(up, r, causal) for r in range(P)]
    for r in range(P):
        stages[r].fwd1(xs[r], wqkv)
    for r in range(P):
        stages[r].fwd2(cos, sin)
    outs = [stages[r].fwd3(wo) for r in range(P)]
    if dys is None:
        return outs, None, None, None
    gwo = torch.zeros_like(wo, dtype=torch.float32)
    gwqkv = torch.zeros_like(wqkv, dtype=torch.float32)
    saved = (getattr(wo, "_sxb_main_grad", None), getattr(wqkv, "_sxb_main_grad", None))
    wo._sxb_main_grad, wqkv._sxb_main_grad = gwo, gwqkv
    wo._sxb_grad_written = wqkv._sxb_grad_written = True      # accumulate over the logical ranks
    try:
        for r in range(P):
            stages[r].bwd1(dys[r].contiguous(), wo)
        for r in range(P):
            stages[r].bwd2()
        dxs = [stages[r].bwd3(xs[r].contiguous(), wqkv, cos, sin)[0] for r in range(P)]
    finally:
        wo._sxb_main_grad, wqkv._sxb_main_grad = saved
        wo._sxb_grad_written = wqkv._sxb_grad_written = False
    return outs, dxs, gwqkv, gwo
