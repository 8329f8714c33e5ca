"""Emulated peers: N logical ranks on ONE GPU.

The peer-memory kernels only see pointer tables (``bufs[world]``, ``sigs[world]``) and a rank index, so the same
device code runs when all ``world`` buffers and signal pads live on one device.  The ranks of a collective wait on
one another's flags, and separate launches on one GPU carry no guarantee of running at the same time — so every
collective here is ONE cooperative grid (``blockIdx.y`` = logical rank, entry points ``sxb_symm_*_emu``): the
launch itself fails if the ``world x ctas`` CTAs cannot be co-resident.  "Peer" loads / stores are ordinary local
accesses; this exercises the flag protocol, the indexing and the arithmetic of every fused collective without a
second GPU.  It says nothing about NVLink ordering or NVLS (no multicast pointer here).

``tests/test_symm_emulated_gpu.py`` compares every kernel with its PyTorch statement in the ``gpu`` pytest tier."""
from __future__ import annotations

import ctypes
from typing import List, Optional, Sequence

import torch

from ..native import lib as _n
from . import symm as _symm      # registers the kernel signatures
from .symm import CH_A2A_FREE, CH_A2A_LANDED, CH_AG_READY, CH_AVG_READ, CH_AVG_READY, CH_GRADS_READ, CH_GRADS_READY

_n.register({
    "sxb_symm_rs_adam_emu": "ppil" "ppppp" "fffff" "iii" "f" "i" "p" "ii" "p",
    "sxb_symm_subset_mean_emu": "ppi" "pp" "pl" "i" "p" "ii" "p",
    "sxb_symm_all_gather_emu": "ppil" "pl" "p" "iii" "p",
    "sxb_symm_a2a_rows_emu": "ppi" "pppp" "ii" "p" "i" "p",
})

MAX_CTAS = 8          # per logical rank: world * MAX_CTAS CTAs run as one cooperative grid


def _ptr_array(tensors: Sequence[Optional[torch.Tensor]]):
    arr = (ctypes.c_void_p * len(tensors))(*[None if t is None else t.data_ptr() for t in tensors])
    return arr, ctypes.cast(arr, ctypes.c_void_p)


class FakePeerBuffer:
    """``world`` same-sized buffers + signal pads on one device, with the pointer tables the kernels expect."""

    def __init__(self, world: int, numel: int, dtype: torch.dtype, device=None):
        if not 1 <= world <= 16:
            raise ValueError("world must be in 1..16")
        dev = device or torch.device("cuda", torch.cuda.current_device())
        self.world, self.numel, self.dtype = world, numel, dtype
        self.tensors: List[torch.Tensor] = [torch.zeros(numel, dtype=dtype, device=dev) for _ in range(world)]
        self.pads: List[torch.Tensor] = [torch.zeros(16 * 16, dtype=torch.int32, device=dev) for _ in range(world)]
        self._bufs = (ctypes.c_void_p * world)(*[t.data_ptr() for t in self.tensors])
        self._sigs = (ctypes.c_void_p * world)(*[p.data_ptr() for p in self.pads])
        self.epochs = [0] * 16
        self.multicast_ptr = 0

    def next_epoch(self, ch: int) -> int:
        self.epochs[ch] += 1
        return self.epochs[ch]

    @property
    def bufs(self):
        return ctypes.cast(self._bufs, ctypes.c_void_p)

    @property
    def sigs(self):
        return ctypes.cast(self._sigs, ctypes.c_void_p)


class FakePeerGroup:
    """Runs each collective for all logical ranks as one cooperative launch on the current stream."""

    def __init__(self, world: int, device=None):
        self.world = world
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.counters = [torch.zeros(16, dtype=torch.int32, device=self.device) for _ in range(world)]

    def _ctrs(self, i: int):
        arr = (ctypes.c_void_p * self.world)(*[c.data_ptr() + 4 * i for c in self.counters])
        return arr, ctypes.cast(arr, ctypes.c_void_p)

    # K1: reduce-scatter (+ AdamW); per-rank optimizer tensors are lists indexed by logical rank
    def rs_adam(self, grads: FakePeerBuffer, part: int, hp, m, v, lp, cfg: dict, step: int, adamw: bool = True,
                grad_scale: float = 1.0, grad_out: Optional[List[torch.Tensor]] = None) -> None:
        epoch = grads.next_epoch(CH_GRADS_READY)
        grads.epochs[CH_GRADS_READ] = epoch
        b1, b2 = cfg["betas"]
        apply = grad_out is None
        none = [None] * self.world
        keep = [_ptr_array(x if x is not None else none) for x in (hp, m, v, lp, grad_out)]
        ck, cp = self._ctrs(0)
        _n.call("sxb_symm_rs_adam_emu", grads.bufs, grads.sigs, self.world, int(part),
                keep[0][1], keep[1][1], keep[2][1], keep[3][1], keep[4][1],
                float(cfg["lr"]), float(b1), float(b2), float(cfg["eps"]), float(cfg["weight_decay"]), int(step),
                int(adamw), int(bool(cfg.get("bias_correction", True))), float(grad_scale), int(epoch), cp,
                int(apply), MAX_CTAS, _n.stream_ptr())

    # K9: every rank averages over ITS ring (``rings``: list of member lists partitioning the ranks)
    def subset_mean(self, src: FakePeerBuffer, rings: List[List[int]], out: List[torch.Tensor], n: int,
                    shared: bool = False) -> None:
        epoch = src.next_epoch(CH_AVG_READY)
        src.epochs[CH_AVG_READ] = epoch
        ring_of = {r: ring for ring in rings for r in ring}
        table = (ctypes.c_int * (self.world * 16))()
        counts = (ctypes.c_int * self.world)()
        for r in range(self.world):
            counts[r] = len(ring_of[r])
            for j, mbr in enumerate(ring_of[r]):
                table[r * 16 + j] = mbr
        ok, op = _ptr_array(out)
        ck, cp = self._ctrs(6 if shared else 1)
        _n.call("sxb_symm_subset_mean_emu", src.bufs, src.sigs, self.world, ctypes.cast(table, ctypes.c_void_p),
                ctypes.cast(counts, ctypes.c_void_p), op, int(n), int(epoch), cp, MAX_CTAS, int(shared),
                _n.stream_ptr())

    # K2: all-gather of ``part_numel`` elements per rank (+ per-chunk arrival flags)
    def all_gather(self, shard: FakePeerBuffer, part_numel: int, dst: List[torch.Tensor], chunk_bytes: int = 1 << 16,
                   chunk_flags: Optional[List[torch.Tensor]] = None, flag_value: int = 0) -> None:
        epoch = shard.next_epoch(CH_AG_READY)
        es = dst[0].element_size()
        dk, dp = _ptr_array(dst)
        fk, fp = _ptr_array(chunk_flags) if chunk_flags is not None else (None, None)
        _n.call("sxb_symm_all_gather_emu", shard.bufs, shard.sigs, self.world, int(part_numel * es), dp,
                int(chunk_bytes), fp, int(epoch), int(flag_value), MAX_CTAS, _n.stream_ptr())

    # row all-to-all: tables[r][d] = (send_off, recv_off, count) of rank r towards destination d
    def a2a_rows(self, dst: FakePeerBuffer, srcs: List[torch.Tensor], send_off, recv_off, count, row_bytes: int) -> None:
        epoch = dst.next_epoch(CH_A2A_FREE)
        dst.epochs[CH_A2A_LANDED] = epoch
        W = self.world
        L = ctypes.c_longlong * (W * W)
        so = L(*[int(send_off[r][d]) for r in range(W) for d in range(W)])
        ro = L(*[int(recv_off[r][d]) for r in range(W) for d in range(W)])
        ct = L(*[int(count[r][d]) for r in range(W) for d in range(W)])
        sk, sp = _ptr_array(srcs)
        ck, cp = self._ctrs(2)
        _n.call("sxb_symm_a2a_rows_emu", dst.bufs, dst.sigs, W, sp, ctypes.cast(so, ctypes.c_void_p),
                ctypes.cast(ro, ctypes.c_void_p), ctypes.cast(ct, ctypes.c_void_p), int(row_bytes), int(epoch), cp,
                MAX_CTAS, _n.stream_ptr())


# ------------------------------------------------------------------------------------------------
# Expert-parallel exchange (csrc/moe_ep.cu) with all logical ranks on one GPU.
_n.register({
    "sxb_moe_ep_plan_emu": "ppi" "pp" "iii" "iii" "p",
    "sxb_moe_ep_push_emu": "ppi" "pppp" "ppp" "iiiii" "i" "i" "pi" "p",
    "sxb_moe_ep_pull_emu": "ppi" "ppp" "ppp" "iiiii" "i" "i" "pi" "p",
})


class FakeEP:
    """``world`` logical expert-parallel ranks: per-rank receive / return / count buffers and tables, and the three
    kernels of ``moe/fused_ep.py`` launched for all ranks at once."""

    def __init__(self, world: int, num_experts: int, hidden: int, cap_rows: int, device=None):
        dev = device or torch.device("cuda", torch.cuda.current_device())
        self.world, self.E, self.L, self.H, self.cap = world, num_experts, num_experts // world, hidden, cap_rows
        self.recv = FakePeerBuffer(world, cap_rows * hidden, torch.bfloat16, dev)
        self.ret = FakePeerBuffer(world, cap_rows * hidden, torch.bfloat16, dev)
        self.counts = FakePeerBuffer(world, world * num_experts, torch.int32, dev)
        self.ret.pads = [torch.zeros(32 * 16, dtype=torch.int32, device=dev) for _ in range(world)]
        self.ret._sigs = (ctypes.c_void_p * world)(*[p.data_ptr() for p in self.ret.pads])
        self.tables = [torch.zeros(num_experts + 2 * self.L + 2, dtype=torch.int32, device=dev) for _ in range(world)]
        self.counters = [torch.zeros(16, dtype=torch.int32, device=dev) for _ in range(world)]
        self.epoch = 0
        self.pull_epoch = 0

    def _ctrs(self, i):
        arr = (ctypes.c_void_p * self.world)(*[c.data_ptr() + 4 * i for c in self.counters])
        return arr, ctypes.cast(arr, ctypes.c_void_p)

    def plan(self, counts: Optional[List[torch.Tensor]]):
        self.epoch += 1
        ck, cp = _ptr_array(counts if counts is not None else [None] * self.world)
        tk, tp = _ptr_array(self.tables)
        _n.call("sxb_moe_ep_plan_emu", self.counts.bufs, self.ret.sigs, self.world, cp, tp, self.E, self.L, self.cap,
                self.epoch, self.pull_epoch, int(counts is not None), _n.stream_ptr())

    def push(self, x, assign, offsets, T, K, weights=None, pulled=None, d_scores=None):
        keep = [_ptr_array(v) if v is not None else (None, None) for v in (x, weights, pulled, d_scores, assign, offsets,
                                                                             self.tables)]
        ck, cp = self._ctrs(3)
        _n.call("sxb_moe_ep_push_emu", self.recv.bufs, self.ret.sigs, self.world, keep[0][1], keep[1][1], keep[2][1],
                keep[3][1], keep[4][1], keep[5][1], keep[6][1], T, K, self.H, self.E, self.L, self.cap, self.epoch, cp,
                MAX_CTAS, _n.stream_ptr())

    def pull(self, out, assign, offsets, T, K, weights=None, keep_rows=None):
        keep = [_ptr_array(v) if v is not None else (None, None) for v in (out, weights, keep_rows, assign, offsets,
                                                                             self.tables)]
        ck, cp = self._ctrs(4)
        _n.call("sxb_moe_ep_pull_emu", self.ret.bufs, self.ret.sigs, self.world, keep[0][1], keep[1][1], keep[2][1],
                keep[3][1], keep[4][1], keep[5][1], T, K, self.H, self.E, self.L, self.cap, self.epoch, cp, MAX_CTAS,
                _n.stream_ptr())
        self.pull_epoch = self.epoch
