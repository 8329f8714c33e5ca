""""Fake peer" mode: N logical ranks on ONE GPU.

The peer-memory kernels only see pointer tables (``bufs[world]``, ``sigs[world]``) and a rank index, so the same
kernels run unchanged when all ``world`` buffers and signal pads live on one device: rank ``r``'s kernel is launched
on its own stream with a small grid (so the ``world`` grids are co-resident — they spin on each other's flags and
would deadlock if one could not be scheduled), and "peer" loads / stores are ordinary local accesses.  This exercises
the flag protocol, the indexing and the arithmetic of every fused collective without a second GPU; it says nothing
about NVLink ordering or NVLS (no multicast pointer here).

Use from a single process on a CUDA device; ``tools/symm_fake_peer_check.py`` compares every kernel with its PyTorch
statement.  (Not yet part of the pytest ``gpu`` tier: written after this round's GPU time was spent.)"""
from __future__ import annotations

import ctypes
from typing import List, Optional

import torch

from ..native import lib as _n
from . import symm as _symm      # registers the kernel signatures
from .symm import CH_A2A_FREE, CH_A2A_LANDED, CH_AG_READY, CH_AVG_READ, CH_AVG_READY, CH_GRADS_READ, CH_GRADS_READY

MAX_CTAS = 8          # per logical rank: world * MAX_CTAS CTAs must be resident together


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
    """Launches one kernel per logical rank, each on its own stream, then joins them."""

    def __init__(self, world: int, device=None):
        self.world = world
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.streams = [torch.cuda.Stream(device=self.device) for _ in range(world)]
        self.counters = torch.zeros(world, 16, dtype=torch.int32, device=self.device)

    def _each_rank(self, fn) -> None:
        cur = torch.cuda.current_stream()
        for r, s in enumerate(self.streams):
            s.wait_stream(cur)
            with torch.cuda.stream(s):
                fn(r)
        for s in self.streams:
            cur.wait_stream(s)

    def _ctr(self, rank: int, i: int) -> int:
        return self.counters.data_ptr() + 4 * (rank * 16 + i)

    # K1: reduce-scatter (+ AdamW); per-rank optimizer tensors are lists indexed by logical rank
    def rs_adam(self, grads: FakePeerBuffer, part: int, hp, m, v, lp, cfg: dict, step: int, adamw: bool = True,
                grad_scale: float = 1.0, grad_out: Optional[List[torch.Tensor]] = None) -> None:
        epoch = grads.next_epoch(CH_GRADS_READY)
        grads.epochs[CH_GRADS_READ] = epoch
        b1, b2 = cfg["betas"]
        apply = grad_out is None

        def launch(r):
            _n.call("sxb_symm_rs_adam", grads.bufs, grads.sigs, r, self.world, int(part),
                    _n.ptr(hp[r]) if apply else None, _n.ptr(m[r]) if apply else None, _n.ptr(v[r]) if apply else None,
                    _n.ptr(lp[r]) if apply else None, None if apply else _n.ptr(grad_out[r]),
                    float(cfg["lr"]), float(b1), float(b2), float(cfg["eps"]), float(cfg["weight_decay"]), int(step),
                    int(adamw), int(bool(cfg.get("bias_correction", True))), float(grad_scale), int(epoch), None,
                    self._ctr(r, 0), int(apply), MAX_CTAS, _n.stream_ptr())
        self._each_rank(launch)

    # K9: mean over a subset of ranks
    def subset_mean(self, src: FakePeerBuffer, members: List[int], out: List[torch.Tensor], n: int) -> None:
        epoch = src.next_epoch(CH_AVG_READY)
        src.epochs[CH_AVG_READ] = epoch
        arr = (ctypes.c_int * len(members))(*members)

        def launch(r):
            if r in members:
                _n.call("sxb_symm_subset_mean", src.bufs, src.sigs, r, self.world, ctypes.cast(arr, ctypes.c_void_p),
                        len(members), out[r].data_ptr(), int(n), _n.dtype_code(out[r].dtype), int(epoch),
                        self._ctr(r, 1), MAX_CTAS, _n.stream_ptr())
        self._each_rank(launch)

    # K2: all-gather of ``part_numel`` elements per rank
    def all_gather(self, shard: FakePeerBuffer, part_numel: int, dst: List[torch.Tensor]) -> None:
        epoch = shard.next_epoch(CH_AG_READY)
        es = dst[0].element_size()

        def launch(r):
            _n.call("sxb_symm_all_gather", shard.bufs, shard.sigs, r, self.world, int(part_numel * es),
                    dst[r].data_ptr(), 1 << 16, None, int(epoch), MAX_CTAS, _n.stream_ptr())
        self._each_rank(launch)

    def barrier(self, buf: FakePeerBuffer, channel: int = CH_AG_READY) -> None:
        epoch = buf.next_epoch(channel)
        self._each_rank(lambda r: _n.call("sxb_symm_barrier", buf.bufs, buf.sigs, r, self.world, int(channel),
                                           int(epoch), _n.stream_ptr()))
