"""ZeRO-3 on the symmetric heap: the product path for parameter gather and gradient reduction.

When ``zero_optimization.fused_collectives`` is on (default: automatically when the data-parallel
group runs NCCL on CUDA with more than one rank and symmetric memory can be set up):

* every unit's low-precision shard lives in ONE symmetric arena; fetching a unit is the pull
  all-gather kernel (``sxb_symm_all_gather``) reading the peers' arena slices over NVLink and
  publishing per-chunk arrival flags — no NCCL all-gather, no staging ``torch.cat``
  (reference path: partition_parameters.py:1221-1261);
* every unit's flat gradient buffer is a slice of one of two rotating symmetric gradient arenas
  that the wgrad GEMMs write into directly; when the unit's last gradient lands, ONE kernel
  (``sxb_symm_rs_adam``) reduces this rank's shard across the peers (NVLS ``multimem.ld_reduce`` when
  the fabric provides multicast, peer loads otherwise), applies 1/n, the unscale/clip factor and
  AdamW on the fp32 master, and writes the bf16 shard straight back into the arena
  (reference path: coalesced_collectives.py:159-218 + stage3.py:1478-1540 + 2137-2152).
"""
from __future__ import annotations

import ctypes
from typing import List, Optional

import torch

from ...comm import symm
from ...utils.logging import logger


class SymmView:
    """A slice [offset, offset+numel) of a :class:`symm.SymmBuffer` with its own peer-pointer table."""

    def __init__(self, parent: symm.SymmBuffer, offset: int, numel: int):
        self.parent, self.offset, self.numel = parent, offset, numel
        es = parent.tensor.element_size()
        self.tensor = parent.tensor[offset: offset + numel]
        self._bufs = (ctypes.c_void_p * parent.world)(*[int(p) + offset * es for p in parent.handle.buffer_ptrs])
        self._sigs = parent._sigs
        self.multicast_ptr = parent.multicast_ptr + offset * es if parent.multicast_ptr else 0
        self.world, self.rank = parent.world, parent.rank

    @property
    def bufs(self):
        return ctypes.cast(self._bufs, ctypes.c_void_p)

    @property
    def sigs(self):
        return ctypes.cast(self._sigs, ctypes.c_void_p)

    @property
    def epochs(self):
        return self.parent.epochs

    def next_epoch(self, ch: int) -> int:
        return self.parent.next_epoch(ch)


class FusedZ3Comm:
    def __init__(self, optimizer, group):
        self.opt = optimizer
        self.sg = symm.SymmGroup.get(group)
        self.world, self.rank = self.sg.world, self.sg.rank
        groups = optimizer.groups
        dtype = optimizer.dtype
        # ---- shard arena -------------------------------------------------------------------------
        total_part = sum(fg.part for fg in groups)
        self.shards = symm.SymmBuffer(total_part, dtype, group)
        off = 0
        for fg in groups:
            view = SymmView(self.shards, off, fg.part)
            view.tensor.copy_(fg.lp_shard)
            fg.lp_shard = view.tensor
            fg._symm_shard = view
            off += fg.part
        # ---- two rotating gradient arenas sized for the largest unit ----------------------------------
        biggest = max(sum(fg.total for fg in u.groups) for u in optimizer.units)
        self.grad_arenas = [symm.SymmBuffer(biggest, dtype, group) for _ in range(2)]
        self._slot = 0
        self.use_multicast = bool(self.grad_arenas[0].multicast_ptr)
        logger.info(f"[rank {self.rank}] ZeRO-3 fused collectives: shard arena {total_part * 2 / 2**20:.0f} MiB, "
                    f"2 x {biggest * 2 / 2**20:.0f} MiB gradient arenas, NVLS multicast={self.use_multicast}")

    # ----------------------------------------------------------------------------------- params
    def gather(self, fg):
        """Full flat buffer of ``fg`` from the peers' shards (kernel-driven, on the current stream)."""
        flat = fg.allocate_full()
        self.sg.all_gather(fg._symm_shard, fg.part, flat, chunk_bytes=1 << 20, chunk_flags=None, max_ctas=48)
        fg._bind_params()

    # ------------------------------------------------------------------------------------ grads
    def attach_unit_grads(self, unit):
        """Give the unit a zeroed slice of the next gradient arena (after every peer finished reading
        that arena's previous occupant) and point the parameters' main-grad views into it."""
        arena = self.grad_arenas[self._slot]
        self._slot ^= 1
        self.sg.wait_readers(arena, symm.CH_GRADS_READ)
        off = 0
        for fg in unit.groups:
            view = SymmView(arena, off, fg.total)
            view.tensor.zero_()
            fg.flat_grad = view.tensor
            fg._symm_grad = view
            off += fg.total
            fg.attach_main_grads()
            fg.mark_grads_unwritten()

    def reduce_and_step(self, fg, fuse_step: bool, accumulate: bool, grad_scale: float = 1.0):
        """K1 for one flat group.  ``fuse_step``: update master + bf16 shard in the same kernel; otherwise
        write the reduced (mean) gradient shard for a deferred step."""
        opt = self.opt
        view = fg._symm_grad
        if fuse_step and not accumulate:
            st = opt.optimizer.flat_state(fg.hp_param)
            st["step"] += 1
            self.sg.rs_adam(view, fg.part, fg.hp_param.data, st["exp_avg"], st["exp_avg_sq"], fg.lp_shard,
                            fg.param_group, st["step"], adamw=bool(getattr(opt.optimizer, "adam_w_mode", 1)),
                            grad_scale=grad_scale, use_multicast=self.use_multicast)
            fg.grad_shard = None
            return True
        out = torch.empty(fg.part, dtype=fg.grad_dtype, device=fg.device)
        self.sg.rs_adam(view, fg.part, None, None, None, None, fg.param_group, 1, grad_out=out,
                        use_multicast=self.use_multicast)
        if accumulate and fg.grad_shard is not None:
            fg.grad_shard.add_(out)
        else:
            fg.grad_shard = out
        return False


def maybe_create(optimizer, group, requested: Optional[bool]):
    """Returns a FusedZ3Comm or None (NCCL path)."""
    if requested is False or optimizer.world <= 1 or optimizer.dtype != torch.bfloat16:
        return None
    if not symm.available(group) or optimizer.offload_optimizer:
        if requested:
            raise RuntimeError("fused_collectives requested but symmetric memory is unavailable")
        return None
    try:
        return FusedZ3Comm(optimizer, group)
    except Exception as e:  # pragma: no cover
        if requested:
            raise
        logger.warning(f"symmetric-memory setup failed ({e}); using the NCCL path")
        return None
