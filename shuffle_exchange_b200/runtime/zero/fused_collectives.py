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
        # ---- gradient arenas ---------------------------------------------------------------------------
        # Units whose gradients stay open for the WHOLE backward (the root unit: LM-head wgrad first, embedding
        # scatter-add last) own a dedicated arena; every other unit takes the next of N rotating arenas sized for
        # the largest of them.  (A long-lived unit must never share a rotating arena: the units that cycle through
        # it would zero and overwrite its early gradients, and its late gradients would race with their reduction.)
        # Unit i of the rotation may only start once EVERY peer has finished reducing the arena's previous occupant
        # (unit i - N).  With N = 2 that couples the compute streams of all ranks one unit apart (measured: 98 ms of
        # exposed waiting per step at 8 GPUs); N = 4 leaves three units of slack.
        long_lived = [u for u in optimizer.units if self._is_long_lived(u)]
        rotating = [u for u in optimizer.units if not self._is_long_lived(u)]
        self.pinned_arenas = {u.index: symm.SymmBuffer(sum(fg.total for fg in u.groups), dtype, group)
                              for u in long_lived}
        biggest = max((sum(fg.total for fg in u.groups) for u in rotating), default=0)
        self.n_arenas = (4 if len(rotating) > 4 else 2) if rotating else 0
        self.grad_arenas = [symm.SymmBuffer(biggest, dtype, group) for _ in range(self.n_arenas)]
        self._slot = 0
        self._gate_seq = 0
        # measurement aids: local_mode serves every collective from this rank's own memory (no NVLink, no cross-rank
        # waits; results are meaningless) so that step_time - step_time(local_mode) = exposed communication;
        # timing collects CUDA-event pairs around every gather / reduce launch
        self.local_mode = False
        self.timing = None
        self._shards_dirty = True        # a step rewrote shards since the last cross-rank barrier
        import os
        self.gather_mode = os.environ.get("SXB_Z3_GATHER", "ce")   # "ce" (copy engines) | "sm" (pull kernel)
        any_arena = (self.grad_arenas + list(self.pinned_arenas.values()))[0]
        self.use_multicast = bool(any_arena.multicast_ptr)
        pinned_mib = sum(b.tensor.numel() for b in self.pinned_arenas.values()) * 2 / 2 ** 20
        logger.info(f"[rank {self.rank}] ZeRO-3 fused collectives: shard arena {total_part * 2 / 2**20:.0f} MiB, "
                    f"{self.n_arenas} x {biggest * 2 / 2**20:.0f} MiB rotating + {pinned_mib:.0f} MiB dedicated gradient "
                    f"arenas, NVLS multicast={self.use_multicast}")

    def _is_long_lived(self, u) -> bool:
        """Gradients of this unit are attached at the start of backward and reduced at its end."""
        return u is self.opt.root_unit or u.module is self.opt.module

    # ----------------------------------------------------------------------------------- params
    def shards_ready_barrier(self):
        """ONE cross-rank barrier per optimizer step (on the current stream): after it, every peer's updated shards
        are visible and stay unchanged until the next step, so all gathers of the step are barrier-free pulls.
        (A peer rewrites its shard of unit u only after every rank finished the backward of u — the gradient
        reduction of u cannot complete earlier — hence no gather of this step can still be reading it.)"""
        if self._shards_dirty:
            if not self.local_mode:
                self.sg.barrier(self.shards)
            self._shards_dirty = False

    def gather(self, fg, gated: bool = False):
        """Full flat buffer of ``fg`` from the peers' shards (on the current stream).

        ``gated`` (K2): every piece of the gather publishes an arrival flag behind itself and the unit's matrix
        parameters carry a :class:`~...ops.gemm_ext.Gate`; the GEMMs that consume them (``ops.functional.linear``,
        forward and dgrad) wait for the weight tiles they need INSIDE the kernel, so the compute stream does not
        wait for the gather as a whole.  The few non-matrix parameters (norm weights) are fetched first; the event
        returned for them is the only thing the compute stream waits on."""
        flat = fg.allocate_full()
        small_event = None
        t0 = self._tick()
        if gated:
            from ...ops.gemm_ext import Gate, stream_write_u32
            self._gate_seq += 1
            seq = self._gate_seq
            if self.gather_mode == "ce":
                chunk_bytes, nflags = fg.part * flat.element_size(), self.world
            else:
                chunk_bytes = 1 << 20
                nflags = (fg.total * flat.element_size() + chunk_bytes - 1) // chunk_bytes
            flags = getattr(fg, "_gate_flags", None)
            if flags is None or flags.numel() != nflags:
                flags = fg._gate_flags = torch.zeros(nflags, dtype=torch.int32, device=flat.device)
            view = fg._symm_shard
            # non-matrix parameters first: tiny copies out of whichever shards hold them
            for i, p in enumerate(fg.params):
                if len(fg.shapes[i]) >= 2:
                    continue
                lo, hi = fg.offsets[i], fg.offsets[i] + fg.numels[i]
                while lo < hi:
                    r = lo // fg.part
                    end = min(hi, (r + 1) * fg.part)
                    src = self.shards.peer_tensor(self.rank if self.local_mode else r)[
                        view.offset + lo - r * fg.part: view.offset + end - r * fg.part]
                    flat[lo:end].copy_(src, non_blocking=True)
                    lo = end
            small_event = torch.cuda.Event()
            small_event.record()
        if self.gather_mode == "ce":
            # copy-engine pull: world-1 peer-to-peer DMA copies + one local copy; leaves every SM to the GEMMs that
            # run concurrently on the compute stream (the SM pull kernel below costs them issue slots and L2)
            view = fg._symm_shard
            for i in range(self.world):
                r = (self.rank + i) % self.world if gated else (self.rank + 1 + i) % self.world
                src = self.shards.peer_tensor(self.rank if self.local_mode else r)[view.offset: view.offset + fg.part]
                flat[r * fg.part: (r + 1) * fg.part].copy_(src, non_blocking=True)
                if gated:
                    stream_write_u32(flags, r, seq)      # front-end write behind the copy: no SM involved
        else:
            self.sg.all_gather(fg._symm_shard, fg.part, flat, chunk_bytes=1 << 20,
                               chunk_flags=flags if gated else None, max_ctas=48, barrier=False,
                               flag_value=seq if gated else 0)
        self._tock("gather", t0, fg.total * flat.element_size() * (self.world - 1) // self.world)
        fg._bind_params()
        if gated:
            gate = Gate(flags, seq, chunk_bytes, flat.data_ptr())
            for i, p in enumerate(fg.params):
                p._sxb_gate = gate if len(fg.shapes[i]) >= 2 else None
        return small_event

    # ------------------------------------------------------------------------------------ grads
    def arena_key(self, unit):
        """Which arena ``attach_unit_grads(unit)`` will use next: ('pin', unit index) or ('rot', slot)."""
        if unit.index in self.pinned_arenas:
            return ("pin", unit.index)
        return ("rot", self._slot)

    def attach_unit_grads(self, unit):
        """Give the unit a slice of its gradient arena (after every peer finished reading the arena's
        previous occupant) and point the parameters' main-grad views into it."""
        if unit.index in self.pinned_arenas:
            arena = self.pinned_arenas[unit.index]
        else:
            arena = self.grad_arenas[self._slot]
            self._slot = (self._slot + 1) % self.n_arenas
        if not self.local_mode:
            self.sg.wait_readers(arena, symm.CH_GRADS_READ)
        off = 0
        for fg in unit.groups:
            view = SymmView(arena, off, fg.total)
            if fg.lazy_grad_zero:
                fg.zero_dead_ranges(view.tensor)   # parameters overwrite on first delivery; see ensure_flat_grad
            else:
                view.tensor.zero_()
            fg.flat_grad = view.tensor
            fg._symm_grad = view
            off += fg.total
            fg.attach_main_grads()
            fg.mark_grads_unwritten()

    def announce(self, fg) -> int:
        """On the compute stream, right behind the unit's last wgrad: publish "my gradients of this group are
        complete" to all peers (one-warp kernel).  The reduce stream then never holds SMs while a slower rank is
        still computing."""
        return None if self.local_mode else self.sg.announce_grads(fg._symm_grad)

    # ------------------------------------------------------------------------------ measurement
    def _tick(self):
        if self.timing is None:
            return None
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        return e

    def _tock(self, what, e0, nvlink_bytes):
        if e0 is None:
            return
        e1 = torch.cuda.Event(enable_timing=True)
        e1.record()
        self.timing.append((what, e0, e1, int(nvlink_bytes)))

    def timing_summary(self):
        """{what: (launches, summed ms, NVLink bytes received)} of the launches recorded since ``timing = []``."""
        torch.cuda.synchronize()
        out = {}
        for what, e0, e1, b in self.timing or []:
            n, ms, tot = out.get(what, (0, 0.0, 0))
            out[what] = (n + 1, ms + e0.elapsed_time(e1), tot + b)
        return out

    def reduce_and_step(self, fg, fuse_step: bool, accumulate: bool, grad_scale: float = 1.0,
                        announced_epoch: Optional[int] = None):
        """K1 for one flat group.  ``fuse_step``: update master + bf16 shard in the same kernel; otherwise
        write the reduced (mean) gradient shard for a deferred step."""
        opt = self.opt
        view = fg._symm_grad
        t0 = self._tick()
        # bytes this rank's reduction pulls over NVLink: its shard of every other rank's gradient buffer
        nvl = fg.part * 2 * (self.world - 1)
        if fuse_step and not accumulate:
            st = opt.optimizer.flat_state(fg.hp_param)
            st["step"] += 1
            self._shards_dirty = True
            self.sg.rs_adam(view, fg.part, fg.hp_param.data, st["exp_avg"], st["exp_avg_sq"], fg.lp_shard,
                            fg.param_group, st["step"], adamw=bool(getattr(opt.optimizer, "adam_w_mode", 1)),
                            grad_scale=grad_scale, use_multicast=self.use_multicast,
                            announced_epoch=announced_epoch, local_only=self.local_mode)
            self._tock("reduce_scatter_adam", t0, nvl)
            fg.grad_shard = None
            return True
        out = torch.empty(fg.part, dtype=fg.grad_dtype, device=fg.device)
        self.sg.rs_adam(view, fg.part, None, None, None, None, fg.param_group, 1, grad_out=out,
                        use_multicast=self.use_multicast, announced_epoch=announced_epoch, local_only=self.local_mode)
        self._tock("reduce_scatter", t0, nvl)
        if accumulate and fg.grad_shard is not None:
            fg.grad_shard.add_(out)
        else:
            fg.grad_shard = out
        return False


def maybe_create(optimizer, group, requested: Optional[bool]):
    """Returns a FusedZ3Comm or None (NCCL path)."""
    if requested is False or optimizer.world <= 1 or optimizer.dtype != torch.bfloat16:
        return None
    zeropp = (getattr(optimizer, "hpz_size", 1) > 1 or getattr(optimizer, "quantized_weights", False)
              or getattr(optimizer, "quantized_gradients", False) or getattr(optimizer, "offload_param", False))
    if not symm.available(group) or optimizer.offload_optimizer or zeropp:
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
