"""FlatGroup — the storage unit shared by every sharded optimizer in this framework.

A FlatGroup owns a list of parameters laid out back to back (each start aligned) in ONE flat
low-precision buffer that is cut into ``world`` equal partitions:

    flat_lp   [world * part]   bf16/fp16 — parameters are views into it (may be released under ZeRO-3)
    lp_shard  [part]           this rank's slice of flat_lp (always resident)
    hp_shard  [part]           fp32 master copy of lp_shard (the tensor the base optimizer steps)
    flat_grad [world * part]   gradients; parameters' main-grad views point here (transient)
    grad_shard[part]           reduced gradient of this rank's slice (comm dtype or fp32 accumulator)

ZeRO-1/2 use one FlatGroup per optimizer param group (reference: flatten + ``2*dp`` alignment,
deepspeed/runtime/zero/stage_1_and_2.py:462-472, partitions :1822-1840, fp32 master :507-534);
ZeRO-3 uses one per parameter "unit" (reference sub-groups/defragmented partitions,
stage3.py:720-801, 849-983); BF16/FP16 optimizers use it with world == dp or 1.  Because partition
``k`` is the contiguous slice ``k`` of the flat buffers, gradient reduction is a single
``reduce_scatter_tensor(grad_shard, flat_grad)`` and parameter refresh a single in-place
``all_gather_into_tensor(flat_lp, lp_shard)`` — or their fused symmetric-memory kernels.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional

import os

import torch

from ...comm import comm as dist

ALIGN = 64  # elements: keeps every parameter 128-byte aligned for TMA/vector access


def _align(n: int, a: int) -> int:
    return (n + a - 1) // a * a


_POISON_RELEASED = os.environ.get("SXB_DEBUG_POISON", "0") == "1"

class _GradHolder:
    """Carrier of ``.grad`` and the LoCo error-feedback buffers for one flat group."""
    grad = None


class FlatGroup:
    def __init__(self, params: List[torch.nn.Parameter], names: List[str], world: int, rank: int, process_group,
                 device: torch.device, lp_dtype: torch.dtype, grad_dtype: Optional[torch.dtype] = None,
                 group_index: int = 0, param_group: Optional[dict] = None, keep_full_params: bool = True,
                 hp_on_cpu: bool = False, pin: bool = False, allocator=None, lp_on_cpu: bool = False,
                 lp_store=None):
        self._lp_store, self._lp_shard = None, None       # see the ``lp_shard`` property
        self._lp_key = f"group{group_index:04d}"
        self.params = list(params)
        self.names = list(names)
        self.world, self.rank, self.pg = world, rank, process_group
        self.device, self.lp_dtype = device, lp_dtype
        self.grad_dtype = grad_dtype or lp_dtype
        self.index = group_index
        self.param_group = param_group
        self.keep_full_params = keep_full_params
        self.hp_on_cpu = hp_on_cpu

        # ---- layout -------------------------------------------------------------------------
        self.offsets: List[int] = []
        self.numels: List[int] = []
        cur = 0
        for p in self.params:
            n = getattr(p, "ds_numel", p.numel())
            self.offsets.append(cur)
            self.numels.append(n)
            cur = _align(cur + n, ALIGN)
        self.unpadded_numel = (self.offsets[-1] + self.numels[-1]) if self.params else 0
        self.total = _align(max(cur, 1), ALIGN * world)
        self.part = self.total // world
        self.padding = self.total - self.unpadded_numel
        self.shapes = [tuple(getattr(p, "ds_shape", p.shape)) for p in self.params]

        # ---- low precision storage --------------------------------------------------------------
        alloc = allocator or (lambda n, dt: torch.zeros(n, dtype=dt, device=device))
        self.flat_lp: Optional[torch.Tensor] = alloc(self.total, lp_dtype)
        for p, off, n, shp in zip(self.params, self.offsets, self.numels, self.shapes):
            if p.device.type == "meta" or p.numel() == 0:
                continue
            self.flat_lp[off: off + n].copy_(p.data.reshape(-1).to(device=device, dtype=lp_dtype))
        self.lp_shard = self.flat_lp[rank * self.part: (rank + 1) * self.part]
        if not keep_full_params:
            self.lp_shard = self.lp_shard.clone()
        if lp_on_cpu and not keep_full_params:
            # offload_param = cpu: the low-precision shard lives in (pinned) host memory between uses
            host = self.lp_shard.to("cpu")
            self.lp_shard = host.pin_memory() if (pin and torch.cuda.is_available()) else host
            if lp_store is not None:
                # offload_param = nvme: the shard's home is a file; it passes through the store's host buffers
                lp_store.put(self._lp_key, host)
                self._lp_shard, self._lp_store = None, lp_store
        self._bind_params()

        # ---- fp32 master ---------------------------------------------------------------------
        hp_dev = torch.device("cpu") if hp_on_cpu else device
        self.hp_shard = self.lp_shard.detach().to(device=hp_dev, dtype=torch.float32).clone()
        if hp_on_cpu and pin and torch.cuda.is_available():
            self.hp_shard = self.hp_shard.pin_memory()
        self.hp_shard.requires_grad_(False)

        self.flat_grad: Optional[torch.Tensor] = None
        self.grad_shard: Optional[torch.Tensor] = None
        self.grads_ready = 0
        self.micro_steps_accumulated = 0
        self.trainable = [p.requires_grad for p in self.params]
        self.n_trainable = sum(self.trainable)
        if not keep_full_params:
            self.release_params()

    # ------------------------------------------------------------------------------ parameters
    @property
    def lp_shard(self) -> Optional[torch.Tensor]:
        """This rank's low-precision shard.  With an NVMe store (``offload_param.device = nvme``) every access goes
        through the store's resident buffers; whoever modifies the returned tensor calls :meth:`lp_shard_written`."""
        if self._lp_store is not None:
            return self._lp_store.get(self._lp_key)
        return self._lp_shard

    @lp_shard.setter
    def lp_shard(self, t: Optional[torch.Tensor]):
        if self._lp_store is not None and t is not None:
            self._lp_store.put(self._lp_key, t)
        else:
            self._lp_shard = t

    def lp_shard_written(self):
        if self._lp_store is not None:
            self._lp_store.written(self._lp_key)

    def _bind_params(self):
        for p, off, n, shp in zip(self.params, self.offsets, self.numels, self.shapes):
            p.data = self.flat_lp[off: off + n].view(shp)

    def params_available(self) -> bool:
        return self.flat_lp is not None

    def allocate_full(self, allocator=None):
        if self.flat_lp is None:
            self.flat_lp = (allocator(self.total, self.lp_dtype) if allocator else
                            torch.empty(self.total, dtype=self.lp_dtype, device=self.device))
        return self.flat_lp

    # ZeRO++ / ZeRO-Infinity switches of the NCCL gather path (set by the stage-3 optimizer from the config):
    quantized_gradients = None    # qgZ: (all-to-all groups dict, LoCo parameters or None)
    quantized_weights = False     # qwZ: int8 group-quantised all-gather (reference partition_parameters.py:1320-1460)
    hpz_group = None              # hpZ: secondary shards inside this sub-group serve the backward gathers (:1747-1797)
    hpz_secondary = None

    def _device_shard(self) -> torch.Tensor:
        """This rank's low-precision shard on the compute device (``offload_param``: staged from pinned host memory)."""
        if self.lp_shard.device == self.device:
            return self.lp_shard
        return self.lp_shard.to(self.device, non_blocking=True)

    def gather_params(self, async_op: bool = False, allocator=None, backward: bool = False):
        """Materialise the full flat buffer from the shards and re-point the parameters."""
        if self.flat_lp is not None and self.keep_full_params:
            return None
        flat = self.allocate_full(allocator)
        handle = None
        if self.world == 1:
            flat.copy_(self._device_shard())
        elif backward and self.hpz_secondary is not None:
            # hpZ: the backward pass re-gathers from the secondary partition kept inside the (intra-node) sub-group
            handle = dist.all_gather_into_tensor(flat, self.hpz_secondary, group=self.hpz_group, async_op=async_op)
        elif self.quantized_weights:
            from ...ops.quantizer import CUDAQuantizer
            qz = CUDAQuantizer()
            q, s = qz.quantize(self._device_shard())
            qa = torch.empty((self.world,) + tuple(q.shape), dtype=q.dtype, device=q.device)
            sa = torch.empty((self.world,) + tuple(s.shape), dtype=s.dtype, device=s.device)
            dist.all_gather_into_tensor(qa.view(-1), q.reshape(-1), group=self.pg)
            dist.all_gather_into_tensor(sa.view(-1), s.reshape(-1), group=self.pg)
            for r in range(self.world):
                flat[r * self.part: (r + 1) * self.part].copy_(qz.dequantize(qa[r], sa[r], self.lp_dtype).reshape(-1))
            # this rank's own slice stays exact
            flat[self.rank * self.part: (self.rank + 1) * self.part].copy_(self._device_shard())
        else:
            handle = dist.all_gather_into_tensor(flat, self._device_shard(), group=self.pg, async_op=async_op)
        if self.hpz_group is not None and not backward and self.hpz_secondary is None:
            if handle is not None:
                handle.wait()
                handle = None
            k = dist.get_world_size(self.hpz_group)
            r = dist.get_rank(self.hpz_group)
            sec = self.total // k
            self.hpz_secondary = flat[r * sec: (r + 1) * sec].clone()
        self._bind_params()
        return handle

    def release_params(self):
        """Free the gathered buffer (ZeRO-3); parameters become empty placeholders."""
        if self.keep_full_params:
            return
        if _POISON_RELEASED and self.flat_lp is not None and self.flat_lp.is_floating_point():
            # debug build (SXB_DEBUG_POISON=1): a released gathered buffer is overwritten with NaN before it goes back
            # to the allocator, so a kernel that still reads it (missing stream wait, stale view) poisons the loss
            # instead of silently using old-but-plausible weights
            self.flat_lp.fill_(float("nan"))
        self.flat_lp = None
        for p in self.params:
            p.data = torch.empty(0, dtype=self.lp_dtype, device=self.device)
            if getattr(p, "_sxb_gate", None) is not None:
                p._sxb_gate = None

    def refresh_full_from_shard(self, async_op: bool = False):
        """Post-step parameter refresh for ZeRO-1/2: in-place all-gather of the updated shards."""
        if self.world == 1 or self.flat_lp is None:
            return None
        return dist.all_gather_into_tensor(self.flat_lp, self.lp_shard, group=self.pg, async_op=async_op)

    # ------------------------------------------------------------------------------- gradients
    def ensure_flat_grad(self, zero: bool = True):
        """The unit's flat gradient buffer.  With ``lazy_grad_zero`` it is NOT filled with zeros: every gradient producer overwrites on its
        first delivery (``_sxb_grad_written``), so only the ranges no trainable parameter covers are cleared here and
        parameters that never received a gradient are cleared right before the reduction
        (:meth:`zero_unwritten_grads`) — a full fill of a Llama-3-8B layer unit cost 0.17 ms, 6 ms per step."""
        if self.flat_grad is None:
            if zero and not self.lazy_grad_zero:
                self.flat_grad = torch.zeros(self.total, dtype=self.grad_dtype, device=self.device)
            else:
                self.flat_grad = torch.empty(self.total, dtype=self.grad_dtype, device=self.device)
                if zero:
                    self.zero_dead_ranges(self.flat_grad)
        return self.flat_grad

    # Opt-in (ZeRO-3 sets it): optimizers whose every reduction path calls ``zero_unwritten_grads`` first.
    lazy_grad_zero = False

    def dead_ranges(self):
        """Element ranges of the flat layout that no trainable parameter covers: alignment gaps, frozen parameters and
        the tail padding."""
        dead = self.__dict__.get("_dead_ranges")
        if dead is None:
            dead, pos = [], 0
            for i in sorted(range(len(self.params)), key=lambda j: self.offsets[j]):
                if not self.trainable[i]:
                    continue
                if self.offsets[i] > pos:
                    dead.append((pos, self.offsets[i]))
                pos = max(pos, self.offsets[i] + self.numels[i])
            if pos < self.total:
                dead.append((pos, self.total))
            self._dead_ranges = dead
        return dead

    def zero_dead_ranges(self, buf: torch.Tensor):
        for a, b in self.dead_ranges():
            buf[a:b].zero_()

    def zero_unwritten_grads(self):
        """Before the buffer is reduced: a trainable parameter nothing delivered a gradient for contributes zeros."""
        if self.flat_grad is None or not self.lazy_grad_zero:
            return
        for i, p in enumerate(self.params):
            if self.trainable[i] and not getattr(p, "_sxb_grad_written", False):
                self.flat_grad[self.offsets[i]: self.offsets[i] + self.numels[i]].zero_()

    def grad_view(self, i: int) -> torch.Tensor:
        g = self.ensure_flat_grad()
        return g[self.offsets[i]: self.offsets[i] + self.numels[i]].view(self.shapes[i])

    def attach_main_grads(self):
        """Point every trainable parameter's main-grad view into the flat gradient buffer."""
        self.ensure_flat_grad()
        for i, p in enumerate(self.params):
            if self.trainable[i]:
                p._sxb_main_grad = self.grad_view(i)
                if not hasattr(p, "_sxb_grad_written"):
                    p._sxb_grad_written = False

    def detach_main_grads(self):
        for p in self.params:
            if hasattr(p, "_sxb_main_grad"):
                p._sxb_main_grad = None
                p._sxb_grad_written = False

    def collect_autograd_grads(self):
        """Copy ``p.grad`` (produced by ops that do not know the main-grad protocol) into the flat
        buffer, accumulating onto anything already written there, and drop ``p.grad``."""
        for i, p in enumerate(self.params):
            if p.grad is None:
                continue
            gv = self.grad_view(i)
            if getattr(p, "_sxb_grad_written", False):
                gv.add_(p.grad.reshape(gv.shape).to(gv.dtype))
            else:
                gv.copy_(p.grad.reshape(gv.shape))
            p._sxb_grad_written = True
            p.grad = None

    def mark_grads_unwritten(self):
        """After a reduction the next micro-step may overwrite instead of accumulate."""
        for p in self.params:
            p._sxb_grad_written = False

    def reduce_grads(self, predivide: float = 1.0, accumulate_into_shard: bool = False, async_op: bool = False,
                     average: bool = True):
        """flat_grad --(reduce-scatter, mean over the group)--> grad_shard."""
        assert self.flat_grad is not None
        self.zero_unwritten_grads()
        if self.grad_shard is None or not accumulate_into_shard:
            if self.grad_shard is None:
                self.grad_shard = torch.zeros(self.part, dtype=self.grad_dtype, device=self.device)
        if self.world == 1:
            src = self.flat_grad[: self.part]
            if accumulate_into_shard:
                self.grad_shard.add_(src)
            else:
                self.grad_shard.copy_(src)
            return None
        if predivide != 1.0:
            self.flat_grad.div_(predivide)
        if self.quantized_gradients is not None and average and predivide == 1.0:
            # qgZ (ZeRO++): int4 hierarchical all-to-all instead of the reduce-scatter; with LoCo error feedback when
            # the config carries ``zeropp_loco_param`` (reference stage3.py:1388-1417, coalesced_collectives.py:30-156)
            from ..comm import coalesced_collectives as cc
            groups, loco = self.quantized_gradients
            cols = 2048 if self.total % 2048 == 0 else self.part
            t2d = self.flat_grad.view(-1, cols)
            if loco is not None:
                holder = self.__dict__.setdefault("_loco_holder", _GradHolder())
                holder.grad = t2d
                red = cc.all_to_all_loco_quant_reduce([holder], groups, loco)[0]
            else:
                red = cc.all_to_all_quant_reduce([t2d], groups)[0]
            red = red.reshape(-1).to(self.grad_dtype)
            if accumulate_into_shard:
                self.grad_shard.add_(red)
            else:
                self.grad_shard.copy_(red)
            return None
        tmp = torch.empty_like(self.grad_shard) if accumulate_into_shard else self.grad_shard
        op = dist.ReduceOp.AVG if (average and dist.get_backend(self.pg) == "nccl" and predivide == 1.0) else dist.ReduceOp.SUM
        h = dist.reduce_scatter_tensor(tmp, self.flat_grad, op=op, group=self.pg, async_op=False)
        if op == dist.ReduceOp.SUM and average:
            tmp.mul_(predivide / self.world)
        if accumulate_into_shard:
            self.grad_shard.add_(tmp)
        return h

    def release_flat_grad(self):
        self.flat_grad = None
        self.detach_main_grads()

    # --------------------------------------------------------------------------------- lookups
    def shard_range_of_param(self, i: int):
        """Intersection of parameter ``i`` with this rank's partition: (start_in_param, start_in_shard, numel)."""
        lo, hi = self.rank * self.part, (self.rank + 1) * self.part
        p_lo, p_hi = self.offsets[i], self.offsets[i] + self.numels[i]
        a, b = max(lo, p_lo), min(hi, p_hi)
        if a >= b:
            return None
        return a - p_lo, a - lo, b - a

    def param_slice_mappings(self) -> Dict[str, tuple]:
        """name -> (start, numel) fragment of the fp32 partition (reference key ``param_slice_mappings``)."""
        out = {}
        for i, name in enumerate(self.names):
            r = self.shard_range_of_param(i)
            if r is not None:
                out[name] = (r[1], r[2])
        return out
