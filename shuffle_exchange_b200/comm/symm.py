"""Symmetric heap + in-kernel NVLink collectives (the product path for the comm-bound hot ops).

A :class:`SymmBuffer` is a device allocation that every rank of a process group has mapped into its
own address space (``torch.distributed._symmetric_memory``: CUDA VMM handles exchanged over the
store, plus an NVLS multicast mapping when the fabric offers one).  The kernels in
``csrc/symm_collectives.cu`` receive the table of peer pointers and the peers' signal pads and do the
transfer themselves — loads/stores over NVLink 5 / NVSwitch issued from inside the kernel that also
does the adjacent math — instead of calling NCCL:

* :meth:`SymmGroup.rs_adam`        gradient reduce-scatter ⊕ scale ⊕ AdamW ⊕ bf16 shard write  (K1)
* :meth:`SymmGroup.subset_mean`    shuffle-exchange replica averaging over a ring given as data (K9)
* :meth:`SymmGroup.all_gather`     pull all-gather with per-chunk arrival flags for a consumer GEMM (K2)
* :meth:`SymmGroup.a2a_rows`       row all-to-all by direct peer stores (MoE / Ulysses transport)  (K4-K7)

The reference only touches symmetric memory in DeepCompile (deepspeed/comm/torch.py:416-422,
compile/init_z3.py:40-42); every collective on its hot paths is an NCCL call.
"""
from __future__ import annotations

import ctypes
from typing import Dict, List, Optional

import torch

from ..native import lib as _n

_n.register({
    "sxb_symm_rs_adam": "ppiil" "ppppp" "fffff" "iii" "f" "i" "pp" "iii" "p",
    "sxb_symm_wait": "ppiiiip",
    "sxb_symm_signal": "ppiiiip",
    "sxb_symm_barrier": "ppiiiip",
    "sxb_symm_subset_mean": "ppiipiplii" "pip",
    "sxb_symm_subset_wsum": "ppiippiplii" "pip",
    "sxb_symm_subset_wsum2": "ppiippiplii" "piip",
    "sxb_symm_all_gather": "ppiilplpiiip",
    "sxb_symm_a2a_rows": "ppiippppiipip",
    "sxb_symm_hang_info": "p",
    "sxb_register_foreign_range": "pl",
    "sxb_range_is_local": "pl",
    "sxb_attn_would_multicast": "pp" "iiiii" "ll",
})

import os as _os
_RS_CTAS = int(_os.environ.get("SXB_RS_CTAS", "148"))   # grid cap of the reduce-scatter+Adam kernel (1 CTA / SM)

CH_GRADS_READY, CH_GRADS_READ, CH_AVG_READY, CH_AVG_READ, CH_AG_READY, CH_A2A_FREE, CH_A2A_LANDED = range(7)


def available(group=None) -> bool:
    if not torch.cuda.is_available():
        return False
    try:
        import importlib
        importlib.import_module("torch.distributed._symmetric_memory")
        return torch.distributed.is_initialized() and torch.distributed.get_backend(group) == "nccl"
    except Exception:
        return False


class SymmBuffer:
    """One symmetric allocation + the pointer tables the kernels need."""

    def __init__(self, numel: int, dtype: torch.dtype, group, device=None):
        import torch.distributed._symmetric_memory as sm
        self.group = group if group is not None else torch.distributed.group.WORLD
        dev = device or torch.device("cuda", torch.cuda.current_device())
        self.tensor = sm.empty(numel, dtype=dtype, device=dev)
        self.handle = sm.rendezvous(self.tensor, self.group.group_name)
        self.rank, self.world = self.handle.rank, self.handle.world_size
        self._bufs = (ctypes.c_void_p * self.world)(*[int(p) for p in self.handle.buffer_ptrs])
        self._sigs = (ctypes.c_void_p * self.world)(*[int(p) for p in self.handle.signal_pad_ptrs])
        mc = getattr(self.handle, "multicast_ptr", 0) or 0
        self.multicast_ptr = int(mc)
        self.epochs = [0] * 16          # per-buffer, per-channel (signal pads belong to the buffer)
        # Every mapping of ANOTHER rank's memory (and the multicast window) is declared foreign to the kernel
        # library: TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from
        # another GPU's memory hung the GPU — kernels that would multicast check this registry and fall back.
        nbytes = int(numel) * self.tensor.element_size()
        for r, p in enumerate(self.handle.buffer_ptrs):
            if r != self.rank:
                _n.call("sxb_register_foreign_range", int(p), nbytes)
        if self.multicast_ptr:
            _n.call("sxb_register_foreign_range", self.multicast_ptr, nbytes)

    def next_epoch(self, ch: int) -> int:
        self.epochs[ch] += 1
        return self.epochs[ch]

    def peer_tensor(self, rank: int) -> torch.Tensor:
        """The whole buffer of ``rank`` as a tensor on THIS device's address space (NVLink peer mapping); copying
        from it with ``Tensor.copy_`` is a ``cudaMemcpyAsync`` served by the copy engines — no SMs involved."""
        cache = self.__dict__.setdefault("_peer_tensors", {})
        t = cache.get(rank)
        if t is None:
            t = cache[rank] = (self.tensor if rank == self.rank else
                               self.handle.get_buffer(rank, tuple(self.tensor.shape), self.tensor.dtype))
        return t

    @property
    def bufs(self):
        return ctypes.cast(self._bufs, ctypes.c_void_p)

    @property
    def sigs(self):
        return ctypes.cast(self._sigs, ctypes.c_void_p)


class SymmGroup:
    """Per-process-group state: epochs per channel, CTA counters, buffer cache."""

    _instances: Dict[int, "SymmGroup"] = {}

    def __init__(self, group=None):
        self.group = group if group is not None else torch.distributed.group.WORLD
        self.rank = torch.distributed.get_rank(self.group)
        self.world = torch.distributed.get_world_size(self.group)
        self.counter = torch.zeros(16, dtype=torch.int32, device="cuda")
        self.buffers: Dict[str, SymmBuffer] = {}

    @classmethod
    def get(cls, group=None) -> "SymmGroup":
        # keyed by the group's NAME (stable for the life of the process group; id() of a collected group object can be
        # re-used by a different group)
        key = getattr(group, "group_name", None) if group is not None else 0
        if key is None:
            key = id(group)
        inst = cls._instances.get(key)
        if inst is None:
            inst = cls._instances[key] = SymmGroup(group)
        return inst

    def buffer(self, name: str, numel: int, dtype: torch.dtype) -> SymmBuffer:
        """Named symmetric buffer, grown on demand (collective: all ranks must call identically)."""
        b = self.buffers.get(name)
        if b is None or b.tensor.numel() < numel or b.tensor.dtype != dtype:
            b = self.buffers[name] = SymmBuffer(numel, dtype, self.group)
        return b

    def _ctr(self, i: int) -> int:
        return self.counter.data_ptr() + 4 * i

    # ------------------------------------------------------------------------------------- K1
    def announce_grads(self, grads: SymmBuffer) -> int:
        """Tell every peer, on the CURRENT stream (the one that produced the gradients), that this rank's
        gradients in ``grads`` are complete.  Returns the epoch to hand to :meth:`rs_adam` — which then only waits
        for the peers with a one-warp kernel instead of spinning an SM-filling grid next to the backward GEMMs."""
        epoch = grads.next_epoch(CH_GRADS_READY)
        grads.epochs[CH_GRADS_READ] = epoch
        _n.call("sxb_symm_signal", grads.bufs, grads.sigs, self.rank, self.world, CH_GRADS_READY, int(epoch),
                _n.stream_ptr())
        return epoch

    def rs_adam(self, grads: SymmBuffer, part: int, hp, m, v, lp, group_cfg: dict, step: int, adamw: bool = True,
                grad_scale: float = 1.0, grad_out: Optional[torch.Tensor] = None, use_multicast: bool = True,
                max_ctas: int = 0, announced_epoch: Optional[int] = None, local_only: bool = False):
        """Reduce this rank's shard of every peer's flat bf16 gradient buffer and (unless ``grad_out`` is
        given) apply AdamW to ``hp/m/v`` writing the bf16 shard ``lp`` — one kernel, no NCCL.
        ``announced_epoch``: epoch returned by :meth:`announce_grads` (handshake split out of the kernel)."""
        bufs = grads.bufs
        if local_only:
            # measurement aid (bench.py exposed-communication figure): every "peer" buffer is this rank's own, no
            # handshake — the kernel does the same arithmetic and memory traffic with NVLink taken out
            epoch = grads.next_epoch(CH_GRADS_READY)
            use_multicast = False
            tbl = (ctypes.c_void_p * self.world)(*([ctypes.cast(grads.bufs, ctypes.POINTER(ctypes.c_void_p))[self.rank]]
                                                   * self.world))
            bufs = ctypes.cast(tbl, ctypes.c_void_p)
            announced_epoch = epoch
        elif announced_epoch is not None:
            epoch = announced_epoch
            _n.call("sxb_symm_wait", grads.bufs, grads.sigs, self.rank, self.world, CH_GRADS_READY, int(epoch),
                    _n.stream_ptr())
        else:
            epoch = grads.next_epoch(CH_GRADS_READY)
            grads.epochs[CH_GRADS_READ] = epoch
        b1, b2 = group_cfg["betas"]
        apply = grad_out is None
        mc = grads.multicast_ptr if use_multicast else 0
        _n.call("sxb_symm_rs_adam", bufs, grads.sigs, self.rank, self.world, int(part),
                _n.ptr(hp), _n.ptr(m), _n.ptr(v), _n.ptr(lp), _n.ptr(grad_out),
                float(group_cfg["lr"]), float(b1), float(b2), float(group_cfg["eps"]),
                float(group_cfg["weight_decay"]), int(step), int(adamw),
                int(bool(group_cfg.get("bias_correction", True))), float(grad_scale), int(epoch),
                mc if mc else None, self._ctr(0), int(apply), int(max_ctas or _RS_CTAS),
                int(announced_epoch is not None), _n.stream_ptr())
        return epoch

    def wait_readers(self, buf: SymmBuffer, channel: int = CH_GRADS_READ, epoch: Optional[int] = None):
        """Block the stream until every peer finished reading ``buf`` (before it is overwritten)."""
        e = buf.epochs[channel] if epoch is None else epoch
        if e > 0:
            _n.call("sxb_symm_wait", buf.bufs, buf.sigs, self.rank, self.world, int(channel), int(e), _n.stream_ptr())

    # ------------------------------------------------------------------------------------- K9
    def subset_mean(self, src: SymmBuffer, members: List[int], out: torch.Tensor, n: int, max_ctas: int = 0,
                    weights: Optional[List[float]] = None, shared: bool = False):
        """``out = sum_k weights[k] * shard(members[k])`` (``weights=None``: the mean) read straight out of the
        members' symmetric buffers.  Every rank of the group must call it in the same step (each with its own member
        list): the kernel handshakes with ALL ranks before reading and tells all of them when it is done.

        ``shared=True`` (every member passes the same list and weights — RR, H-RR): two-phase form, each member
        reduces one slice and the reduced slices are exchanged; ``src`` is overwritten with reduced values."""
        epoch = src.next_epoch(CH_AVG_READY)
        src.epochs[CH_AVG_READ] = epoch
        arr = (ctypes.c_int * len(members))(*members)
        warr = (ctypes.c_float * len(members))(*weights) if weights is not None else None
        two_phase = bool(shared) and len(members) > 1 and n % 8 == 0
        _n.call("sxb_symm_subset_wsum2", src.bufs, src.sigs, self.rank, self.world,
                ctypes.cast(arr, ctypes.c_void_p), ctypes.cast(warr, ctypes.c_void_p) if warr is not None else None,
                len(members), out.data_ptr(), int(n), _n.dtype_code(out.dtype), int(epoch),
                self._ctr(6 if two_phase else 1), int(max_ctas), int(two_phase), _n.stream_ptr())

    # ------------------------------------------------------------------------------------- K2
    def barrier(self, buf: SymmBuffer, channel: int = CH_AG_READY):
        """Stream-ordered barrier among all ranks of the group on ``buf``'s signal pad."""
        epoch = buf.next_epoch(channel)
        _n.call("sxb_symm_barrier", buf.bufs, buf.sigs, self.rank, self.world, int(channel), int(epoch),
                _n.stream_ptr())
        return epoch

    def all_gather(self, shard: SymmBuffer, part_numel: int, dst: torch.Tensor, chunk_bytes: int = 1 << 20,
                   chunk_flags: Optional[torch.Tensor] = None, max_ctas: int = 32, barrier: bool = True,
                   flag_value: int = 0):
        """``barrier=False``: pure pull, no cross-rank handshake (the shards were declared ready by an earlier
        :meth:`barrier` and do not change until the next one).  ``chunk_flags[c]`` is set to ``flag_value`` (release)
        when chunk ``c`` of ``dst`` is complete — the arrival signal the gated GEMM polls (K2)."""
        epoch = shard.next_epoch(CH_AG_READY) if barrier else 0
        es = dst.element_size()
        _n.call("sxb_symm_all_gather", shard.bufs, shard.sigs, self.rank, self.world, int(part_numel * es),
                dst.data_ptr(), int(chunk_bytes), _n.ptr(chunk_flags), int(epoch), int(max_ctas), int(flag_value),
                _n.stream_ptr())
        return epoch

    # ------------------------------------------------------------------------------- a2a rows
    def a2a_rows(self, dst: SymmBuffer, src: torch.Tensor, send_off, recv_off, count, row_bytes: int,
                 max_ctas: int = 0):
        """Store ``count[d]`` rows of ``src`` starting at ``send_off[d]`` into peer ``d``'s ``dst`` buffer
        at row ``recv_off[d]``; returns the epoch to pass to :meth:`wait_landed` on the receiving side."""
        epoch = dst.next_epoch(CH_A2A_FREE)
        dst.epochs[CH_A2A_LANDED] = epoch
        L = ctypes.c_longlong * self.world
        so, ro, ct = L(*send_off), L(*recv_off), L(*count)
        _n.call("sxb_symm_a2a_rows", dst.bufs, dst.sigs, self.rank, self.world, src.data_ptr(),
                ctypes.cast(so, ctypes.c_void_p), ctypes.cast(ro, ctypes.c_void_p), ctypes.cast(ct, ctypes.c_void_p),
                int(row_bytes), int(epoch), self._ctr(2), int(max_ctas), _n.stream_ptr())
        return epoch

    def wait_landed(self, buf: SymmBuffer, epoch: int):
        _n.call("sxb_symm_wait", buf.bufs, buf.sigs, self.rank, self.world, CH_A2A_LANDED, int(epoch), _n.stream_ptr())


class ReplicaAverageKernel:
    """Adapter used by ``runtime/zero/replica_averaging.py``: shard -> symmetric staging -> subset mean."""

    def __init__(self, group, numel: int, dtype: torch.dtype):
        self.sg = SymmGroup.get(group)
        self.buf = self.sg.buffer("replica_avg", numel, dtype)

    @classmethod
    def maybe_create(cls, group, shard: torch.Tensor):
        if not available(group) or shard.dtype not in (torch.bfloat16, torch.float16):
            return None
        try:
            return cls(group, shard.numel(), shard.dtype)
        except Exception:
            return None

    def fits(self, shard: torch.Tensor) -> bool:
        return shard.numel() <= self.buf.tensor.numel() and shard.dtype == self.buf.tensor.dtype

    def run(self, shard: torch.Tensor, members: List[int], weights: Optional[List[float]] = None,
            shared: bool = False):
        """``shared``: every member averages over the same list with the same weights (two-phase kernel)."""
        n = shard.numel()
        self.sg.wait_readers(self.buf, CH_AVG_READ)           # peers finished reading the previous contents
        self.buf.tensor[:n].copy_(shard.view(-1))
        self.sg.subset_mean(self.buf, members, shard.view(-1), n, weights=weights, shared=shared)
