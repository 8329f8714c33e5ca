"""Activation checkpointing.

Capabilities of the reference module (deepspeed/runtime/activation_checkpointing/checkpointing.py:
``CheckpointFunction`` :488-702, partitioned activations :377-430 with backward all-gather :266-303,
CPU checkpointing, contiguous buffers, ``CudaRNGStatesTracker`` :124-194, ``non_reentrant_checkpoint``
:704-946, ``configure`` :1029), rebuilt around one small state object:

* ``checkpoint(fn, *args)``            reentrant recompute with RNG replay;
* ``non_reentrant_checkpoint(fn, *a)`` saved-tensor-hook based variant (works with ZeRO-3 units and
  with inputs that do not require grad);
* ``partition_activations``            each tensor-parallel rank keeps 1/tp of every saved input and
  re-assembles it with one ``all_gather_into_tensor`` in backward;
* ``cpu_checkpointing``                saved inputs live in pinned host memory between fwd and bwd;
* ``contiguous_memory_optimization``   saved partitions are packed into one preallocated arena per
  checkpoint slot (``number_checkpoints`` slots), avoiding allocator fragmentation at 128K context.
"""
from __future__ import annotations

import contextlib
from typing import Any, List, Optional

import torch

from ...comm import comm as dist
from ...device import get_accelerator
from ...utils.logging import logger
from ...utils.timer import SynchronizedWallClockTimer

_MODEL_PARALLEL_RNG_TRACKER_NAME = "model-parallel-rng"


class _State:
    def __init__(self):
        self.reset()

    def reset(self):
        self.configured = False
        self.mpu = None
        self.mp_rank = 0
        self.mp_size = 1
        self.mp_group = None
        self.partition_activations = False
        self.cpu_checkpoint = False
        self.contiguous = False
        self.num_layers = None
        self.synchronize = False
        self.profile = False
        self.arenas: List[Optional[torch.Tensor]] = []
        self.arena_cursor = 0
        self.slot = 0
        self.timers = None


_S = _State()
transport_stream = None


# ------------------------------------------------------------------------------ RNG tracking
class CudaRNGStatesTracker:
    """Named device RNG states so dropout inside tensor-parallel regions differs per TP rank while
    everything else stays identical (Megatron semantics)."""

    def __init__(self):
        self.states_ = {}
        self.seeds_ = set()

    def reset(self):
        self.states_ = {}
        self.seeds_ = set()

    def get_states(self):
        return dict(self.states_)

    def set_states(self, states):
        self.states_ = states

    def add(self, name, seed):
        if seed in self.seeds_:
            raise Exception(f"seed {seed} already exists")
        self.seeds_.add(seed)
        if name in self.states_:
            raise Exception(f"cuda rng state {name} already exists")
        acc = get_accelerator()
        orig = acc.get_rng_state()
        acc.manual_seed(seed)
        self.states_[name] = acc.get_rng_state()
        acc.set_rng_state(orig)

    @contextlib.contextmanager
    def fork(self, name=_MODEL_PARALLEL_RNG_TRACKER_NAME):
        if name not in self.states_:
            raise Exception(f"cuda rng state {name} is not added")
        acc = get_accelerator()
        orig = acc.get_rng_state()
        acc.set_rng_state(self.states_[name])
        try:
            yield
        finally:
            self.states_[name] = acc.get_rng_state()
            acc.set_rng_state(orig)


_CUDA_RNG_STATE_TRACKER = CudaRNGStatesTracker()


def get_cuda_rng_tracker():
    return _CUDA_RNG_STATE_TRACKER


def model_parallel_cuda_manual_seed(seed: int):
    """Default state seeded identically on all TP ranks; the tracked state offset by the TP rank."""
    tp_rank = _S.mp_rank
    if _S.mpu is not None:
        tp_rank = _mpu_rank(_S.mpu)
    offset = seed + 2718
    _CUDA_RNG_STATE_TRACKER.reset()
    get_accelerator().manual_seed(seed)
    _CUDA_RNG_STATE_TRACKER.add(_MODEL_PARALLEL_RNG_TRACKER_NAME, offset + tp_rank)


model_parallel_reconfigure_tp_seed = model_parallel_cuda_manual_seed


def _mpu_rank(mpu):
    for n in ("get_tensor_model_parallel_rank", "get_model_parallel_rank"):
        if hasattr(mpu, n):
            return getattr(mpu, n)()
    return 0


def _mpu_size(mpu):
    for n in ("get_tensor_model_parallel_world_size", "get_model_parallel_world_size"):
        if hasattr(mpu, n):
            return getattr(mpu, n)()
    return 1


def _mpu_group(mpu):
    for n in ("get_tensor_model_parallel_group", "get_model_parallel_group"):
        if hasattr(mpu, n):
            return getattr(mpu, n)()
    return None


# ---------------------------------------------------------------------- saved-input handling
class _Saved:
    """One checkpointed input: possibly partitioned across TP ranks and/or parked on the host."""
    __slots__ = ("data", "shape", "dtype", "device", "numel", "partitioned", "on_cpu", "requires_grad", "is_tensor")

    def __init__(self, value):
        self.is_tensor = torch.is_tensor(value)
        if not self.is_tensor:
            self.data = value
            return
        self.shape, self.dtype, self.device = value.shape, value.dtype, value.device
        self.numel = value.numel()
        self.requires_grad = value.requires_grad
        t = value.detach()
        self.partitioned = bool(_S.partition_activations and _S.mp_size > 1 and t.is_floating_point()
                                and self.numel >= _S.mp_size)
        if self.partitioned:
            part = (self.numel + _S.mp_size - 1) // _S.mp_size
            flat = t.contiguous().view(-1)
            lo = min(_S.mp_rank * part, self.numel)
            hi = min(lo + part, self.numel)
            piece = flat.new_zeros(part)
            piece[: hi - lo] = flat[lo:hi]
            t = piece
        self.on_cpu = bool(_S.cpu_checkpoint and t.is_floating_point() and t.device.type != "cpu")
        if self.on_cpu:
            host = torch.empty(t.shape, dtype=t.dtype, device="cpu", pin_memory=get_accelerator().is_available())
            host.copy_(t, non_blocking=True)
            t = host
        elif _S.contiguous and self.partitioned:
            t = _arena_store(t)
        elif self.partitioned:
            t = t.clone()
        self.data = t

    def restore(self):
        if not self.is_tensor:
            return self.data
        t = self.data
        if self.on_cpu:
            t = t.to(self.device, non_blocking=True)
        if self.partitioned:
            full = torch.empty(t.numel() * _S.mp_size, dtype=t.dtype, device=t.device)
            dist.all_gather_into_tensor(full, t.contiguous(), group=_S.mp_group)
            t = full[: self.numel].view(self.shape)
        t = t.detach()
        t.requires_grad_(self.requires_grad)
        return t


def _arena_store(piece: torch.Tensor) -> torch.Tensor:
    slot = _S.slot % max(1, _S.num_layers or 1)
    while len(_S.arenas) <= slot:
        _S.arenas.append(None)
    need = _S.arena_cursor + piece.numel()
    arena = _S.arenas[slot]
    if arena is None or arena.numel() < need or arena.dtype != piece.dtype or arena.device != piece.device:
        grown = torch.empty(max(need, 2 * (arena.numel() if arena is not None and arena.dtype == piece.dtype else 0)),
                            dtype=piece.dtype, device=piece.device)
        if arena is not None and arena.dtype == piece.dtype and _S.arena_cursor:
            grown[: _S.arena_cursor].copy_(arena[: _S.arena_cursor])
        _S.arenas[slot] = arena = grown
    view = arena[_S.arena_cursor: need]
    view.copy_(piece)
    _S.arena_cursor = need
    return view


def _capture_rng():
    acc = get_accelerator()
    return (torch.get_rng_state(), acc.get_rng_state() if acc.is_available() else None,
            _CUDA_RNG_STATE_TRACKER.get_states())


def _replay_rng(state):
    cpu, dev, tracked = state
    torch.set_rng_state(cpu)
    if dev is not None:
        get_accelerator().set_rng_state(dev)
    _CUDA_RNG_STATE_TRACKER.set_states(tracked)


class CheckpointFunction(torch.autograd.Function):
    """Reentrant checkpoint: forward runs under ``no_grad`` keeping only (possibly partitioned /
    offloaded) inputs; backward replays RNG, recomputes with grad and back-propagates."""

    @staticmethod
    def forward(ctx, run_function, n_args, *args):
        ctx.run_function = run_function
        ctx.rng = _capture_rng()
        if _S.profile and _S.timers is not None:
            _S.timers("forward").start()
        _S.arena_cursor = 0
        ctx.saved = [_Saved(a) for a in args]
        _S.slot += 1
        with torch.no_grad():
            outputs = run_function(*args)
        if _S.profile and _S.timers is not None:
            _S.timers("forward").stop()
        if _S.synchronize:
            get_accelerator().synchronize()
        ctx.single = torch.is_tensor(outputs)
        if ctx.single:
            return outputs
        outs = tuple(outputs)
        ctx.non_tensor_outputs = [(i, o) for i, o in enumerate(outs) if not torch.is_tensor(o)]
        for o in outs:
            if torch.is_tensor(o) and not o.is_floating_point():
                ctx.mark_non_differentiable(o)
        return outs

    @staticmethod
    def backward(ctx, *grads):
        if not torch.autograd._is_checkpoint_valid():
            raise RuntimeError("Checkpointing is not compatible with .grad(), please use .backward() if possible")
        inputs = [s.restore() for s in ctx.saved]
        now = _capture_rng()
        _replay_rng(ctx.rng)
        with torch.enable_grad():
            outputs = ctx.run_function(*inputs)
        _replay_rng(now)
        if torch.is_tensor(outputs):
            outputs = (outputs,)
        out_t, grad_t = [], []
        for o, g in zip(outputs, grads):
            if torch.is_tensor(o) and o.requires_grad and g is not None:
                out_t.append(o)
                grad_t.append(g)
        if _S.synchronize:
            get_accelerator().synchronize()
        if out_t:
            torch.autograd.backward(out_t, grad_t)
        ctx.saved = None
        in_grads = tuple(i.grad if torch.is_tensor(i) and i.requires_grad else None for i in inputs)
        return (None, None) + in_grads


def checkpoint(function, *args):
    """Checkpoint ``function(*args)`` (non-tensor positional arguments are passed through)."""
    out = CheckpointFunction.apply(function, len(args), *args)
    return out


def non_reentrant_checkpoint(function, *args):
    """Saved-tensor-hook checkpoint: nothing inside ``function`` is kept; on first unpack in backward
    the whole segment is recomputed once with the forward RNG state and the saved tensors are served
    from that recomputation."""
    rng = _capture_rng()
    frame = {"recomputed": None, "order": 0}
    detached = [a.detach().requires_grad_(a.requires_grad) if torch.is_tensor(a) else a for a in args]

    class _Holder:
        pass

    def pack(t):
        h = _Holder()
        h.idx = frame["order"]
        frame["order"] += 1
        return h

    def unpack(h):
        if frame["recomputed"] is None:
            store = []

            def inner_pack(t):
                store.append(t.detach())
                return len(store) - 1

            def inner_unpack(i):  # pragma: no cover - recomputation graph is never back-propagated
                return store[i]

            now = _capture_rng()
            _replay_rng(rng)
            with torch.enable_grad(), torch.autograd.graph.saved_tensors_hooks(inner_pack, inner_unpack):
                function(*detached)
            _replay_rng(now)
            frame["recomputed"] = store
        return frame["recomputed"][h.idx]

    with torch.autograd.graph.saved_tensors_hooks(pack, unpack):
        return function(*args)


# ------------------------------------------------------------------------------ configuration
def partition_activations_in_checkpoint(partition_activation):
    _S.partition_activations = partition_activation


def set_num_layers(nlayers):
    _S.num_layers = nlayers


def reset():
    """Drop the contiguous arenas (call between iterations when shapes change)."""
    _S.arenas = []
    _S.arena_cursor = 0
    _S.slot = 0


def _configure_using_config_file(config, mpu=None):
    from ..config import DeepSpeedConfig
    c = DeepSpeedConfig(config, mpu=mpu).activation_checkpointing_config
    _S.partition_activations = c.partition_activations
    _S.contiguous = c.contiguous_memory_optimization
    _S.num_layers = c.number_checkpoints
    _S.cpu_checkpoint = c.cpu_checkpointing
    _S.synchronize = c.synchronize_checkpoint_boundary
    _S.profile = c.profile


def configure(mpu_, deepspeed_config=None, partition_activations=None, contiguous_checkpointing=None,
              num_checkpoints=None, checkpoint_in_cpu=None, synchronize=None, profile=None):
    """Configure from a config (path/dict) and/or explicit overrides."""
    if deepspeed_config is not None:
        _configure_using_config_file(deepspeed_config, mpu=mpu_)
    _S.mpu = mpu_
    if mpu_ is not None:
        _S.mp_rank, _S.mp_size, _S.mp_group = _mpu_rank(mpu_), _mpu_size(mpu_), _mpu_group(mpu_)
    for name, val in (("partition_activations", partition_activations), ("contiguous", contiguous_checkpointing),
                      ("num_layers", num_checkpoints), ("cpu_checkpoint", checkpoint_in_cpu),
                      ("synchronize", synchronize), ("profile", profile)):
        if val is not None:
            setattr(_S, name, val)
    if _S.contiguous:
        assert _S.partition_activations, "Contiguous Checkpointing is only available with partitioned activations. Set partitioned activations to true in deepspeed config"
        assert _S.num_layers is not None, "Must specify the number of layers with contiguous memory checkpointing"
    if _S.profile:
        _S.timers = SynchronizedWallClockTimer()
    _S.configured = True


def is_configured():
    return _S.configured


# ============================================================ reference-named helpers (``checkpointing.py:196-486``)
# The checkpoint function above keeps its inputs as `_Saved` records; these are the same operations as free functions,
# for code that drives partitioned / host-resident activations itself.
def detach_variable(inputs, device=None):
    """Detached copies (same ``requires_grad``) of a tensor or tuple of tensors, optionally moved to ``device``."""
    if isinstance(inputs, tuple):
        out = []
        for inp in inputs:
            if not torch.is_tensor(inp):
                out.append(inp)
                continue
            x = inp.to(device=device) if device is not None else inp
            x = x.detach()
            x.requires_grad = inp.requires_grad
            out.append(x)
        return tuple(out)
    raise RuntimeError("Only tuple of tensors is supported. Got Unsupported input type: ", type(inputs).__name__)


def get_partition_size(item) -> int:
    """Elements of ``item`` each model-parallel rank keeps (ceil division: the last rank's piece is zero-padded)."""
    size = item.numel()
    return (size + _S.mp_size - 1) // _S.mp_size if _S.mp_size > 1 else size


def get_partition_start(item) -> int:
    return min(get_partition_size(item) * _S.mp_rank, item.numel())


def is_activation_to_checkpoint(item) -> bool:
    """Tensors worth partitioning / offloading: floating point and at least one element per rank."""
    return torch.is_tensor(item) and item.is_floating_point() and item.numel() >= max(1, _S.mp_size)


def extract_tensors(all_objects):
    """Split a tuple into ``(tensors, non_tensors, tensor_flags)`` so the tensors can go through autograd's
    ``save_for_backward`` while everything else is kept aside."""
    tensors = [o for o in all_objects if torch.is_tensor(o)]
    non_tensors = [o for o in all_objects if not torch.is_tensor(o)]
    flags = [torch.is_tensor(o) for o in all_objects]
    if isinstance(all_objects, tuple):
        return tuple(tensors), tuple(non_tensors), tuple(flags)
    return tensors, non_tensors, flags


def merge_tensors(tensor_objects, non_tensor_objects, tensor_flags):
    """Inverse of :func:`extract_tensors`."""
    merged, ti, ni = [], 0, 0
    real_flags = tensor_flags[:-1] if len(tensor_flags) > len(tensor_objects) + len(non_tensor_objects) else tensor_flags
    for is_tensor in real_flags:
        if is_tensor:
            merged.append(tensor_objects[ti]); ti += 1
        else:
            merged.append(non_tensor_objects[ni]); ni += 1
    return tuple(merged)


def partition_activations(args, cpu_checkpoint=False, contiguous_checkpoint=False):
    """``[piece_0, size_0, piece_1, size_1, ...]``: this rank's slice of every checkpointable tensor followed by its
    original shape (as a tensor), the layout :func:`gather_partitioned_activations` consumes."""
    out = []
    for item in args:
        if not is_activation_to_checkpoint(item):
            out.append(item)
            continue
        part = get_partition_size(item)
        lo = get_partition_start(item)
        flat = item.detach().contiguous().view(-1)
        piece = flat.new_zeros(part)
        hi = min(lo + part, flat.numel())
        piece[: hi - lo] = flat[lo:hi]
        if cpu_checkpoint:
            piece = piece.to("cpu")
        out.append(piece)
        out.append(torch.tensor(list(item.shape), dtype=torch.long))
    return out


def gather_partitioned_activations(tensors, device=None):
    """Rebuild full activations from ``[piece, size, piece, size, ...]`` (all-gather over the model-parallel group)."""
    assert len(tensors) % 2 == 0, f"Expected even count of tensors, instead got {len(tensors)}"
    out = []
    for i in range(len(tensors) // 2):
        piece, size = tensors[2 * i], tensors[2 * i + 1]
        if not is_activation_to_checkpoint(piece) and piece.numel() * max(1, _S.mp_size) < max(1, _S.mp_size):
            out.append(piece)
            continue
        shape = [int(s) for s in size.tolist()]
        numel = 1
        for s in shape:
            numel *= s
        if device is not None:
            piece = piece.to(device)
        if _S.mp_size > 1:
            full = torch.empty(piece.numel() * _S.mp_size, dtype=piece.dtype, device=piece.device)
            dist.all_gather_into_tensor(full, piece.contiguous(), group=_S.mp_group)
        else:
            full = piece
        out.append(full[:numel].view(shape))
    return out


def get_partitioned_activations_for_backward(args, inputs, contiguous_checkpoint=False):
    """What a checkpoint with partitioned activations saves: for every checkpointable input its piece and shape, other
    inputs unchanged followed by a ``None`` placeholder (keeps the pairs aligned)."""
    new_args = []
    for i, (arg, inp) in enumerate(zip(args, inputs)):
        if not is_activation_to_checkpoint(arg):
            new_args += [arg, None]
            continue
        piece = partition_activations([arg])[0]
        new_args += [piece, torch.tensor(list(arg.shape), dtype=torch.long)]
    return new_args


def get_cpu_activations_for_backward(args, inputs):
    """Host copies of the checkpointable inputs (``cpu_checkpointing`` without partitioning)."""
    new_args = []
    for arg, inp in zip(args, inputs):
        if not is_activation_to_checkpoint(arg):
            new_args.append(arg)
        else:
            new_args.append(inp.detach().to("cpu"))
    return new_args
