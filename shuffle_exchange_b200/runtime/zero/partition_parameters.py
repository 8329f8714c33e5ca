"""``zero.Init`` / ``GatheredParameters`` / external-parameter registry.

Reference: deepspeed/runtime/zero/partition_parameters.py — ``Init`` patches ``nn.Module.__init__`` so
every parameter is partitioned the moment its module is constructed (:320-660, 879-1103);
``GatheredParameters`` temporarily reassembles parameters, broadcasting edits made by
``modifier_rank`` (:2193-2329); ``register_external_parameter`` (:128-212) tells the coordinator that a
module uses a parameter it does not own.

Here partition-at-construction keeps the same user contract with a simpler mechanism: while the
context is active every ``nn.Module.__init__`` is wrapped so that, when the *outermost* constructor of
a sub-tree returns, each newly created parameter is replaced by its 1/world shard (``param.ds_tensor``)
and an empty placeholder (``ds_status == NOT_AVAILABLE``).  The ZeRO-3 optimizer later forms its
unit-flat shards from these (``stage3._materialize_pre_partitioned``), one unit at a time, so peak
memory is max(largest module, largest unit) — never the whole model.
"""
from __future__ import annotations

import functools
import os
import itertools
from enum import Enum
from typing import Iterable, List, Optional

import torch
import torch.nn as nn

from ...comm import comm as dist
from ...device import get_accelerator
from ...utils.logging import logger

_param_id = itertools.count()
zero_init_context = 0
_active_init: Optional["Init"] = None
_suspended: Optional["Init"] = None


class ZeroParamType(Enum):
    NORMAL = 1
    PARTITIONED = 2
    REMOTE = 3


class ZeroParamStatus(Enum):
    AVAILABLE = 1
    NOT_AVAILABLE = 2
    INFLIGHT = 3


def is_zero_param(p) -> bool:
    return torch.is_tensor(p) and hasattr(p, "ds_id")


def _shard_of(t: torch.Tensor, world: int, rank: int) -> torch.Tensor:
    n = t.numel()
    per = (n + world - 1) // world
    flat = t.detach().reshape(-1)
    out = flat.new_zeros(per)
    lo, hi = min(rank * per, n), min((rank + 1) * per, n)
    out[: hi - lo] = flat[lo:hi]
    return out


class Init:
    """Context manager: modules built inside have their parameters sharded at construction.

    Two granularities (``Init.granularity`` / env ``SXB_ZERO_INIT_GRANULARITY``):

    * ``"model"`` (default): the model is sharded when its OUTERMOST constructor returns.  Any initialisation code
      inside constructors (loops over ``self.parameters()``, ``self.apply(...)``, weight tying) sees ordinary tensors;
      the price is that one rank holds the whole model (on the construction device, normally the host) for a moment.
    * ``"module"``: every sub-module's own parameters are sharded as soon as ITS constructor returns, as the reference
      does (``partition_parameters.py:1066-1100``), so no rank ever holds more than one sub-module whole — this is what
      builds models larger than one device / one host.  Code that runs later in a parent constructor then sees empty
      tensors; the one common pattern, ``module.apply(init_fn)`` (Hugging Face ``post_init``), is kept working by
      gathering each visited module's own parameters around ``init_fn`` and re-sharding rank 0's result
      (``override_module_apply`` of the reference); anything else needs ``GatheredParameters``.
    """
    granularity = os.environ.get("SXB_ZERO_INIT_GRANULARITY", "model")

    def __init__(self, module=None, data_parallel_group=None, mem_efficient_linear=True, remote_device=None,
                 pin_memory=False, config_dict_or_path=None, config=None, enabled=True, dtype=None, mpu=None,
                 zero_param_parallel_group=None, zero_quantized_weights=False,
                 zero_quantized_nontrainable_weights=False, sequence_data_parallel_group=None, param_swapper=None):
        self.enabled = enabled
        self.dtype = dtype
        self.remote_device = remote_device
        self.pin_memory = pin_memory
        self.group = data_parallel_group or sequence_data_parallel_group
        self.zero_param_process_group = zero_param_parallel_group
        self.depth = 0
        self._orig_init = None
        self._seen = set()
        cfg = config_dict_or_path if config_dict_or_path is not None else config
        if cfg is not None and dtype is None:
            try:
                from ..config import DeepSpeedConfig
                c = DeepSpeedConfig(cfg, mpu)
                self.dtype = torch.float16 if c.fp16_enabled else (torch.bfloat16 if c.bfloat16_enabled else None)
            except Exception:
                pass
        if module is not None and enabled:
            self._convert_module(module)

    # -------------------------------------------------------------------------------- context
    # ---- partitioning geometry (reference ``partition_parameters.py:1100-1140``)
    def get_dp_process_group(self):
        """Group the parameters are sharded over (the hpZ secondary group when one was given)."""
        return self.zero_param_process_group or self.group or (dist.get_world_group() if dist.is_initialized() else None)

    def get_partition_dp_group(self, param=None):
        return getattr(param, "ds_process_group", None) or self.get_dp_process_group()

    def get_partition_rank(self) -> int:
        g = self.get_dp_process_group()
        return dist.get_rank(group=g) if dist.is_initialized() else 0

    @property
    def num_partitions(self) -> int:
        g = self.get_dp_process_group()
        return dist.get_world_size(group=g) if dist.is_initialized() else 1

    def __enter__(self):
        global _active_init, zero_init_context
        if not self.enabled:
            return self
        if not dist.is_initialized():
            dist.init_distributed()
        zero_init_context += 1
        _active_init = self
        self._patch()
        return self

    def __exit__(self, exc_type, exc, tb):
        global _active_init, zero_init_context
        if not self.enabled:
            return False
        zero_init_context -= 1
        self._unpatch()
        _active_init = None
        return False

    def _patch(self):
        outer = self
        self._orig_init = nn.Module.__init__
        self._orig_subclass_inits = {}

        def wrap(cls):
            if "__init__" not in cls.__dict__ or cls in self._orig_subclass_inits:
                return
            orig = cls.__dict__["__init__"]
            self._orig_subclass_inits[cls] = orig

            @functools.wraps(orig)
            def wrapped(module, *a, **k):
                outer.depth += 1
                try:
                    orig(module, *a, **k)
                finally:
                    outer.depth -= 1
                if outer.granularity == "module":
                    outer._convert_module(module, recurse=False)       # children sharded theirs when they returned
                elif outer.depth == 0:
                    outer._convert_module(module)

            cls.__init__ = wrapped

        def all_subclasses(c):
            for s in c.__subclasses__():
                yield s
                yield from all_subclasses(s)

        for c in list(all_subclasses(nn.Module)):
            wrap(c)
        self._orig_init_subclass = nn.Module.__dict__.get("__init_subclass__")

        def init_subclass(cls, **kw):
            wrap(cls)

        nn.Module.__init_subclass__ = classmethod(init_subclass)
        self._orig_apply = None
        if self.granularity == "module":
            self._orig_apply = nn.Module.apply

            def sharded_apply(module, fn):
                """``module.apply(fn)`` over sharded parameters: children first (as torch does), then this module with
                its OWN parameters made whole; rank 0's result is what every rank keeps."""
                for child in module.children():
                    child.apply(fn)
                own = [p for p in module.parameters(recurse=False)
                       if is_zero_param(p) and getattr(p, "ds_tensor", None) is not None and p.numel() == 0]
                if own:
                    with GatheredParameters(own, modifier_rank=0):
                        fn(module)
                else:
                    fn(module)
                return module

            nn.Module.apply = sharded_apply
            self._patch_init_functions()

    _INIT_FUNCTIONS = ("uniform_", "normal_", "trunc_normal_", "constant_", "ones_", "zeros_", "eye_", "dirac_",
                       "xavier_uniform_", "xavier_normal_", "kaiming_uniform_", "kaiming_normal_", "orthogonal_",
                       "sparse_")

    def _patch_init_functions(self):
        """``torch.nn.init.<fn>(param, ...)`` on a sharded parameter: make it whole, initialise, keep rank 0's values.
        Libraries that captured the functions in a table before this context was entered (``transformers >= 5``:
        ``initialization.TORCH_INIT_FUNCTIONS``, which its ``_init_weights`` goes through) get the table patched too.
        In-place methods on ``param.data`` cannot be seen from here — such code needs ``GatheredParameters``."""
        def wrap(orig):
            @functools.wraps(orig)
            def init_fn(tensor, *a, **k):
                if is_zero_param(tensor) and getattr(tensor, "ds_tensor", None) is not None and tensor.numel() == 0:
                    with GatheredParameters([tensor], modifier_rank=0):
                        orig(tensor, *a, **k)
                    return tensor
                return orig(tensor, *a, **k)
            return init_fn

        self._init_patches = []
        tables = [vars(torch.nn.init)]
        try:
            import sys
            hf = sys.modules.get("transformers.initialization")
            if hf is not None and isinstance(getattr(hf, "TORCH_INIT_FUNCTIONS", None), dict):
                tables.append(hf.TORCH_INIT_FUNCTIONS)
        except Exception:
            pass
        for table in tables:
            for name in self._INIT_FUNCTIONS:
                orig = table.get(name)
                if callable(orig):
                    self._init_patches.append((table, name, orig))
                    if table is vars(torch.nn.init):
                        setattr(torch.nn.init, name, wrap(orig))
                    else:
                        table[name] = wrap(orig)

    def _unpatch_init_functions(self):
        for table, name, orig in getattr(self, "_init_patches", []):
            if table is vars(torch.nn.init):
                setattr(torch.nn.init, name, orig)
            else:
                table[name] = orig
        self._init_patches = []

    def _unpatch(self):
        self._unpatch_init_functions()
        for cls, orig in self._orig_subclass_inits.items():
            cls.__init__ = orig
        self._orig_subclass_inits = {}
        if getattr(self, "_orig_apply", None) is not None:
            nn.Module.apply = self._orig_apply
            self._orig_apply = None
        if self._orig_init_subclass is None:
            try:
                del nn.Module.__init_subclass__
            except AttributeError:
                pass
        else:
            nn.Module.__init_subclass__ = self._orig_init_subclass

    # ------------------------------------------------------------------------------ conversion
    def _world(self):
        if not dist.is_initialized():
            return 1, 0
        return dist.get_world_size(self.group), dist.get_rank(self.group)

    def _convert_module(self, module: nn.Module, recurse: bool = True):
        world, rank = self._world()
        dev = torch.device(self.remote_device) if self.remote_device in ("cpu",) else \
            torch.device(get_accelerator().current_device_name())
        for name, p in module.named_parameters(recurse=recurse):
            if is_zero_param(p) or id(p) in self._seen:
                continue
            self._seen.add(id(p))
            self._convert_param(p, world, rank, dev)

    def _convert_param(self, p, world, rank, dev):
        data = p.data
        if self.dtype is not None and data.is_floating_point():
            data = data.to(self.dtype)
        if world > 1 and dist.is_initialized():
            # every rank must start from rank 0's initial values (reference: broadcast at :1109-1116)
            full = data.to(dev)
            dist.broadcast(full, dist.get_global_rank(self.group, 0) if self.group is not None else 0, group=self.group)
            data = full
        p.ds_id = next(_param_id)
        p.ds_shape = tuple(p.shape)
        p.ds_numel = p.numel()
        p.ds_tensor = _shard_of(data.to(dev), world, rank)
        if self.pin_memory and p.ds_tensor.device.type == "cpu":
            p.ds_tensor = p.ds_tensor.pin_memory()
        p.ds_status = ZeroParamStatus.NOT_AVAILABLE
        p.ds_persist = False
        p.ds_process_group = self.group
        p.data = torch.empty(0, dtype=data.dtype, device=dev)
        p.partition_numel = lambda p=p: p.ds_tensor.numel() if p.ds_tensor is not None else 0


def shutdown_init_context():
    """Temporarily lift an active ``zero.Init`` (``initialize()`` runs its own constructors)."""
    global _suspended, _active_init
    if _active_init is not None:
        _suspended = _active_init
        _active_init._unpatch()
        _active_init = None


def restore_init_context():
    global _suspended, _active_init
    if _suspended is not None:
        _suspended._patch()
        _active_init = _suspended
        _suspended = None


# ------------------------------------------------------------------------------ external params
def register_external_parameter(module: nn.Module, parameter: nn.Parameter) -> None:
    """Declare that ``module.forward`` uses ``parameter`` owned by another module.  With unit-level
    gathering this only matters when the owner lives in a different unit: the owner unit is then
    fetched together with ``module``'s unit."""
    if not isinstance(parameter, nn.Parameter):
        raise RuntimeError("Parameter is not a torch.nn.Parameter")
    if not hasattr(module, "_external_params"):
        module._external_params = {}
    module._external_params[id(parameter)] = parameter


def unregister_external_parameter(module: nn.Module, parameter: nn.Parameter) -> None:
    if not hasattr(module, "_external_params") or id(parameter) not in module._external_params:
        raise RuntimeError("Parameter is not a registered external parameter of module.")
    del module._external_params[id(parameter)]


class GatheredParameters:
    """Temporarily make ZeRO-3 parameters whole::

        with zero.GatheredParameters(layer.weight, modifier_rank=0):
            if dist.get_rank() == 0:
                layer.weight.zero_()

    On exit, if ``modifier_rank`` is given, its values are broadcast and written back to the shards
    (low-precision shard and fp32 master)."""

    def __init__(self, params, modifier_rank: Optional[int] = None, fwd_module: Optional[nn.Module] = None,
                 enabled: bool = True):
        self.enabled = enabled
        if not enabled:
            return
        if isinstance(params, nn.Module):
            params = list(params.parameters())
        elif torch.is_tensor(params):
            params = [params]
        self.params: List[nn.Parameter] = [p for p in params if p is not None]
        self.src_rank = modifier_rank
        self.units = []
        self.raw = []
        if fwd_module is not None:
            for p in self.params:
                register_external_parameter(fwd_module, p)

    def __enter__(self):
        if not self.enabled:
            return
        seen = set()
        for p in self.params:
            lk = getattr(p, "_sxb_link", None)
            if lk is not None and hasattr(lk.optimizer, "_fetch"):
                u = lk.group.unit
                if id(u) not in seen:
                    seen.add(id(u))
                    self.units.append((lk.optimizer, u, u.available))
                    lk.optimizer._fetch(u)
            elif is_zero_param(p) and getattr(p, "ds_tensor", None) is not None and p.numel() == 0:
                self.raw.append(p)
                self._gather_raw(p)

    def _gather_raw(self, p):
        grp = p.ds_process_group
        world = dist.get_world_size(grp) if dist.is_initialized() else 1
        shard = p.ds_tensor
        full = torch.empty(shard.numel() * world, dtype=shard.dtype, device=shard.device)
        if world > 1:
            dist.all_gather_into_tensor(full, shard, group=grp)
        else:
            full.copy_(shard)
        p.data = full[: p.ds_numel].view(p.ds_shape)
        p.ds_status = ZeroParamStatus.AVAILABLE

    def __exit__(self, *exc):
        if not self.enabled:
            return False
        if self.src_rank is not None and dist.is_initialized():
            for p in self.params:
                if p.numel() > 0:
                    grp = getattr(p, "ds_process_group", None)
                    lk = getattr(p, "_sxb_link", None)
                    if lk is not None:
                        grp = lk.group.pg
                    src = dist.get_global_rank(grp, self.src_rank) if grp is not None else self.src_rank
                    dist.broadcast(p.data, src, group=grp)
        if self.src_rank is not None:
            done = set()
            for p in self.params:
                lk = getattr(p, "_sxb_link", None)
                if lk is None or id(lk.group) in done:
                    continue
                done.add(id(lk.group))
                fg = lk.group
                if fg.flat_lp is not None:
                    sl = fg.flat_lp[fg.rank * fg.part: (fg.rank + 1) * fg.part]
                    if sl.data_ptr() != fg.lp_shard.data_ptr():
                        fg.lp_shard.copy_(sl)
                    fg.hp_param.data.copy_(fg.lp_shard.to(fg.hp_param.device))
        for opt, u, was_available in self.units:
            if not was_available:
                opt._release(u)
        for p in self.raw:
            if self.src_rank is not None:
                world = dist.get_world_size(p.ds_process_group) if dist.is_initialized() else 1
                rank = dist.get_rank(p.ds_process_group) if dist.is_initialized() else 0
                p.ds_tensor = _shard_of(p.data, world, rank)
            p.data = torch.empty(0, dtype=p.dtype, device=p.device)
            p.ds_status = ZeroParamStatus.NOT_AVAILABLE
        return False
