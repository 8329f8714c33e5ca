"""Per-module FLOPs / MACs / latency / parameter profile of one forward pass.

Reference: ``deepspeed/profiling/flops_profiler/profiler.py:30-1271`` — same public surface
(``FlopsProfiler.start_profile/stop_profile/reset_profile/end_profile``, ``get_total_*``, ``print_model_profile``,
``print_model_aggregated_profile``, ``get_model_profile``) and the same engine integration (profile at
``flops_profiler.profile_step``, reference engine.py:2105-2150).

Counting mechanism (re-designed): instead of monkey-patching ~60 ``torch.nn.functional`` / ``Tensor`` methods, one
``TorchDispatchMode`` sees every ATen op that actually runs (matmul family, convolution, scaled-dot-product attention
variants, elementwise ops) and the framework's own native kernels — invisible to the dispatcher — report through
``ops.functional.flop_sink``.  Work is attributed to the innermost module on a hook-maintained stack and rolled up to
the ancestors at print time.
"""
from __future__ import annotations

import math
import sys
import time
from collections import OrderedDict
from functools import partial
from typing import List, Optional

import torch
import torch.nn as nn
from torch.utils._python_dispatch import TorchDispatchMode

DEFAULT_PRECISION = 2
aten = torch.ops.aten


def _prod(dims):
    p = 1
    for d in dims:
        p *= int(d)
    return p


def _mm_flops(args):
    a, b = args[0], args[1]
    return 2 * _prod(a.shape) * b.shape[-1]


def _addmm_flops(args):
    a, b = args[1], args[2]
    return 2 * _prod(a.shape) * b.shape[-1] + _prod(args[0].shape)


def _bmm_flops(args):
    a, b = args[0], args[1]
    return 2 * _prod(a.shape) * b.shape[-1]


def _conv_flops(args):
    x, w = args[0], args[1]
    transposed = bool(args[6]) if len(args) > 6 else False
    out_ch_per_group_k = _prod(w.shape[1:])           # (C_in/groups) * kernel volume  (C_out/groups for transposed)
    batch = x.shape[0]
    if transposed:
        return 2 * batch * _prod(x.shape[1:]) * out_ch_per_group_k
    stride = args[3] if len(args) > 3 else [1] * (x.dim() - 2)
    padding = args[4] if len(args) > 4 else [0] * (x.dim() - 2)
    dilation = args[5] if len(args) > 5 else [1] * (x.dim() - 2)
    out_sp = 1
    for i, n in enumerate(x.shape[2:]):
        k = w.shape[2 + i]
        out_sp *= (n + 2 * padding[i] - dilation[i] * (k - 1) - 1) // stride[i] + 1
    return 2 * batch * w.shape[0] * out_sp * out_ch_per_group_k


def _sdpa_flops(args):
    q, k = args[0], args[1]
    b_h = _prod(q.shape[:-2])
    return 4 * b_h * q.shape[-2] * k.shape[-2] * q.shape[-1]


_MATMUL_LIKE = {
    aten.mm.default: _mm_flops, aten.addmm.default: _addmm_flops, aten.bmm.default: _bmm_flops,
    aten.baddbmm.default: lambda a: _bmm_flops(a[1:]) + _prod(a[0].shape),
    aten.convolution.default: _conv_flops,
}
for _name in ("_scaled_dot_product_flash_attention", "_scaled_dot_product_efficient_attention",
              "_scaled_dot_product_cudnn_attention", "_scaled_dot_product_flash_attention_for_cpu"):
    if hasattr(aten, _name):
        _MATMUL_LIKE[getattr(aten, _name).default] = _sdpa_flops

_ELEMENTWISE_1 = ("add", "sub", "mul", "div", "relu", "gelu", "silu", "tanh", "sigmoid", "exp", "neg", "rsqrt", "sqrt",
                  "pow", "leaky_relu", "elu", "hardtanh", "where", "maximum", "minimum", "abs", "clamp")
_NORM_LIKE = {"native_layer_norm": 5, "native_batch_norm": 4, "native_group_norm": 5, "_softmax": 3, "_log_softmax": 3}


class _Counter(TorchDispatchMode):
    def __init__(self, profiler):
        super().__init__()
        self.prof = profiler

    def __torch_dispatch__(self, func, types, args=(), kwargs=None):
        out = func(*args, **(kwargs or {}))
        fn = _MATMUL_LIKE.get(func)
        if fn is not None:
            try:
                fl = fn(args)
                self.prof._add(fl, fl / 2)
            except Exception:
                pass
            return out
        name = func._schema.name.split("::")[-1]
        base = name.rstrip("_")
        if base in _NORM_LIKE:
            ref = out[0] if isinstance(out, (tuple, list)) else out
            if torch.is_tensor(ref):
                self.prof._add(_NORM_LIKE[base] * ref.numel(), 0)
        elif base in _ELEMENTWISE_1:
            if torch.is_tensor(out):
                self.prof._add(out.numel(), 0)
        return out


class FlopsProfiler:
    def __init__(self, model, ds_engine=None, recompute_fwd_factor: float = 0.0):
        self.model = model
        self.ds_engine = ds_engine
        self.recompute_fwd_factor = recompute_fwd_factor
        self.started = False
        self.func_patched = False
        self._stack: List[nn.Module] = []
        self._hooks = []
        self._mode: Optional[_Counter] = None

    # ------------------------------------------------------------------------------ accounting
    def _add(self, flops, macs):
        if self._stack:
            m = self._stack[-1]
            m.__flops__ += flops
            m.__macs__ += macs
        else:
            self.model.__flops__ += flops
            self.model.__macs__ += macs

    def _native_sink(self, kind, flops):
        self._add(flops, flops / 2)

    def _sync(self):
        if torch.cuda.is_available():
            torch.cuda.synchronize()

    # -------------------------------------------------------------------------------- lifecycle
    def start_profile(self, ignore_list=None):
        self.reset_profile()
        ignore = tuple(ignore_list or ())

        def pre(mod, inp):
            self._stack.append(mod)
            self._sync()
            mod.__start_time__ = time.time()

        def post(mod, inp, out):
            self._sync()
            mod.__duration__ += time.time() - mod.__start_time__
            if self._stack and self._stack[-1] is mod:
                self._stack.pop()

        for mod in self.model.modules():
            if ignore and isinstance(mod, ignore):
                continue
            self._hooks.append(mod.register_forward_pre_hook(pre))
            self._hooks.append(mod.register_forward_hook(post))
        from ...ops import functional as F
        F.flop_sink = self._native_sink
        self._mode = _Counter(self)
        self._mode.__enter__()
        self.started = self.func_patched = True

    def stop_profile(self):
        if self.started and self._mode is not None:
            self._mode.__exit__(None, None, None)
            self._mode = None
            from ...ops import functional as F
            F.flop_sink = None
        for h in self._hooks:
            h.remove()
        self._hooks = []
        self._stack = []
        self.func_patched = False

    def reset_profile(self):
        def count(p):
            return p.ds_numel if hasattr(p, "ds_numel") else p.numel()

        for mod in self.model.modules():
            own = list(mod.parameters(recurse=False))
            mod.__params__ = sum(count(p) for p in own if not (hasattr(p, "allreduce") and not p.allreduce))
            mod.__expert_params__ = sum(count(p) for p in own if hasattr(p, "allreduce") and not p.allreduce)
            mod.__flops__ = 0
            mod.__macs__ = 0
            mod.__duration__ = 0.0

    def end_profile(self):
        if not self.started:
            return
        self.stop_profile()
        self.started = False
        for mod in self.model.modules():
            for attr in ("__flops__", "__macs__", "__params__", "__expert_params__", "__start_time__", "__duration__"):
                if hasattr(mod, attr):
                    delattr(mod, attr)

    # ------------------------------------------------------------------------------------- totals
    @staticmethod
    def _tree_sum(mod, attr):
        return getattr(mod, attr, 0) + sum(FlopsProfiler._tree_sum(c, attr) for c in mod.children())

    def get_total_flops(self, as_string=False):
        t = self._tree_sum(self.model, "__flops__")
        return number_to_string(t) if as_string else t

    def get_total_macs(self, as_string=False):
        t = self._tree_sum(self.model, "__macs__")
        return macs_to_string(t) if as_string else t

    def get_total_duration(self, as_string=False):
        t = getattr(self.model, "__duration__", 0.0)
        if t == 0.0:      # the root was called through a wrapper: fall back to the children
            t = sum(getattr(c, "__duration__", 0.0) for c in self.model.children())
        return duration_to_string(t) if as_string else t

    def get_total_params(self, as_string=False):
        total = self._tree_sum(self.model, "__params__")
        ep = self._tree_sum(self.model, "__expert_params__")
        if ep:
            from ...utils import groups
            tp = getattr(groups, "expert_tensor_parallel_world_size", 1) or 1
            total += ep * (tp if self.is_expert_tensor_parallelism_enabled() else 1)
        return params_to_string(total) if as_string else total

    def is_expert_tensor_parallelism_enabled(self):
        for mod in self.model.modules():
            if getattr(mod, "enable_expert_tensor_parallelism", False):
                return True
        return False

    # ------------------------------------------------------------------------------------ report
    def print_model_profile(self, profile_step=1, module_depth=-1, top_modules=1, detailed=True, output_file=None):
        if not self.started:
            return
        out = open(output_file, "w") if output_file else sys.stdout
        p = partial(print, file=out)
        total_flops, total_macs = self.get_total_flops(), self.get_total_macs()
        total_dur, total_params = self.get_total_duration(), self.get_total_params()
        self.flops, self.macs, self.params = total_flops, total_macs, total_params
        p("\n-------------------------- DeepSpeed Flops Profiler --------------------------")
        p(f"Profile Summary at step {profile_step}:")
        p("Notations:\ndata parallel size (dp_size), model parallel size(mp_size),\nnumber of parameters (params), "
          "number of multiply-accumulate operations(MACs),\nnumber of floating-point operations (flops), floating-point "
          "operations per second (FLOPS),\nfwd latency (forward propagation latency), bwd latency (backward propagation "
          "latency),\nstep (weights update latency), iter latency (sum of fwd, bwd and step latency)\n")
        line = lambda k, v: p("{:<70}  {:<8}".format(k, v))
        eng = self.ds_engine
        if eng is not None:
            line("world size: ", eng.world_size)
            line("data parallel size: ", eng.dp_world_size)
            line("model parallel size: ", eng.mp_world_size)
            line("batch size per GPU: ", eng.train_micro_batch_size_per_gpu())
        line("params per GPU: ", params_to_string(total_params))
        line("params of model = params per GPU * mp_size: ", params_to_string(total_params * (eng.mp_world_size if eng else 1)))
        line("fwd MACs per GPU: ", macs_to_string(total_macs))
        line("fwd flops per GPU: ", number_to_string(total_flops))
        line("fwd flops of model = fwd flops per GPU * mp_size: ", number_to_string(total_flops * (eng.mp_world_size if eng else 1)))
        line("fwd latency: ", duration_to_string(total_dur))
        line("fwd FLOPS per GPU = fwd flops per GPU / fwd latency: ", flops_to_string(total_flops / total_dur if total_dur else 0))
        if eng is not None and getattr(eng, "wall_clock_breakdown", lambda: False)():
            try:
                from ...utils.timer import BACKWARD_GLOBAL_TIMER, FORWARD_GLOBAL_TIMER, STEP_GLOBAL_TIMER
                fwd = eng.timers(FORWARD_GLOBAL_TIMER).elapsed(False) / 1000.0
                bwd = eng.timers(BACKWARD_GLOBAL_TIMER).elapsed(False) / 1000.0
                stp = eng.timers(STEP_GLOBAL_TIMER).elapsed(False) / 1000.0
                line("bwd latency: ", duration_to_string(bwd))
                line("bwd FLOPS per GPU = 2 * fwd flops per GPU / bwd latency: ",
                     flops_to_string((2 + self.recompute_fwd_factor) * total_flops / bwd if bwd else 0))
                line("fwd+bwd FLOPS per GPU = 3 * fwd flops per GPU / (fwd+bwd latency): ",
                     flops_to_string((3 + self.recompute_fwd_factor) * total_flops / (fwd + bwd) if fwd + bwd else 0))
                line("step latency: ", duration_to_string(stp))
                it = fwd + bwd + stp
                line("iter latency: ", duration_to_string(it))
                line("FLOPS per GPU = 3 * fwd flops per GPU / iter latency: ",
                     flops_to_string((3 + self.recompute_fwd_factor) * total_flops / it if it else 0))
                line("samples/second: ", round(eng.train_micro_batch_size_per_gpu() * eng.world_size / it, DEFAULT_PRECISION) if it else 0)
            except Exception:
                pass
        self.print_model_aggregated_profile(module_depth=module_depth, top_modules=top_modules, file=out)
        if detailed:
            p("\n------------------------------ Detailed Profile per GPU ------------------------------")
            p("Each module profile is listed after its name in the following order: \nparams, percentage of total "
              "params, MACs, percentage of total MACs, fwd latency, percentage of total fwd latency, fwd FLOPS\n")

            def describe(mod):
                fl, mc = self._tree_sum(mod, "__flops__"), self._tree_sum(mod, "__macs__")
                pr = self._tree_sum(mod, "__params__") + self._tree_sum(mod, "__expert_params__")
                du = getattr(mod, "__duration__", 0.0) or sum(getattr(c, "__duration__", 0.0) for c in mod.children())
                return ", ".join([params_to_string(pr), "{:.2%} Params".format(pr / total_params if total_params else 0),
                                  macs_to_string(mc), "{:.2%} MACs".format(mc / total_macs if total_macs else 0),
                                  duration_to_string(du), "{:.2%} latency".format(du / total_dur if total_dur else 0),
                                  flops_to_string(fl / du if du else 0)])

            def walk(mod, name, depth):
                p("  " * depth + f"({name}): {type(mod).__name__}({describe(mod)})")
                if module_depth < 0 or depth < module_depth:
                    for cn, c in mod.named_children():
                        walk(c, cn, depth + 1)

            walk(self.model, type(self.model).__name__, 0)
        p("------------------------------------------------------------------------------")
        if output_file:
            out.close()

    def print_model_aggregated_profile(self, module_depth=-1, top_modules=1, file=None):
        """Top ``top_modules`` module classes by params / MACs / latency at every depth."""
        p = partial(print, file=file or sys.stdout)
        info = {}

        def walk(mod, depth):
            d = info.setdefault(depth, {})
            e = d.setdefault(type(mod).__name__, [0, 0, 0.0])
            e[0] += self._tree_sum(mod, "__macs__")
            e[1] += self._tree_sum(mod, "__params__") + self._tree_sum(mod, "__expert_params__")
            e[2] += getattr(mod, "__duration__", 0.0)
            for c in mod.children():
                walk(c, depth + 1)

        walk(self.model, 0)
        depth = max(info) if module_depth == -1 else min(module_depth, max(info))
        p("\n----------------------------- Aggregated Profile per GPU -----------------------------")
        p(f"Top {top_modules} modules in terms of params, MACs or fwd latency at different model depths:")
        for d in range(depth + 1):
            n = min(top_modules, len(info[d]))
            by = lambda i: {k: v[i] for k, v in sorted(info[d].items(), key=lambda kv: kv[1][i], reverse=True)[:n]}
            p(f"depth {d}:")
            p(f"    params      - { {k: params_to_string(v) for k, v in by(1).items()} }")
            p(f"    MACs        - { {k: macs_to_string(v) for k, v in by(0).items()} }")
            p(f"    fwd latency - { {k: duration_to_string(v) for k, v in by(2).items()} }")


# ------------------------------------------------------------------------------------ formatting
def number_to_string(num, units=None, precision=DEFAULT_PRECISION):
    if units is None:
        for scale, u in ((1e12, "T"), (1e9, "G"), (1e6, "M"), (1e3, "K")):
            if abs(num) >= scale:
                return f"{round(num / scale, precision):g} {u}"
        if abs(num) >= 1 or num == 0:
            return f"{round(num, precision):g} "
        for scale, u in ((1e-3, "m"), (1e-6, "u")):
            if abs(num) >= scale:
                return f"{round(num / scale, precision):g} {u}"
        return f"{num:.{precision}e} "
    scale = {"T": 1e12, "G": 1e9, "M": 1e6, "K": 1e3, "": 1, "m": 1e-3, "u": 1e-6}[units]
    return f"{round(num / scale, precision):g} {units}"


def macs_to_string(macs, units=None, precision=DEFAULT_PRECISION):
    return f"{number_to_string(macs, units=units, precision=precision)}MACs"


def flops_to_string(flops, units=None, precision=DEFAULT_PRECISION):
    return f"{number_to_string(flops, units=units, precision=precision)}FLOPS"


def bytes_to_string(b, units=None, precision=DEFAULT_PRECISION):
    return f"{number_to_string(b, units=units, precision=precision)}B"


def params_to_string(params_num, units=None, precision=DEFAULT_PRECISION):
    units = units.replace("B", "G") if units else units
    return number_to_string(params_num, units=units, precision=precision).replace("G", "B").strip()


def duration_to_string(duration, units=None, precision=DEFAULT_PRECISION):
    return f"{number_to_string(duration, units=units, precision=precision)}s"


def get_module_flops(module):
    return FlopsProfiler._tree_sum(module, "__flops__")


def get_module_macs(module):
    return FlopsProfiler._tree_sum(module, "__macs__")


def get_module_duration(module):
    d = getattr(module, "__duration__", 0.0)
    return d if d else sum(get_module_duration(c) for c in module.children())


def get_model_profile(model, input_shape=None, args=[], kwargs={}, print_profile=True, detailed=True, module_depth=-1,
                      top_modules=1, warm_up=1, as_string=True, output_file=None, ignore_modules=None, mode="forward"):
    """Profile one forward (or ``generate``) call of ``model``; returns ``(flops, macs, params)``."""
    assert isinstance(model, nn.Module), "model must be a PyTorch module"
    prof = FlopsProfiler(model)
    model.eval()
    if input_shape is not None:
        assert isinstance(input_shape, tuple) and len(input_shape) >= 1
        try:
            ref = next(model.parameters())
            inp = torch.ones(()).new_empty(input_shape, dtype=ref.dtype, device=ref.device)
        except StopIteration:
            inp = torch.ones(()).new_empty(input_shape)
        args = [inp]
    assert (len(args) > 0) or (len(kwargs) > 0), "args and/or kwargs must be specified if input_shape is None"
    call = model.generate if mode == "generate" else model
    with torch.no_grad():
        for _ in range(warm_up):
            call(*args, **kwargs)
        prof.start_profile(ignore_list=ignore_modules)
        call(*args, **kwargs)
    flops, macs, params = prof.get_total_flops(), prof.get_total_macs(), prof.get_total_params()
    if print_profile:
        prof.print_model_profile(profile_step=warm_up, module_depth=module_depth, top_modules=top_modules,
                                 detailed=detailed, output_file=output_file)
    prof.end_profile()
    if as_string:
        return number_to_string(flops), macs_to_string(macs), params_to_string(params)
    return flops, macs, params
