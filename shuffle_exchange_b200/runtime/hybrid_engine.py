"""Hybrid engine: one engine object that trains (ZeRO / LoRA) and generates (RLHF actor roll-outs).

Reference: ``deepspeed/runtime/hybrid_engine.py:30-445`` — ``DeepSpeedHybridEngine(DeepSpeedEngine)`` with
``generate()`` (gather ZeRO-3 parameters, fuse LoRA, optional inference TP, run ``module.generate``, unfuse),
``eval()/train()`` mode switches with latency accounting, and ``step()`` bookkeeping.  The reference swaps every
transformer layer for a kernel-injection "inference container"; here the training modules already run on the framework's
sm_100a kernels (tcgen05 GEMM, flash attention, fused norms), so generation reuses the SAME module objects and the
mode switch is only about *where the weights live*:

* ZeRO-3 + ``pin_parameters``: all shards are gathered once per ``generate`` call (copy-engine all-gather over NVLink,
  180 GB of HBM holds a full bf16 replica of anything up to ~70B) instead of layer by layer per decoded token;
  without pinning the ordinary ZeRO-3 fetch/release hooks serve every forward.
* LoRA: adapters are folded into the base weight (``W += scale * B A``) for the duration of the roll-out so decode
  GEMMs are single launches, then unfolded exactly (the delta is recomputed, not subtracted from a rounded sum, when a
  master copy exists).
* ``inference_tp_size > 1``: prompts are all-gathered over the TP group so every TP rank decodes the same batch, and the
  caller gets back its own slice.
* ``release_inference_cache``: the KV/workspace arena is returned to the allocator between roll-outs.
"""
from __future__ import annotations

import gc
import math
import time
from typing import Any, Dict, List, Optional

import torch

from .. import comm as dist
from ..utils.logging import logger
from .engine import DeepSpeedEngine
from .zero.partition_parameters import GatheredParameters


def _lora_triplets(module: torch.nn.Module):
    """Yield ``(owner, base_weight, right, left, scaling)`` for every LoRA-carrying submodule.  Two conventions are
    recognised: DeepSpeed-Chat's ``lora_right_weight [in, r]`` / ``lora_left_weight [r, out]`` / ``lora_scaling`` on a
    linear-like module with ``.weight [out, in]``, and this framework's ``LoRAOptimizedLinear`` (``lora_weight_1 =
    A [r, in]``, ``lora_weight_2 = B [out, r]``)."""
    for m in module.modules():
        if hasattr(m, "lora_right_weight") and hasattr(m, "lora_left_weight") and hasattr(m, "weight"):
            yield m, m.weight, m.lora_right_weight, m.lora_left_weight, float(getattr(m, "lora_scaling", 1.0))
        elif hasattr(m, "lora_weight_1") and hasattr(m, "lora_weight_2") and hasattr(m, "weight") \
                and getattr(m, "weight", None) is not None and m.weight.dim() == 2:
            a, b = m.lora_weight_1.weight, m.lora_weight_2.weight          # [r, in], [out, r]
            cfg = getattr(m, "lora_config", None)
            scale = float(cfg.lora_alpha / cfg.lora_r) if cfg is not None else 1.0
            yield m, m.weight, a.t(), b.t(), scale


class DeepSpeedHybridEngine(DeepSpeedEngine):
    """Training engine with a fast ``generate`` path."""
    inference_mp_group = None

    def __init__(self, args, model, **kwargs):
        super().__init__(args, model, **kwargs)
        if torch.cuda.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            # all ranks must sample identically when they decode replicated prompts
            state = torch.cuda.get_rng_state().cuda()
            dist.broadcast(state, 0)
            torch.cuda.set_rng_state(state.cpu())
        he = self._config.hybrid_engine
        self.Z3_enabled = (self._config.zero_optimization_stage == 3)
        self.gather_all_layers = he.pin_parameters
        self._inference_tp = max(1, int(he.inference_tp_size))
        self.mp_group = None
        if self._inference_tp > 1:
            self._create_inference_tp_groups()
        # populated lazily: LoRA owners and their parameters
        self._lora: List[Any] = list(_lora_triplets(self.module))
        self.all_lora_params = [p for (_, _, r, l, _) in self._lora for p in (r, l)]
        self.is_lora_fused = False
        self._lora_backup: Dict[int, torch.Tensor] = {}
        # performance stats (same breakdown the reference prints)
        self._t_start = None
        self._total_latency = 0.0
        self._iters = 0
        self._training_start_time = None
        self._generate_latency = 0.0
        self._training_latency = 0.0
        self._gather_latency = 0.0
        self._total_batch_size = None
        self._workspace_released = False

    # ------------------------------------------------------------------------------------------ groups
    def _create_inference_tp_groups(self):
        world = dist.get_world_size()
        tp = self._inference_tp
        assert world % tp == 0, f"inference_tp_size {tp} must divide the world size {world}"
        rank = dist.get_rank()
        for i in range(world // tp):
            ranks = list(range(i * tp, (i + 1) * tp))
            g = dist.new_group(ranks)
            if rank in ranks:
                self.mp_group = g
        DeepSpeedHybridEngine.inference_mp_group = self.mp_group

    # ------------------------------------------------------------------------------- inference containers
    # Optional fused-layer path (reference ``hybrid_engine.py:78-176``): every layer an injection policy recognises gets
    # a hybrid container whose fused inference layer is refreshed from the training weights before generation.
    # ``generate`` uses the training module directly unless ``create_inference_module`` was called.
    def populate_all_inference_policies(self):
        from ..module_inject.replace_policy import generic_policies, replace_policies
        self.inference_policies = {}
        for pol in replace_policies:
            name = getattr(pol, "_orig_layer_class_name", None)
            if name:
                self.inference_policies[name] = pol
        self.inference_policies.update({nn_cls: self._replace_linear_layer for nn_cls in (torch.nn.Linear,)})
        self._generic_policies = list(generic_policies)
        return self.inference_policies

    def _replace_linear_layer(self, child, name=None):
        return child

    def new_inference_container(self, orig_layer, policy_cls, layer_id):
        from ..inference.config import DeepSpeedInferenceConfig
        from ..module_inject.containers import POLICY_TO_CONTAINER
        from ..module_inject.containers.features import HybridEngineContainer
        base = POLICY_TO_CONTAINER[policy_cls]
        hybrid_cls = type("Hybrid" + base.__name__, (HybridEngineContainer, base), {})
        dtype = torch.bfloat16 if self.bfloat16_enabled() else (torch.float16 if self.fp16_enabled() else torch.float32)
        icfg = DeepSpeedInferenceConfig(dtype=dtype, tensor_parallel={"tp_size": self._inference_tp},
                                        max_out_tokens=self._config.hybrid_engine.max_out_tokens)
        if self.mp_group is not None:
            icfg.tensor_parallel.tp_group = self.mp_group
        container = hybrid_cls(policy_cls(orig_layer, inference=True), icfg, getattr(self.module, "config", None),
                               layer_id, orig_layer)
        container.create_ds_model_config()
        container.initialize_tensors()
        container.apply_tensor_parallelism()
        container.create_module()
        return container

    def create_inference_containers(self, module, layer_id=0):
        if not hasattr(self, "inference_policies"):
            self.populate_all_inference_policies()
            self._inference_containers, self._orig_layers = [], []
        for name, child in module.named_children():
            pol = self.inference_policies.get(type(child).__name__)
            if isinstance(pol, type) and pol.match(child):
                c = self.new_inference_container(child, pol, layer_id)
                self._inference_containers.append(c)
                self._orig_layers.append((module, name, child))
                layer_id += 1
            else:
                layer_id = self.create_inference_containers(child, layer_id)
        return layer_id

    def create_inference_module(self):
        """Build the containers (once).  ``refresh_inference_module`` re-reads the training weights; swapping the fused
        layers into the module tree for a generation phase is ``use_inference_layers(True)``."""
        self._inference_containers, self._orig_layers = [], []
        self.populate_all_inference_policies()
        self.create_inference_containers(self.module)
        return len(self._inference_containers)

    def refresh_inference_module(self):
        for c in getattr(self, "_inference_containers", ()):
            c.refresh()

    def use_inference_layers(self, enable: bool):
        from ..module_inject.containers import HFLayerAdapter
        for c, (parent, name, orig) in zip(getattr(self, "_inference_containers", ()), getattr(self, "_orig_layers", ())):
            if enable:
                setattr(parent, name, HFLayerAdapter(c.module, returns_tuple=type(c.policy).returns_tuple))
            else:
                setattr(parent, name, orig)

    def convert_to_linear_transposed(self, model):
        """Replace ``nn.Linear`` layers by ``TLinear`` (weight stored transposed: a ZeRO-3 gathered weight feeds
        ``matmul`` without a transpose copy)."""
        from .utils import TLinear

        def walk(m):
            for name, child in list(m.named_children()):
                if isinstance(child, torch.nn.Linear) and not isinstance(child, TLinear):
                    setattr(m, name, TLinear(child, name=name))
                else:
                    walk(child)
        walk(model)
        return model

    # ------------------------------------------------------------------------------------------ LoRA
    def fuse_lora_weight(self):
        if self.is_lora_fused:
            return
        with torch.no_grad():
            for owner, w, right, left, scale in self._lora:
                if w.numel() == 0:           # ZeRO-3 shard not gathered: nothing to fold into
                    continue
                delta = (scale * torch.matmul(left.t().float(), right.t().float())).to(w.dtype)   # [out, in]
                self._lora_backup[id(owner)] = delta
                w.add_(delta)
        self.is_lora_fused = True

    def unfuse_lora_weight(self):
        if not self.is_lora_fused:
            return
        with torch.no_grad():
            for owner, w, right, left, scale in self._lora:
                delta = self._lora_backup.pop(id(owner), None)
                if delta is not None and w.numel() == delta.numel():
                    w.sub_(delta)
        self.is_lora_fused = False

    def unfuse_lora_weight_non_pinned(self):
        params = [p for (_, w, r, l, _) in self._lora for p in (w, r, l)]
        with GatheredParameters(params, enabled=self.Z3_enabled):
            self.unfuse_lora_weight()

    # ------------------------------------------------------------------------------------------ workspace
    def retake_inference_cache(self):
        """Nothing is pre-reserved: KV caches are allocated by ``module.generate`` from the caching allocator; after a
        release we only make sure fragmentation does not make the first allocation fail."""
        if self._config.hybrid_engine.release_inference_cache and self._workspace_released:
            self._workspace_released = False

    def _release_inference_cache(self):
        if self._config.hybrid_engine.release_inference_cache:
            gc.collect()
            if torch.cuda.is_available():
                torch.cuda.empty_cache()
            self._workspace_released = True

    # ------------------------------------------------------------------------------------------ generate
    def _generate(self, *inputs, **kwargs):
        was_training = self.module.training
        self.module.eval()
        try:
            with torch.no_grad():
                return self.module.generate(*inputs, **kwargs)
        finally:
            if was_training:
                self.module.train()

    def _all_params(self) -> List[torch.nn.Parameter]:
        return list(self.module.parameters())

    def generate(self, *inputs, **kwargs):
        ids = inputs[0] if len(inputs) > 0 else kwargs["input_ids"]
        if self._total_batch_size is None:
            world = dist.get_world_size() if dist.is_initialized() else 1
            self._total_batch_size = ids.shape[0] * world
        self._t0 = time.time()
        self._gather_latency = 0.0
        local_bs = ids.shape[0]
        if self._inference_tp > 1:
            gathered = torch.zeros((local_bs * self._inference_tp,) + tuple(ids.shape[1:]), dtype=ids.dtype,
                                   device=ids.device)
            dist.all_gather_into_tensor(gathered, ids.contiguous(), group=self.mp_group)
            if len(inputs) > 0:
                inputs = (gathered, *inputs[1:])
            else:
                kwargs["input_ids"] = gathered
            am = kwargs.get("attention_mask")
            if torch.is_tensor(am):
                gm = torch.zeros((local_bs * self._inference_tp,) + tuple(am.shape[1:]), dtype=am.dtype,
                                 device=am.device)
                dist.all_gather_into_tensor(gm, am.contiguous(), group=self.mp_group)
                kwargs["attention_mask"] = gm

        has_lora = len(self._lora) > 0
        if self.Z3_enabled and self.gather_all_layers:
            # one gather for the whole roll-out; layer groups bound the transient peak when asked to
            part = int(self._config.hybrid_engine.tp_gather_partition_size)
            params = self._all_params()
            with GatheredParameters(params):
                if torch.cuda.is_available():
                    torch.cuda.synchronize()
                self._gather_latency = time.time() - self._t0
                if has_lora:
                    self.fuse_lora_weight()
                self.retake_inference_cache()
                out = self._generate(*inputs, **kwargs)
                if has_lora:
                    self.unfuse_lora_weight()
            del part
        else:
            if has_lora and not self.Z3_enabled:
                self.fuse_lora_weight()
            self.retake_inference_cache()
            out = self._generate(*inputs, **kwargs)
            if has_lora:
                if not self.Z3_enabled:
                    self.unfuse_lora_weight()
                else:
                    self.unfuse_lora_weight_non_pinned()
                self.is_lora_fused = False
        if self._inference_tp > 1 and torch.is_tensor(out):
            r = dist.get_rank(group=self.mp_group)
            out = out[local_bs * r:local_bs * (r + 1)]
        self._release_inference_cache()
        self._generate_latency = time.time() - self._t0 - self._gather_latency
        return out

    # ------------------------------------------------------------------------------------------ mode switches
    def eval(self):
        if self._t_start is not None:
            latency = time.time() - self._t_start
            self._total_latency += latency
            self._iters += 1
            if not dist.is_initialized() or dist.get_rank() == 0:
                msg = (f"|E2E latency={latency:.2f}s "
                       f"|Gather latency={self._gather_latency:.2f}s ({self._gather_latency / latency * 100:.2f}%) "
                       f"|Generate time={self._generate_latency:.2f}s ({self._generate_latency / latency * 100:.2f}%) "
                       f"|Training time={self._training_latency:.2f}s ({self._training_latency / latency * 100:.2f}%) ")
                others = latency - (self._generate_latency + self._training_latency)
                msg += f"|Others={others:.2f} ({others / latency * 100:.2f}%)"
                if self._total_batch_size is not None:
                    msg += f"|CurSamplesPerSec={1 / latency * self._total_batch_size:.2f} "
                    msg += f"|AvgSamplesPerSec={1 / (self._total_latency / self._iters) * self._total_batch_size:.2f}"
                logger.info(msg)
            self._t_start = time.time()
        self._training_latency = 0.0
        super().eval()
        if self.Z3_enabled:
            gc.collect()
            if torch.cuda.is_available():
                torch.cuda.empty_cache()
        if self._t_start is None:
            self._t_start = time.time()
        return self

    def train(self, mode=True):
        if mode and self.is_lora_fused:
            self.unfuse_lora_weight()
        super().train(mode)
        if mode:
            self._training_start_time = time.time()
        return self

    def step(self, lr_kwargs=None):
        super().step(lr_kwargs=lr_kwargs)
        if self._training_start_time is not None:
            self._training_latency += time.time() - self._training_start_time
            self._training_start_time = time.time()

    # ------------------------------------------------------------------------------------------ introspection
    def latency_breakdown(self) -> Dict[str, float]:
        return {"gather": self._gather_latency, "generate": self._generate_latency,
                "training": self._training_latency, "total": self._total_latency, "iters": float(self._iters)}
