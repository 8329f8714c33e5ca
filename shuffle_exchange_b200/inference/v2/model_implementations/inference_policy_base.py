"""Policies: turn a HuggingFace config + parameter stream into a :class:`DSTransformerModelBase`.

Reference: ``inference/v2/model_implementations/inference_policy_base.py`` (``InferenceV2Policy.build_model`` pulls
parameters from a checkpoint engine into per-layer containers through a ``ContainerMap`` of name prefixes) plus
``layer_container_base.py`` / ``parameter_base.py`` (declarative fused parameters such as ``UnfusedQKVParameter``).
Same flow here with plain data: a policy declares

* ``spec(model_config) -> ArchSpec``;
* ``layer_prefix`` (e.g. ``"model.layers."``) and ``layer_rules``: ``{target: source}`` where ``source`` is one HF name
  (relative to the layer) or a tuple of names that are concatenated along dim 0 (unfused q/k/v, gate/up);
* ``global_rules`` for embeddings, final norm and unembedding;
* optional hooks (``transform_layer``) for layouts that need more than concatenation (Falcon's grouped QKV, experts).
"""
from __future__ import annotations

import re
from abc import ABC, abstractmethod
from typing import Any, Dict, Iterable, List, Optional, Tuple, Union

import torch

from ..checkpoint.base_engine import CheckpointEngineBase
from ..config_v2 import RaggedInferenceEngineConfig
from .inference_transformer_base import ArchSpec, DSTransformerModelBase, LayerWeights

Source = Union[str, Tuple[str, ...]]


class ContainerMap:
    """Name-prefix bookkeeping kept for API parity: records which checkpoint names were consumed so that unmapped
    parameters can be reported (``validate``)."""

    def __init__(self) -> None:
        self.consumed: set = set()
        self.unmapped: List[str] = []

    def mark(self, name: str) -> None:
        self.consumed.add(name)

    def set_unmapped_params(self, names: Iterable[str]) -> None:
        self.unmapped = list(names)

    def validate(self, allowed_unmapped: Iterable[str] = ()) -> None:
        pats = [re.compile(p) for p in allowed_unmapped]
        bad = [n for n in self.unmapped if not any(p.search(n) for p in pats)]
        if bad:
            raise RuntimeError(f"checkpoint parameters without a destination: {bad[:8]}{' ...' if len(bad) > 8 else ''}")


class InferenceV2Policy(ABC):
    """``policy(state_manager)`` builds the model (the engine calls it twice: once with ``None`` to size the KV
    cache, once with the real state manager)."""

    layer_prefix: str = "model.layers."
    layer_rules: Dict[str, Source] = {}
    global_rules: Dict[str, Source] = {}
    ignore_patterns: Tuple[str, ...] = (r"rotary_emb\.inv_freq", r"masked_bias", r"(^|\.)attn\.bias$")

    def __init__(self, model_config: Any, checkpoint_engine: Optional[CheckpointEngineBase] = None,
                 inf_checkpoint_path: Optional[str] = None) -> None:
        self._model_config = model_config
        self._checkpoint_engine = checkpoint_engine
        self._inf_checkpoint_path = inf_checkpoint_path
        self._weights: Optional[Dict] = None
        self._engine_config: Optional[RaggedInferenceEngineConfig] = None
        self.container_map = ContainerMap()
        self.dtype_override: Optional[torch.dtype] = None      # e.g. the dtype of an in-memory model's parameters
        self._mp_group = None

    def set_mp_group(self, group) -> None:
        """Tensor-parallel group the engine created; models built afterwards shard their weights over it."""
        self._mp_group = group

    def resolved_spec(self) -> ArchSpec:
        spec = self.spec(self._model_config)
        if self.dtype_override is not None:
            spec.dtype = self.dtype_override
        return spec

    # ------------------------------------------------------------------------------------------ to implement
    @abstractmethod
    def spec(self, model_config: Any) -> ArchSpec:
        ...

    def transform_layer(self, lw: LayerWeights, raw: Dict[str, torch.Tensor], spec: ArchSpec) -> None:
        """Hook for layouts that need more than concatenation; ``raw`` holds every tensor of the layer by its
        layer-relative HF name."""

    def transform_globals(self, weights: Dict, raw: Dict[str, torch.Tensor], spec: ArchSpec) -> None:
        pass

    # ------------------------------------------------------------------------------------------ loading
    def _device(self):
        return torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")

    def _gather(self, raw: Dict[str, torch.Tensor], src: Source) -> Optional[torch.Tensor]:
        names = (src,) if isinstance(src, str) else src
        if any(n not in raw for n in names):
            return None
        return raw[names[0]] if len(names) == 1 else torch.cat([raw[n] for n in names], dim=0)

    def _load_flat(self, rank: int, size: int) -> Dict:
        """Weights written by ``InferenceEngineV2.serialize``: already fused, sharded for ``(rank, size)`` and
        transformed, restored as views into one buffer (reference ``inference_policy_base.py:156-180``)."""
        from .flat_model_helpers import (ModelMetadata, make_metadata_filename, make_param_filename,
                                         restore_inference_model)
        base = self._inf_checkpoint_path
        with open(make_metadata_filename(base, rank, size)) as f:
            md = ModelMetadata.parse_raw(f.read())
        buf = torch.load(make_param_filename(base, rank, size), map_location="cpu", weights_only=True).to(self._device())
        weights = restore_inference_model(buf, md)
        weights["_transformed"] = True
        weights["_presharded"] = (rank, size)
        weights["_flat"] = (buf, md)
        if isinstance(weights.get("embed"), torch.Tensor) and weights["embed"].is_floating_point():
            self.dtype_override = weights["embed"].dtype
        return weights

    def load_weights(self) -> Dict:
        if self._weights is not None:
            return self._weights
        if self._inf_checkpoint_path is not None:
            import torch.distributed as dist
            g = self._mp_group
            rank, size = (dist.get_rank(g), dist.get_world_size(g)) if g is not None else (0, 1)
            self._weights = self._load_flat(rank, size)
            return self._weights
        spec = self.resolved_spec()
        dev, dt = self._device(), spec.dtype
        per_layer: List[Dict[str, torch.Tensor]] = [dict() for _ in range(spec.num_layers)]
        glob: Dict[str, torch.Tensor] = {}
        pat = re.compile("^" + re.escape(self.layer_prefix) + r"(\d+)\.(.+)$")
        ignore = [re.compile(p) for p in self.ignore_patterns]
        unmapped = []
        for name, t in self._checkpoint_engine.parameters():
            if any(p.search(name) for p in ignore):
                continue
            t = t.detach()
            t = t.to(dev, dt) if t.is_floating_point() else t.to(dev)
            m = pat.match(name)
            if m:
                per_layer[int(m.group(1))][m.group(2)] = t
            else:
                glob[name] = t
        weights: Dict[str, Any] = {"layers": []}
        used_global = set()
        for target, src in self.global_rules.items():
            v = self._gather(glob, src)
            if v is not None:
                weights[target] = v
                used_global.update((src,) if isinstance(src, str) else src)
        self.transform_globals(weights, glob, spec)
        if "unembed" not in weights:                       # tied embeddings
            weights["unembed"] = weights["embed"]
        unmapped += [n for n in glob if n not in used_global and n not in getattr(self, "_extra_used_globals", ())]
        for li, raw in enumerate(per_layer):
            lw = LayerWeights()
            used = set()
            for target, src in self.layer_rules.items():
                v = self._gather(raw, src)
                if v is not None:
                    lw[target] = v.contiguous()
                    used.update((src,) if isinstance(src, str) else src)
            before = set(raw)
            self.transform_layer(lw, raw, spec)
            used.update(getattr(self, "_extra_used_layer", ()))
            unmapped += [f"{self.layer_prefix}{li}.{n}" for n in before if n not in used]
            weights["layers"].append(lw)
        self.container_map.set_unmapped_params(unmapped)
        self._weights = weights
        return weights

    # ------------------------------------------------------------------------------------------ build
    def build_model(self, engine_config: RaggedInferenceEngineConfig, mp_group=None,
                    state_manager=None) -> DSTransformerModelBase:
        self._engine_config = engine_config
        spec = self.resolved_spec()
        if mp_group is not None:
            self._mp_group = mp_group
        weights = self.load_weights()
        spec = self.resolved_spec()             # a flat checkpoint fixes the activation dtype
        model = DSTransformerModelBase(spec, weights, engine_config, state_manager, self._mp_group)
        model.model_config = self._model_config
        return model

    def __call__(self, state_manager):
        return self.build_model(self._engine_config or RaggedInferenceEngineConfig(), self._mp_group, state_manager)
