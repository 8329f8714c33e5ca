"""Declarative layer containers (reference: ``inference/v2/model_implementations/layer_container_base.py``).

::

    class Llama2TransformerContainer(LayerContainer):
        qkv_w: UnfusedQKVParameter
        attn_out_w: AttentionOutputParameter
        mlp_1_w: GatedMLPParameter
        ...
        PARAM_MAPPING = {"self_attn.q_proj.weight": "qkv_w.q_params", ...,
                         "block_sparse_moe.experts.*.w1.weight": "moe_mlp_1.gating_experts"}   # '*' = list index

Annotated attributes are instantiated per container; ``set_dependency(checkpoint_name, tensor)`` routes a checkpoint
tensor to ``attribute.dependency`` (one name may feed several targets); finished parameters appear as attributes holding
the final tensor.  ``Optional[...]`` annotations may stay empty (e.g. biases some checkpoints lack).
``to_layer_weights()`` hands the result to :class:`DSTransformerModelBase` in its fused layout."""
from __future__ import annotations

import re
import typing
from typing import Any, Dict, List, Optional, Union

import torch

from .parameter_base import ParameterBase

# container attribute (reference naming) -> key of the generic model's per-layer dictionary
LAYER_KEYS = {
    "qkv_w": "wqkv", "qkv_b": "bqkv", "attn_out_w": "wo", "attn_out_b": "bo", "mlp_1_w": "w_in", "mlp_1_b": "b_in",
    "mlp_2_w": "w_out", "mlp_2_b": "b_out", "attn_norm_gamma": "ln1_w", "attn_norm_beta": "ln1_b",
    "mlp_norm_gamma": "ln2_w", "mlp_norm_beta": "ln2_b", "ln_gamma": "ln1_w", "ln_beta": "ln1_b",
    "moe_gate": "moe_gate", "moe_mlp_1": "moe_w1", "moe_mlp_2": "moe_w2", "shared_moe_mlp_1": "shared_w_in",
    "shared_moe_mlp_2": "shared_w_out", "shared_moe_gate": "shared_gate",
    # non-transformer container
    "word_emb": "embed", "word_emb_pos": "pos_embed", "word_unembed": "unembed", "word_unembed_bias": "unembed_b",
    "final_norm": "final_norm_w", "final_norm_w": "final_norm_w", "final_norm_b": "final_norm_b",
    "final_norm_gamma": "final_norm_w", "final_norm_beta": "final_norm_b",
}


def _unwrap_optional(ann):
    if typing.get_origin(ann) is Union:
        args = [a for a in typing.get_args(ann) if a is not type(None)]
        return (args[0], True) if len(args) == 1 else (ann, False)
    return ann, False


class LayerContainer:
    PARAM_MAPPING: Dict[str, Union[str, List[str]]] = {}

    def __init_subclass__(cls, **kw):
        super().__init_subclass__(**kw)
        params: Dict[str, type] = dict(getattr(cls, "_param_types", {}))
        optional = set(getattr(cls, "_optional_params", ()))
        hints = typing.get_type_hints(cls) if cls.__dict__.get("__annotations__") else {}
        for name, ann in hints.items():
            t, opt = _unwrap_optional(ann)
            if isinstance(t, type) and issubclass(t, ParameterBase):
                params[name] = t
                if opt:
                    optional.add(name)
        cls._param_types, cls._optional_params = params, frozenset(optional)
        # compile the mapping once: checkpoint-name pattern -> [(attr, dep)]
        compiled = []
        for src, targets in cls.PARAM_MAPPING.items():
            targets = [targets] if isinstance(targets, str) else list(targets)
            pairs = []
            for t in targets:
                attr, _, dep = t.partition(".")
                if attr not in params:
                    raise TypeError(f"{cls.__name__}.PARAM_MAPPING targets unknown parameter {attr!r}")
                if dep and dep not in params[attr]._tensor_deps and dep not in params[attr]._list_deps:
                    raise TypeError(f"{params[attr].__name__} has no dependency {dep!r}")
                pairs.append((attr, dep))
            rx = re.compile("^" + re.escape(src).replace(r"\*", r"(\d+)") + "$")
            compiled.append((src, rx, pairs))
        cls._compiled_mapping = compiled

    def __init__(self, model: Any) -> None:
        self.inference_model = model
        self._finalized_tensors: Dict[str, torch.Tensor] = {}
        self._params: Dict[str, ParameterBase] = {}
        for name, t in self._param_types.items():
            p = t(model, self)
            p._name_in_parent = name
            self._params[name] = p

    # ------------------------------------------------------------------------------------- feeding
    def _parameter_finalized(self, name: str, value: torch.Tensor) -> None:
        self._finalized_tensors[name] = value

    def direct_injection(self, name: str, tensor: torch.Tensor) -> None:
        """Bypass assembly: ``tensor`` IS the final value of parameter ``name`` (reloading a serialized model)."""
        if name not in self._param_types:
            raise ValueError(f"cannot inject unknown parameter {name}")
        self._finalized_tensors[name] = tensor

    def set_dependency(self, dep_name: str, dep_value: torch.Tensor) -> bool:
        """``dep_name``: checkpoint name relative to this container (layer prefix stripped).  Returns whether any
        mapping consumed it."""
        hit = False
        for src, rx, pairs in self._compiled_mapping:
            m = rx.match(dep_name)
            if not m:
                continue
            hit = True
            for attr, dep in pairs:
                p = self._params[attr]
                if m.groups():
                    getattr(p, dep)[int(m.group(1))] = dep_value
                else:
                    setattr(p, dep, dep_value)
        return hit

    # ------------------------------------------------------------------------------------- state
    def __getattr__(self, name):
        ft = self.__dict__.get("_finalized_tensors", {})
        if name in ft:
            return ft[name]
        if name in self.__dict__.get("_params", {}):
            return None                     # declared, not (yet) finalized
        raise AttributeError(name)

    @property
    def n_params(self) -> int:
        return len(self._param_types)

    @property
    def annotation_attrs(self) -> list:
        return list(self._param_types)

    @property
    def mapping_params(self) -> dict:
        return dict(self.PARAM_MAPPING)

    def unfinished(self) -> List[str]:
        return [n for n in self._param_types if n not in self._finalized_tensors and n not in self._optional_params]

    @property
    def is_populated(self) -> bool:
        """Every required parameter has its final tensor."""
        return not self.unfinished()

    @property
    def is_initialized(self) -> bool:
        """Populated, and every tensor lives on the serving device."""
        if not self.is_populated:
            return False
        dev = "cuda" if torch.cuda.is_available() else "cpu"
        return all(t.device.type == dev for t in self._finalized_tensors.values())

    def to_layer_weights(self):
        from .inference_transformer_base import LayerWeights
        lw = LayerWeights()
        for name, t in self._finalized_tensors.items():
            lw[LAYER_KEYS.get(name, name)] = t
        return lw
