"""Post-training INT8/INT4 weight quantisation of transformer checkpoints and live models.

Reference: ``deepspeed/runtime/weight_quantizer.py:11-153`` — group-wise symmetric quantisation of the four big
matrices of every layer (QKV, attention output, MLP in/out; MLP matrices optionally with twice the groups), scales
collected per layer in the order ``[qkv, dense, h4h, 4hh]`` and merged into one ``[layers, 4, max_groups]`` tensor that
the int8 inference kernels consume; ``merge_scales_split`` for TP-split checkpoints.  Quantisation here is one
vectorised pass per tensor (no Python loop over groups)."""
from __future__ import annotations

from typing import Dict, List, Optional

import torch

from ..module_inject.replace_policy import HFBertLayerPolicy, replace_policies


def _dev():
    return torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")


class WeightQuantization:

    def __init__(self, mlp_extra_grouping=True, mp_size=1):
        self.dense_scales: List[torch.Tensor] = []
        self.qkv_scales: List[torch.Tensor] = []
        self.mlp4hh_scales: List[torch.Tensor] = []
        self.mlph4h_scales: List[torch.Tensor] = []
        self.mlp_extra_grouping = mlp_extra_grouping
        self.mp_size = mp_size

    def quantize_data(self, data, quantize_bits, groups, key=None):
        """-> ``(int8 tensor of data.shape, scale [groups, 1])`` with ``q = round(x * scale)``,
        ``scale = 2^bits / (2 * absmax + 1e-5)``."""
        flat = data.float().reshape(groups, -1)
        absmax = torch.maximum(flat.amax(dim=1), flat.amin(dim=1).abs())
        scale = float(1 << quantize_bits) / (2 * absmax + 1e-5)
        lo, hi = -(1 << (quantize_bits - 1)), (1 << (quantize_bits - 1)) - 1
        q = (flat * scale[:, None]).round().clamp(lo, hi).reshape(data.shape).to(torch.int8)
        return q, scale.view(-1, 1)

    def is_mlp(self, data, merge_count=1):
        return ((self.mp_size * data.shape[0] * merge_count) / data.shape[1] == 4
                or (self.mp_size * data.shape[1] * merge_count) / data.shape[0] == 4)

    def is_qkv(self, data):
        return ((self.mp_size * data.shape[0]) / data.shape[1] == 3
                or (self.mp_size * data.shape[1]) / data.shape[0] == 3)

    def Quantize(self, value_list, quantize_bits, groups, key, merge_dim=0):
        if self.mlp_extra_grouping and self.is_mlp(value_list[0], merge_count=len(value_list)):
            groups *= 2
        scales = []
        for i, data in enumerate(value_list):
            q, s = self.quantize_data(data, quantize_bits, groups, key)
            scales.append(s)
            value_list[i] = q
        inv = (1 / torch.cat(scales, dim=merge_dim).to(_dev()).view(-1).unsqueeze(0))
        if "mlp.dense_4h_to_h.weight" in key:
            self.mlp4hh_scales.append(inv)
        elif "mlp.dense_h_to_4h.weight" in key:
            self.mlph4h_scales.append(inv)
        elif "attention.query_key_value.weight" in key:
            self.qkv_scales.append(inv)
        else:
            self.dense_scales.append(inv)
        return value_list

    def merge_layer_scales(self, layer_scales):
        width = max(s.shape[-1] for s in layer_scales)
        padded = [torch.nn.functional.pad(s, (0, width - s.shape[-1])) for s in layer_scales]
        return torch.cat(padded).unsqueeze(0)

    def merge_scales(self):
        per_layer = [self.merge_layer_scales([qkv, dense, h4h, fhh]) for dense, qkv, fhh, h4h in
                     zip(self.dense_scales, self.qkv_scales, self.mlp4hh_scales, self.mlph4h_scales)]
        return torch.cat(per_layer)

    def merge_scales_split(self, split_count):
        out: List[List[torch.Tensor]] = [[] for _ in range(split_count)]
        for dense, qkv, fhh, h4h in zip(self.dense_scales, self.qkv_scales, self.mlp4hh_scales, self.mlph4h_scales):
            cut = lambda t: torch.split(t, t.numel() // split_count)   # noqa: E731
            dense_s, qkv_s, fhh_s, h4h_s = cut(dense), cut(qkv), cut(fhh), cut(h4h)
            for s in range(split_count):
                out[s].append(torch.cat([torch.cat((qkv_s[s], torch.zeros_like(qkv_s[s])), dim=1),
                                         torch.cat((dense_s[s], torch.zeros_like(dense_s[s])), dim=1),
                                         h4h_s[s], fhh_s[s]]).unsqueeze(0))
        return out

    _MEGATRON_KEYS = ("attention.dense.weight", "mlp.dense_4h_to_h.weight", "mlp.dense_h_to_4h.weight",
                      "attention.query_key_value.weight")

    def sd_quantize_megatron(self, sd, quantize_bits, groups):
        for key in list(sd.keys()):
            if any(k in key for k in self._MEGATRON_KEYS):
                sd[key] = self.Quantize([sd[key]], quantize_bits, groups, key=key)[0]
        return sd, self.merge_scales()

    def model_quantize(self, model, quantize_policy, quantize_bits, groups):
        """Quantise, in place, the 2-D linear weights of every layer matched by a policy; returns
        ``(model, scales [layers, n_mats, max_groups])``.  ``quantize_policy``: ``{layer_class: policy}`` or ``None``
        for the built-in policy list."""
        policies: Dict[type, object] = {}
        by_name = {}
        if quantize_policy is not None:
            policies.update(quantize_policy)
        else:
            by_name = {p._orig_layer_class_name: p for p in replace_policies}
        all_scales = []

        def quantize_layer(layer, policy_cls):
            layer_scales = []
            for _, lin in layer.named_modules():
                w = getattr(lin, "weight", None)
                if not isinstance(lin, torch.nn.Linear) or w is None or w.dim() != 2:
                    continue
                g = groups
                if self.mlp_extra_grouping and self.is_mlp(w):
                    g = groups * 2
                elif policy_cls is HFBertLayerPolicy and self.is_qkv(w):
                    g = groups * 3
                if w.numel() % g != 0:
                    continue
                q, s = self.quantize_data(w.data, quantize_bits, g)
                w.data.copy_(q.to(w.dtype))
                layer_scales.append(1 / s.to(_dev()).view(-1).unsqueeze(0))
            if layer_scales:
                all_scales.append(self.merge_layer_scales(layer_scales))

        def walk(m):
            for _, child in m.named_children():
                pol = policies.get(type(child)) or by_name.get(type(child).__name__)
                if pol is not None:
                    quantize_layer(child, pol)
                else:
                    walk(child)

        walk(model)
        width = max((s.shape[-1] for s in all_scales), default=0)
        rows = max((s.shape[1] for s in all_scales), default=0)
        padded = [torch.nn.functional.pad(s, (0, width - s.shape[-1], 0, rows - s.shape[1])) for s in all_scales]
        return model, (torch.cat(padded) if padded else torch.empty(0))
