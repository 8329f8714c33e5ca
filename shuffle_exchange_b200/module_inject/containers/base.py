"""Kernel injection: turn a HuggingFace / Megatron transformer layer into a :class:`DeepSpeedTransformerInference`.

Reference: ``module_inject/containers/base.py`` (``BaseTransformerContainer``: ``create_ds_model_config``,
``initialize_tensors``, ``apply_tensor_parallelism``, ``copy_data_to_new_module``, ``create_module``) +
``module_inject/policy.py`` (``TransformerPolicy`` with ``get_hidden_heads / attention / mlp / layernorm``) and the
feature mix-ins ``containers/features/{meta_tensor,hybrid_engine,megatron,split_qkv,gated_mlp}.py``.

The structure here is flatter: a *policy* reads one client layer and answers with plain tensors in ``nn.Linear``
orientation (``[out, in]``) plus a handful of architecture switches; one container class does the rest (fused QKV
assembly, tensor-parallel slicing, transposition into the ``x @ W`` layout of the fused layer, module creation).  An
architecture is therefore one small policy class (``containers/<arch>.py``); nothing else varies.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Optional, Tuple

import torch
import torch.nn as nn

from ...utils.types import ActivationFuncType, NormType


@dataclass
class AttentionTensors:
    """q/k/v either fused (``qkvw [(hq + 2 hkv) d, H]`` in q|k|v order) or split; biases may be ``None``."""
    qkvw: Optional[torch.Tensor] = None
    qkvb: Optional[torch.Tensor] = None
    qw: Optional[torch.Tensor] = None
    qb: Optional[torch.Tensor] = None
    kw: Optional[torch.Tensor] = None
    kb: Optional[torch.Tensor] = None
    vw: Optional[torch.Tensor] = None
    vb: Optional[torch.Tensor] = None
    ow: Optional[torch.Tensor] = None
    ob: Optional[torch.Tensor] = None


@dataclass
class MLPTensors:
    """``inter_w [I, H]`` (or ``up_w`` + ``gate_w`` for gated activations), ``out_w [H, I]``."""
    inter_w: Optional[torch.Tensor] = None
    inter_b: Optional[torch.Tensor] = None
    up_w: Optional[torch.Tensor] = None
    gate_w: Optional[torch.Tensor] = None
    out_w: Optional[torch.Tensor] = None
    out_b: Optional[torch.Tensor] = None


@dataclass
class NormTensors:
    """``attn_*``: the norm in front of the MLP (post-attention); ``input_*``: the norm in front of attention (pre-LN)
    or the layer's final norm (post-LN)."""
    attn_nw: Optional[torch.Tensor] = None
    attn_nb: Optional[torch.Tensor] = None
    input_nw: Optional[torch.Tensor] = None
    input_nb: Optional[torch.Tensor] = None


class TransformerPolicy:
    """Describes one architecture's layer.  Class attributes are the architecture switches the fused layer needs;
    instances wrap one client layer (``client_module``)."""

    _orig_layer_class_name: Optional[str] = None
    # switches (reference ``policy.py:43-86``)
    inference: bool = True
    linear_layer: bool = True            # False: Conv1D-style [in, out] weights (GPT-2)
    scale_attention: bool = True
    is_megatron_v2: bool = False
    use_mup: bool = False
    mlp_act_func_type: ActivationFuncType = ActivationFuncType.GELU
    pre_attn_norm: bool = True
    use_load_prefix: bool = False
    split_qkv: bool = True
    norm_type: NormType = NormType.LayerNorm
    triangular_masking: bool = True
    mlp_after_attn: bool = True
    bigscience_bloom: bool = False
    rotate_half: bool = False
    rotate_every_two: bool = False
    window_size: int = 1
    local_attention: bool = False
    returns_tuple: bool = False          # client model indexes the layer output (``outputs[0]``)

    def __init__(self, client_module=None, inference: bool = True):
        self.client_module = client_module
        self.inference = inference

    @classmethod
    def match(cls, module) -> bool:
        return cls._orig_layer_class_name is not None and type(module).__name__ == cls._orig_layer_class_name

    # ------------------------------------------------------------------ to implement per architecture
    def get_hidden_heads(self) -> Tuple[int, int, float, int]:
        """(hidden_size, num_heads, layer-norm eps, intermediate_size)"""
        raise NotImplementedError

    def num_kv_heads(self) -> int:
        return -1

    def rotary(self) -> Tuple[int, float]:
        """(rotary_dim or -1, theta)"""
        return -1, 10000.0

    def attention(self) -> AttentionTensors:
        raise NotImplementedError

    def mlp(self) -> MLPTensors:
        raise NotImplementedError

    def layernorm(self) -> NormTensors:
        raise NotImplementedError

    # ------------------------------------------------------------------ helpers for subclasses
    @staticmethod
    def _wb(linear) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
        return linear.weight.data, (linear.bias.data if getattr(linear, "bias", None) is not None else None)

    @staticmethod
    def _conv1d(conv) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
        """HF ``Conv1D`` keeps ``[in, out]``: hand back the ``[out, in]`` view."""
        return conv.weight.data.t(), (conv.bias.data if getattr(conv, "bias", None) is not None else None)

    @staticmethod
    def _deinterleave_qkv(w: torch.Tensor, b: Optional[torch.Tensor], heads: int):
        """``[heads, 3, d]``-ordered fused QKV rows (BLOOM, GPT-NeoX, Megatron v2) -> q|k|v blocks."""
        d = w.shape[0] // (3 * heads)
        w = w.view(heads, 3, d, -1).permute(1, 0, 2, 3).reshape(3 * heads * d, -1)
        if b is not None:
            b = b.view(heads, 3, d).permute(1, 0, 2).reshape(-1)
        return w, b


DSPolicy = TransformerPolicy          # reference name


class HFLayerAdapter(nn.Module):
    """Sits where the client layer sat and speaks its calling convention: the first argument / ``hidden_states`` is the
    input, the first tensor among the remaining positionals or ``attention_mask`` is the mask (boolean keep-masks become
    additive), cache objects are ignored (the fused layer keeps its own KV cache), the output is a tensor or the client's
    tuple shape."""

    def __init__(self, ds_layer: nn.Module, returns_tuple: bool = False):
        super().__init__()
        self.ds_layer = ds_layer
        self.returns_tuple = returns_tuple

    @property
    def config(self):
        return self.ds_layer.config

    def forward(self, *args, **kwargs):
        hidden = args[0] if args else kwargs.pop("hidden_states", kwargs.pop("x", kwargs.pop("input", None)))
        mask = kwargs.get("attention_mask", kwargs.get("attn_mask"))
        if mask is None:
            mask = next((a for a in args[1:] if isinstance(a, torch.Tensor) and a.dim() >= 2), None)
        if mask is not None:
            if mask.dtype == torch.bool:
                mask = torch.zeros(mask.shape, dtype=hidden.dtype, device=mask.device).masked_fill(
                    ~mask, torch.finfo(hidden.dtype).min)
            elif not mask.is_floating_point():
                mask = (1 - mask.to(hidden.dtype)) * torch.finfo(hidden.dtype).min
            else:
                mask = mask.to(hidden.dtype)
        # positions come from the fused layer's own cache length; ALiBi slopes are the client's
        extra = {"alibi": kwargs["alibi"]} if kwargs.get("alibi") is not None else {}
        out = self.ds_layer(hidden, attention_mask=mask, **extra)
        if isinstance(out, tuple):
            out = out[0]
        return (out,) if self.returns_tuple else out


class BaseTransformerContainer:
    """Builds the fused layer for one client layer described by ``policy``."""

    def __init__(self, policy: TransformerPolicy, config, model_config, layer_id: int, child: nn.Module):
        self.policy = policy
        self.config = config                # DeepSpeedInferenceConfig (engine-level)
        self.model_config = model_config
        self.layer_id = layer_id
        self.child = child
        tp = getattr(config, "tensor_parallel", None)
        self.mp_size = getattr(tp, "tp_size", 1) if tp is not None else 1
        self.mp_group = getattr(tp, "tp_group", None) if tp is not None else None
        self.mp_rank = 0
        if self.mp_size > 1:
            from ... import comm as dist
            self.mp_rank = dist.get_rank(group=self.mp_group)
        self.dtype = getattr(config, "dtype", torch.float32)
        self.hidden_size, self.num_attention_heads, self.layernorm_epsilon, self.intermediate_size = \
            policy.get_hidden_heads()
        self.num_kv = policy.num_kv_heads()
        self.rotary_dim, self.rope_theta = policy.rotary()
        self.attn: Optional[AttentionTensors] = None
        self.mlp_t: Optional[MLPTensors] = None
        self.norms: Optional[NormTensors] = None
        self.module = None
        self.ds_model_config = None

    # -------------------------------------------------------------------------------------- config
    def create_ds_model_config(self):
        from ...ops.transformer.inference.config import DeepSpeedInferenceConfig
        p = self.policy
        self.ds_model_config = DeepSpeedInferenceConfig(
            hidden_size=self.hidden_size, intermediate_size=self.intermediate_size, heads=self.num_attention_heads,
            layer_norm_eps=self.layernorm_epsilon, dtype=self.dtype, pre_layer_norm=p.pre_attn_norm,
            norm_type=p.norm_type, mp_size=self.mp_size, scale_attention=p.scale_attention,
            triangular_masking=p.triangular_masking, local_attention=p.local_attention, window_size=p.window_size,
            rotary_dim=self.rotary_dim, rotate_half=p.rotate_half, rotate_every_two=p.rotate_every_two,
            return_tuple=False, mlp_after_attn=p.mlp_after_attn, mlp_act_func_type=p.mlp_act_func_type,
            bigscience_bloom=p.bigscience_bloom, use_mup=p.use_mup,
            max_out_tokens=getattr(self.config, "max_out_tokens", 1024),
            min_out_tokens=getattr(self.config, "min_out_tokens", 1), num_kv=self.num_kv, rope_theta=self.rope_theta,
            return_single_tuple=False, set_empty_params=False, transposed_mode=False, invert_mask=False)
        return self.ds_model_config

    # -------------------------------------------------------------------------------------- tensors
    def initialize_tensors(self):
        self.attn, self.mlp_t, self.norms = self.policy.attention(), self.policy.mlp(), self.policy.layernorm()
        a = self.attn
        if a.qkvw is None:          # split projections: fuse in q|k|v order
            a.qkvw = torch.cat([a.qw, a.kw, a.vw], dim=0)
            if a.qb is not None or a.kb is not None or a.vb is not None:
                z = lambda w, b: b if b is not None else torch.zeros(w.shape[0], dtype=w.dtype, device=w.device)  # noqa: E731
                a.qkvb = torch.cat([z(a.qw, a.qb), z(a.kw, a.kb), z(a.vw, a.vb)], dim=0)
        m = self.mlp_t
        if m.inter_w is None:       # gated: fused layer expects (up | gate)
            m.inter_w = torch.cat([m.up_w, m.gate_w], dim=0)

    def apply_tensor_parallelism(self):
        """Column-parallel QKV / MLP-in (each of q, k, v and up, gate cut separately), row-parallel attention-out /
        MLP-out (reference ``base.py:241-281`` + ``features/split_qkv.py``)."""
        if self.mp_size == 1:
            return
        r, n = self.mp_rank, self.mp_size
        a, m = self.attn, self.mlp_t
        d = self.hidden_size // self.num_attention_heads
        hq = self.num_attention_heads
        hkv = self.num_kv if self.num_kv > 0 else hq
        if hq % n or hkv % n:
            raise ValueError(f"heads ({hq} q / {hkv} kv) must divide by the tensor-parallel size {n}")

        def cut_rows(t, sizes):
            if t is None:
                return None
            parts, o = [], 0
            for s in sizes:
                parts.append(t[o + r * (s // n): o + (r + 1) * (s // n)])
                o += s
            return torch.cat(parts, dim=0).contiguous()

        a.qkvw = cut_rows(a.qkvw, [hq * d, hkv * d, hkv * d])
        a.qkvb = cut_rows(a.qkvb, [hq * d, hkv * d, hkv * d])
        a.ow = a.ow[:, r * (hq * d // n):(r + 1) * (hq * d // n)].contiguous()
        gated = m.up_w is not None
        inter = self.intermediate_size
        m.inter_w = cut_rows(m.inter_w, [inter, inter] if gated else [inter])
        m.inter_b = cut_rows(m.inter_b, [inter, inter] if gated else [inter])
        m.out_w = m.out_w[:, r * (inter // n):(r + 1) * (inter // n)].contiguous()
        # row-parallel biases (attention output, MLP output) stay whole on every rank: the fused layer adds them AFTER
        # the all-reduce of the partial products

    # -------------------------------------------------------------------------------------- module
    def create_module(self, config=None):
        from ...model_implementations.transformers.ds_transformer import DeepSpeedTransformerInference
        cfg = config or self.ds_model_config or self.create_ds_model_config()
        self.module = DeepSpeedTransformerInference(cfg, mp_group=self.mp_group)
        self.copy_data_to_new_module()
        return self.module

    def copy_data_to_new_module(self):
        mod, a, m, nm = self.module, self.attn, self.mlp_t, self.norms
        dt = torch.half if self.dtype == torch.int8 else self.dtype

        def put(param, value, transpose=False):
            if value is None:
                param.data.zero_()
                return
            v = value.t() if transpose else value
            assert tuple(param.shape) == tuple(v.shape), (tuple(param.shape), tuple(v.shape))
            param.data.copy_(v.to(dt))

        quantizer = None
        if self.dtype == torch.int8:
            from ..replace_module import GroupQuantizer
            q = getattr(self.config, "quant", None)
            wq = getattr(q, "weight", None)
            n_groups = int(getattr(wq, "q_groups", 1) or 1)
            # default (q_groups = 1 in the config): one step size per row of the stored [in, out] weight, as the
            # reference's GroupQuantizer() does; q_groups > 1 asks for exactly that many groups
            quantizer = GroupQuantizer(q_int8=True, num_bits=int(getattr(wq, "num_bits", 8) or 8), group_size=1,
                                       num_groups=n_groups if n_groups > 1 else 0)

        def put_gemm(owner, name, value):
            """The four GEMM weights: kept in the compute dtype, or as int8 + per-group step sizes."""
            param = getattr(owner, name)
            if quantizer is None:
                return put(param, value, True)
            assert tuple(param.shape) == tuple(value.t().shape), (tuple(param.shape), tuple(value.t().shape))
            setattr(owner, name, quantizer.quantize(value.t().contiguous().to(param.device)))

        put_gemm(mod.attention, "attn_qkvw", a.qkvw)
        put(mod.attention.attn_qkvb, a.qkvb)
        put_gemm(mod.attention, "attn_ow", a.ow)
        put(mod.attention.attn_ob, a.ob)
        put_gemm(mod.mlp, "inter_w", m.inter_w)
        put(mod.mlp.inter_b, m.inter_b)
        put_gemm(mod.mlp, "output_w", m.out_w)
        put(mod.mlp.output_b, m.out_b)
        if nm.attn_nw is None:       # single-norm parallel block (GPT-J): MLP reads the attention's normed input
            mod.mlp.attn_nw = mod.mlp.attn_nb = None
        else:
            put(mod.mlp.attn_nw, nm.attn_nw)
            put(mod.mlp.attn_nb, nm.attn_nb)
        put(mod.norm_w, nm.input_nw)
        put(mod.norm_b, nm.input_nb)
