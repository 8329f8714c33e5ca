"""Per-architecture policies (reference ``inference/v2/model_implementations/{llama_v2,mistral,mixtral,falcon,opt,phi,
phi3,qwen,qwen_v2,qwen_v2_moe}/{policy,container,model}.py``).  Each class states the :class:`ArchSpec` derived from
the HuggingFace config and the name mapping onto the fused layout of :class:`DSTransformerModelBase`."""
from __future__ import annotations

import torch

from ..inference_utils import ActivationType, NormTypeEnum
from ..modules.configs import PositionalEmbeddingType
from .inference_policy_base import InferenceV2Policy
from .inference_transformer_base import ArchSpec


def _dtype(cfg):
    dt = getattr(cfg, "torch_dtype", None) or getattr(cfg, "dtype", None)
    if isinstance(dt, str):
        dt = getattr(torch, dt.replace("torch.", ""), None)
    return dt if dt in (torch.bfloat16, torch.float16, torch.float32) else torch.bfloat16


def _rope(cfg, key, default):
    """transformers >= 5 keeps rotary settings in ``config.rope_parameters``; older configs have flat attributes."""
    v = getattr(cfg, key, None)
    if v is None:
        v = (getattr(cfg, "rope_parameters", None) or {}).get(key)
    return default if v is None else v


def _base(cfg, **kw) -> ArchSpec:
    heads = cfg.num_attention_heads
    d = dict(num_layers=cfg.num_hidden_layers, model_dim=cfg.hidden_size, vocab_size=cfg.vocab_size, n_heads=heads,
             n_heads_kv=getattr(cfg, "num_key_value_heads", None) or heads,
             head_size=getattr(cfg, "head_dim", None) or cfg.hidden_size // heads,
             intermediate_dim=getattr(cfg, "intermediate_size", 4 * cfg.hidden_size),
             norm_eps=getattr(cfg, "rms_norm_eps", getattr(cfg, "layer_norm_eps", 1e-5)),
             rope_theta=float(_rope(cfg, "rope_theta", 10000.0)), dtype=_dtype(cfg),
             max_seq_len=getattr(cfg, "max_position_embeddings", 4096))
    d.update(kw)
    return ArchSpec(**d)


_LLAMA_LAYER = {
    "ln1_w": "input_layernorm.weight",
    "wqkv": ("self_attn.q_proj.weight", "self_attn.k_proj.weight", "self_attn.v_proj.weight"),
    "bqkv": ("self_attn.q_proj.bias", "self_attn.k_proj.bias", "self_attn.v_proj.bias"),
    "wo": "self_attn.o_proj.weight",
    "ln2_w": "post_attention_layernorm.weight",
    "w_in": ("mlp.gate_proj.weight", "mlp.up_proj.weight"),
    "w_out": "mlp.down_proj.weight",
}
_LLAMA_GLOBAL = {"embed": "model.embed_tokens.weight", "final_norm_w": "model.norm.weight",
                 "unembed": "lm_head.weight"}


class Llama2Policy(InferenceV2Policy):
    layer_rules, global_rules = _LLAMA_LAYER, _LLAMA_GLOBAL

    def spec(self, cfg) -> ArchSpec:
        return _base(cfg)


class MistralPolicy(Llama2Policy):
    pass


class Qwen2Policy(Llama2Policy):
    """Llama layout + biases on q/k/v (picked up by the ``bqkv`` rule)."""


class Phi3Policy(InferenceV2Policy):
    layer_rules = {"ln1_w": "input_layernorm.weight", "wqkv": "self_attn.qkv_proj.weight",
                   "wo": "self_attn.o_proj.weight", "ln2_w": "post_attention_layernorm.weight",
                   "w_in": "mlp.gate_up_proj.weight", "w_out": "mlp.down_proj.weight"}
    global_rules = _LLAMA_GLOBAL

    def spec(self, cfg) -> ArchSpec:
        return _base(cfg)


def _stack_experts(raw, fmt_in, fmt_out, n):
    """-> (w1 [E, 2F, H] gate|up, w2 [E, H, F]) and the consumed names.  Accepts per-expert tensors (checkpoint files)
    or the pre-fused ``mlp.experts.gate_up_proj`` / ``mlp.experts.down_proj`` of in-memory transformers >= 5 models."""
    if "mlp.experts.gate_up_proj" in raw:
        return (raw["mlp.experts.gate_up_proj"].contiguous(), raw["mlp.experts.down_proj"].contiguous(),
                ["mlp.experts.gate_up_proj", "mlp.experts.down_proj"])
    w1, w2, used = [], [], []
    for e in range(n):
        names = [f.format(e=e) for f in fmt_in]
        w1.append(torch.cat([raw[x] for x in names], dim=0))
        w2.append(raw[fmt_out.format(e=e)])
        used += names + [fmt_out.format(e=e)]
    return torch.stack(w1).contiguous(), torch.stack(w2).contiguous(), used


class MixtralPolicy(InferenceV2Policy):
    layer_rules = {k: v for k, v in _LLAMA_LAYER.items() if k not in ("w_in", "w_out")}
    layer_rules["moe_gate"] = "block_sparse_moe.gate.weight"
    global_rules = _LLAMA_GLOBAL

    def spec(self, cfg) -> ArchSpec:
        return _base(cfg, n_experts=cfg.num_local_experts, top_k=cfg.num_experts_per_tok,
                     normalize_expert_scores=True)

    def transform_layer(self, lw, raw, spec):
        lw["moe_w1"], lw["moe_w2"], used = _stack_experts(
            raw, ("block_sparse_moe.experts.{e}.w1.weight", "block_sparse_moe.experts.{e}.w3.weight"),
            "block_sparse_moe.experts.{e}.w2.weight", spec.n_experts)
        if "moe_gate" not in lw and "mlp.gate.weight" in raw:       # transformers >= 5 module naming
            lw["moe_gate"] = raw["mlp.gate.weight"]
            used.append("mlp.gate.weight")
        self._extra_used_layer = used


class Qwen2MoePolicy(InferenceV2Policy):
    layer_rules = {k: v for k, v in _LLAMA_LAYER.items() if k not in ("w_in", "w_out")}
    layer_rules.update({"moe_gate": "mlp.gate.weight",
                        "shared_w_in": ("mlp.shared_expert.gate_proj.weight", "mlp.shared_expert.up_proj.weight"),
                        "shared_w_out": "mlp.shared_expert.down_proj.weight",
                        "shared_gate": "mlp.shared_expert_gate.weight"})
    global_rules = _LLAMA_GLOBAL

    def spec(self, cfg) -> ArchSpec:
        return _base(cfg, n_experts=cfg.num_experts, top_k=cfg.num_experts_per_tok,
                     normalize_expert_scores=bool(getattr(cfg, "norm_topk_prob", False)),
                     moe_intermediate_dim=cfg.moe_intermediate_size,
                     shared_expert_dim=cfg.shared_expert_intermediate_size)

    def transform_layer(self, lw, raw, spec):
        lw["moe_w1"], lw["moe_w2"], used = _stack_experts(
            raw, ("mlp.experts.{e}.gate_proj.weight", "mlp.experts.{e}.up_proj.weight"),
            "mlp.experts.{e}.down_proj.weight", spec.n_experts)
        self._extra_used_layer = used


class QwenPolicy(InferenceV2Policy):
    """Qwen (v1): fused ``c_attn`` with bias; MLP ``c_proj(w1(x) * silu(w2(x)))`` -> gate = w2, up = w1."""
    layer_prefix = "transformer.h."
    layer_rules = {"ln1_w": "ln_1.weight", "wqkv": "attn.c_attn.weight", "bqkv": "attn.c_attn.bias",
                   "wo": "attn.c_proj.weight", "ln2_w": "ln_2.weight", "w_in": ("mlp.w2.weight", "mlp.w1.weight"),
                   "w_out": "mlp.c_proj.weight"}
    global_rules = {"embed": "transformer.wte.weight", "final_norm_w": "transformer.ln_f.weight",
                    "unembed": "lm_head.weight"}

    def spec(self, cfg) -> ArchSpec:
        heads = cfg.num_attention_heads
        return ArchSpec(num_layers=cfg.num_hidden_layers, model_dim=cfg.hidden_size, vocab_size=cfg.vocab_size,
                        n_heads=heads, n_heads_kv=heads, head_size=getattr(cfg, "kv_channels", cfg.hidden_size // heads),
                        intermediate_dim=cfg.intermediate_size // 2, norm_eps=getattr(cfg, "layer_norm_epsilon", 1e-6),
                        rope_theta=float(getattr(cfg, "rotary_emb_base", 10000.0)), dtype=_dtype(cfg),
                        max_seq_len=getattr(cfg, "seq_length", 8192))


class OPTPolicy(InferenceV2Policy):
    """OPT: LayerNorm with bias, learned positions stored with an offset of 2, ReLU MLP, biases everywhere, tied
    unembedding.  Only ``do_layer_norm_before=True`` checkpoints (everything except 350m)."""
    layer_prefix = "model.decoder.layers."
    layer_rules = {
        "ln1_w": "self_attn_layer_norm.weight", "ln1_b": "self_attn_layer_norm.bias",
        "wqkv": ("self_attn.q_proj.weight", "self_attn.k_proj.weight", "self_attn.v_proj.weight"),
        "bqkv": ("self_attn.q_proj.bias", "self_attn.k_proj.bias", "self_attn.v_proj.bias"),
        "wo": "self_attn.out_proj.weight", "bo": "self_attn.out_proj.bias",
        "ln2_w": "final_layer_norm.weight", "ln2_b": "final_layer_norm.bias",
        "w_in": "fc1.weight", "b_in": "fc1.bias", "w_out": "fc2.weight", "b_out": "fc2.bias"}
    global_rules = {"embed": "model.decoder.embed_tokens.weight", "pos_embed": "model.decoder.embed_positions.weight",
                    "final_norm_w": "model.decoder.final_layer_norm.weight",
                    "final_norm_b": "model.decoder.final_layer_norm.bias", "unembed": "lm_head.weight"}

    def spec(self, cfg) -> ArchSpec:
        if not getattr(cfg, "do_layer_norm_before", True):
            raise ValueError("OPT with do_layer_norm_before=False (the 350m model) is not supported")
        act = {"relu": ActivationType.RELU, "gelu": ActivationType.GELU}[cfg.activation_function]
        return _base(cfg, intermediate_dim=cfg.ffn_dim, norm_type=NormTypeEnum.LayerNorm, norm_eps=1e-5,
                     mlp_activation=act, positional=PositionalEmbeddingType.none, learned_positions=True,
                     positional_offset=2)


class PhiPolicy(InferenceV2Policy):
    """Phi-1.5 / Phi-2: one LayerNorm per layer feeding attention and MLP in parallel, partial rotary, GELU MLP,
    biases everywhere including the unembedding."""
    layer_rules = {
        "ln1_w": "input_layernorm.weight", "ln1_b": "input_layernorm.bias",
        "wqkv": ("self_attn.q_proj.weight", "self_attn.k_proj.weight", "self_attn.v_proj.weight"),
        "bqkv": ("self_attn.q_proj.bias", "self_attn.k_proj.bias", "self_attn.v_proj.bias"),
        "wo": "self_attn.dense.weight", "bo": "self_attn.dense.bias",
        "w_in": "mlp.fc1.weight", "b_in": "mlp.fc1.bias", "w_out": "mlp.fc2.weight", "b_out": "mlp.fc2.bias"}
    global_rules = {"embed": "model.embed_tokens.weight", "final_norm_w": "model.final_layernorm.weight",
                    "final_norm_b": "model.final_layernorm.bias", "unembed": "lm_head.weight",
                    "unembed_b": "lm_head.bias"}

    def spec(self, cfg) -> ArchSpec:
        hs = cfg.hidden_size // cfg.num_attention_heads
        rot = int(hs * float(_rope(cfg, "partial_rotary_factor", 0.5)))
        return _base(cfg, norm_type=NormTypeEnum.LayerNorm, norm_eps=cfg.layer_norm_eps,
                     mlp_activation=ActivationType.GELU, rotary_dim=rot, parallel_attn_mlp=True)


class FalconPolicy(InferenceV2Policy):
    """Falcon: fused ``query_key_value``; 7b-style (multi-query, one LayerNorm, parallel attention + MLP) and the new
    decoder architecture (grouped KV interleaved per group, separate ``ln_attn`` / ``ln_mlp``)."""
    layer_prefix = "transformer.h."
    layer_rules = {"wo": "self_attention.dense.weight", "bo": "self_attention.dense.bias",
                   "w_in": "mlp.dense_h_to_4h.weight", "b_in": "mlp.dense_h_to_4h.bias",
                   "w_out": "mlp.dense_4h_to_h.weight", "b_out": "mlp.dense_4h_to_h.bias"}
    global_rules = {"embed": "transformer.word_embeddings.weight", "final_norm_w": "transformer.ln_f.weight",
                    "final_norm_b": "transformer.ln_f.bias", "unembed": "lm_head.weight"}

    def _kv_heads(self, cfg) -> int:
        if getattr(cfg, "new_decoder_architecture", False):
            return cfg.num_kv_heads
        return 1 if getattr(cfg, "multi_query", True) else cfg.num_attention_heads

    def spec(self, cfg) -> ArchSpec:
        if getattr(cfg, "alibi", False):
            raise ValueError("Falcon checkpoints with ALiBi are not supported")
        if not getattr(cfg, "parallel_attn", True):
            raise ValueError("Falcon checkpoints with parallel_attn=False are not supported")
        heads = cfg.num_attention_heads
        return ArchSpec(num_layers=cfg.num_hidden_layers, model_dim=cfg.hidden_size, vocab_size=cfg.vocab_size,
                        n_heads=heads, n_heads_kv=self._kv_heads(cfg), head_size=cfg.hidden_size // heads,
                        intermediate_dim=4 * cfg.hidden_size, norm_type=NormTypeEnum.LayerNorm,
                        norm_eps=cfg.layer_norm_epsilon, mlp_activation=ActivationType.GELU,
                        rope_theta=float(_rope(cfg, "rope_theta", 10000.0)), parallel_attn_mlp=True,
                        separate_mlp_norm=bool(getattr(cfg, "new_decoder_architecture", False)), dtype=_dtype(cfg),
                        max_seq_len=getattr(cfg, "max_position_embeddings", 2048))

    def transform_layer(self, lw, raw, spec):
        used = ["self_attention.query_key_value.weight"]
        w = raw["self_attention.query_key_value.weight"]
        nq, nkv, d = spec.n_heads, spec.n_heads_kv, spec.head_size
        if spec.separate_mlp_norm:
            # [n_kv groups, (q per group + k + v), d, H] -> all q | all k | all v
            g = nq // nkv
            w4 = w.view(nkv, g + 2, d, w.shape[-1])
            w = torch.cat([w4[:, :g].reshape(nq * d, -1), w4[:, g].reshape(nkv * d, -1),
                           w4[:, g + 1].reshape(nkv * d, -1)], dim=0)
            lw["ln1_w"], lw["ln1_b"] = raw["ln_attn.weight"], raw.get("ln_attn.bias")
            lw["ln2_w"], lw["ln2_b"] = raw["ln_mlp.weight"], raw.get("ln_mlp.bias")
            used += ["ln_attn.weight", "ln_attn.bias", "ln_mlp.weight", "ln_mlp.bias"]
        else:
            if nkv == nq:      # multi-head, interleaved per head: [heads, 3, d, H]
                w3 = w.view(nq, 3, d, w.shape[-1])
                w = torch.cat([w3[:, i].reshape(nq * d, -1) for i in range(3)], dim=0)
            lw["ln1_w"], lw["ln1_b"] = raw["input_layernorm.weight"], raw.get("input_layernorm.bias")
            used += ["input_layernorm.weight", "input_layernorm.bias"]
        lw["wqkv"] = w.contiguous()
        self._extra_used_layer = used


# The same architectures assembled through the declarative containers (``containers.py``); selectable with
# ``engine_factory`` keyword ``use_containers=True`` or by instantiating these classes directly.
from .containers import (ContainerPolicyMixin, FalconNewArchTransformerContainer, FalconNonTransformerContainer,  # noqa: E402
                         Llama2NonTransformerContainer, Llama2TransformerContainer, MixtralNonTransformerContainer,
                         MixtralTransformerContainer, OPTNonTransformerContainer, OPTTransformerContainer)


class Llama2ContainerPolicy(ContainerPolicyMixin, Llama2Policy):
    transformer_container, non_transformer_container = Llama2TransformerContainer, Llama2NonTransformerContainer


class MixtralContainerPolicy(ContainerPolicyMixin, MixtralPolicy):
    transformer_container, non_transformer_container = MixtralTransformerContainer, MixtralNonTransformerContainer


class OPTContainerPolicy(ContainerPolicyMixin, OPTPolicy):
    transformer_container, non_transformer_container = OPTTransformerContainer, OPTNonTransformerContainer


class FalconNewArchContainerPolicy(ContainerPolicyMixin, FalconPolicy):
    transformer_container, non_transformer_container = FalconNewArchTransformerContainer, FalconNonTransformerContainer


CONTAINER_POLICIES = {"llama": Llama2ContainerPolicy, "mistral": Llama2ContainerPolicy, "qwen2": Llama2ContainerPolicy,
                      "mixtral": MixtralContainerPolicy, "opt": OPTContainerPolicy}

POLICIES = {"llama": Llama2Policy, "mistral": MistralPolicy, "mixtral": MixtralPolicy, "falcon": FalconPolicy,
            "opt": OPTPolicy, "phi": PhiPolicy, "phi3": Phi3Policy, "qwen": QwenPolicy, "qwen2": Qwen2Policy,
            "qwen2_moe": Qwen2MoePolicy}
