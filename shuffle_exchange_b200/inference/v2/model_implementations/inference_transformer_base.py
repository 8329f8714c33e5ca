"""Generic ragged-batch transformer for the v2 engine, driven by an architecture description.

Reference: ``inference/v2/model_implementations/{inference_model_base,inference_transformer_base}.py`` define an
abstract model whose subclasses (one per architecture) wire modules picked by ``modules/heuristics.py``; every
architecture then repeats the same forward with small variations.  Here those variations are data
(:class:`ArchSpec`): norm type, sequential vs parallel attention/MLP, gated vs plain MLP and its activation, rotary
(full / partial) vs learned positions, which projections carry biases, dense vs top-k MoE (optional shared expert),
tied or separate unembedding.  One forward serves Llama-2/3, Mistral, Mixtral, Falcon, OPT, Phi, Phi-3, Qwen, Qwen2 and
Qwen2-MoE; per-architecture packages only map HuggingFace parameter names onto the fused layout used here
(``wqkv [ (hq + 2 hkv) * d, H ]``, ``w_gate_up [2F, H]`` = gate | up halves).
"""
from __future__ import annotations

from abc import ABC, abstractmethod
from dataclasses import dataclass, field
from typing import Dict, List, Optional

import torch

from ....ops import functional as F
from ..config_v2 import RaggedInferenceEngineConfig
from ..inference_utils import ActivationType, DtypeEnum, NormTypeEnum, ceil_div
from ..modules import heuristics
from ..modules.configs import (DSEmbeddingsConfig, DSLinearConfig, DSMoEConfig, DSNormConfig, DSSelfAttentionConfig,
                               DSUnembedConfig, MaskingType, PositionalEmbeddingType, RotateHalfConfig)
from ..ragged.kv_cache import KVCacheConfig
from . import sharding


@dataclass
class ArchSpec:
    num_layers: int
    model_dim: int
    vocab_size: int
    n_heads: int
    n_heads_kv: int
    head_size: int
    intermediate_dim: int
    norm_type: NormTypeEnum = NormTypeEnum.RMSNorm
    norm_eps: float = 1e-5
    mlp_activation: ActivationType = ActivationType.SiGLU
    positional: PositionalEmbeddingType = PositionalEmbeddingType.rotate_half
    rope_theta: float = 10000.0
    rotary_dim: Optional[int] = None            # partial rotary (Phi); None = whole head
    learned_positions: bool = False             # OPT
    positional_offset: int = 0
    parallel_attn_mlp: bool = False             # Falcon, Phi: attention and MLP read the same normed input
    separate_mlp_norm: bool = False             # Falcon new decoder architecture: ln_attn + ln_mlp
    n_experts: int = 0
    top_k: int = 0
    normalize_expert_scores: bool = True
    shared_expert_dim: int = 0                  # Qwen2-MoE
    moe_intermediate_dim: int = 0
    dtype: torch.dtype = torch.bfloat16
    max_seq_len: int = 4096
    attn_scale: Optional[float] = None


class DSInferenceModelBase(torch.nn.Module, ABC):
    """Contract the engine relies on (reference ``inference_model_base.py:56-270``)."""

    @abstractmethod
    def kv_cache_config(self, block_size: int) -> KVCacheConfig:
        ...

    @abstractmethod
    def get_kv_requirements(self, sequence, max_new_tokens: int, max_new_blocks: int):
        ...

    @abstractmethod
    def maybe_allocate_kv(self, sequence, n_new_tokens: int) -> None:
        ...

    @abstractmethod
    def forward(self, wrapped_batch) -> torch.Tensor:
        ...


class LayerWeights(dict):
    """Named tensors of one transformer layer (attribute access for readability)."""

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError:
            return None


class DSTransformerModelBase(DSInferenceModelBase):
    """Concrete generic model.  ``weights``: ``{"embed", "pos_embed"?, "final_norm_w", "final_norm_b"?, "unembed",
    "unembed_b"?, "layers": [LayerWeights, ...]}``; each layer has ``ln1_w/ln1_b``, ``wqkv/bqkv``, ``wo/bo``,
    ``ln2_w/ln2_b`` (absent when ``parallel_attn_mlp`` without a separate MLP norm), and either ``w_in/b_in`` +
    ``w_out/b_out`` (dense MLP; ``w_in`` is gate|up for gated activations) or ``moe_gate``, ``moe_w1 [E, .., H]``,
    ``moe_w2 [E, H, F]`` (+ ``shared_w_in``, ``shared_w_out``, ``shared_gate``)."""

    def __init__(self, spec: ArchSpec, weights: Dict, engine_config: RaggedInferenceEngineConfig, state_manager=None,
                 base_mp_group=None) -> None:
        super().__init__()
        self.spec = spec
        self._engine_config = engine_config or RaggedInferenceEngineConfig()
        self.state_manager = state_manager
        self._base_mp_group = base_mp_group
        self._tp_size, self._tp_rank = 1, 0
        if base_mp_group is not None:
            import torch.distributed as dist
            self._tp_size, self._tp_rank = dist.get_world_size(base_mp_group), dist.get_rank(base_mp_group)
        self.flattened_params = self.flattened_param_metadata = None
        if weights.get("_flat") is not None:
            self.flattened_params, self.flattened_param_metadata = weights["_flat"]
        if weights.get("_presharded") is not None:
            if tuple(weights["_presharded"]) != (self._tp_rank, self._tp_size):
                raise ValueError(f"checkpoint was serialized for TP rank/size {weights['_presharded']}, "
                                 f"engine runs {(self._tp_rank, self._tp_size)}")
        elif self._tp_size > 1:
            # tensor parallelism (reference ``inference_transformer_base.py:199-560``): QKV / MLP-in column-parallel,
            # attention-out / MLP-out row-parallel with one all-reduce each, vocabulary-parallel unembedding
            cache = weights.setdefault("_tp_local", {})
            key = (self._tp_rank, self._tp_size)
            if key not in cache:
                cache[key] = sharding.shard_weights(weights, spec, self._tp_rank, self._tp_size)
            weights = cache[key]
        self.weights = weights
        self._build_modules()

    # ------------------------------------------------------------------------------------------ properties
    @property
    def num_layers(self): return self.spec.num_layers
    @property
    def model_dim(self): return self.spec.model_dim
    @property
    def vocab_size(self): return self.spec.vocab_size
    @property
    def head_size(self): return self.spec.head_size
    @property
    def n_heads(self): return self.spec.n_heads
    @property
    def n_heads_q_local(self):
        s = self.spec
        return sharding.get_local_heads(self._tp_rank, self._tp_size, s.n_heads,
                                        None if s.n_heads_kv == s.n_heads else s.n_heads_kv)[0]
    @property
    def n_heads_kv(self): return self.spec.n_heads_kv
    @property
    def n_heads_kv_local(self):
        s = self.spec
        return sharding.get_local_heads(self._tp_rank, self._tp_size, s.n_heads,
                                        None if s.n_heads_kv == s.n_heads else s.n_heads_kv)[1]
    @property
    def intermediate_dim(self): return self.spec.intermediate_dim
    @property
    def activation_dtype(self): return DtypeEnum(self.spec.dtype)
    @property
    def mlp_activation_fn(self): return self.spec.mlp_activation
    @property
    def norm_type(self): return self.spec.norm_type
    @property
    def positional_embedding_type(self): return self.spec.positional
    @property
    def max_sequence_length(self): return self.spec.max_seq_len
    @property
    def tp_size(self): return self._tp_size
    @property
    def tp_rank(self): return self._tp_rank

    def _local_dim(self, n: int) -> int:
        return sharding.sharded_intermediate_dim(n, self._tp_size, self._tp_rank) if self._tp_size > 1 else n

    def _reduce(self, x: torch.Tensor) -> torch.Tensor:
        """Sum the partial outputs of a row-parallel projection over the tensor-parallel group."""
        if self._tp_size > 1:
            import torch.distributed as dist
            dist.all_reduce(x, group=self._base_mp_group)
        return x

    def _gather_logits(self, local: torch.Tensor) -> torch.Tensor:
        """Concatenate vocabulary shards (shards may differ by one row: pad to the largest, trim after)."""
        if self._tp_size == 1:
            return local
        import torch.distributed as dist
        sizes = [sharding.sharded_unembed_dim(self.spec.vocab_size, r, self._tp_size) for r in range(self._tp_size)]
        width = max(sizes)
        send = local if local.shape[-1] == width else torch.nn.functional.pad(local, (0, width - local.shape[-1]))
        parts = [torch.empty_like(send) for _ in sizes]
        dist.all_gather(parts, send.contiguous(), group=self._base_mp_group)
        return torch.cat([p[..., :n] for p, n in zip(parts, sizes)], dim=-1)

    # ------------------------------------------------------------------------------------------ module construction
    def _build_modules(self) -> None:
        s, ec = self.spec, self._engine_config
        sm = ec.state_manager
        dt = DtypeEnum(s.dtype)
        mt = sm.max_ragged_batch_size
        rot = None
        if s.positional == PositionalEmbeddingType.rotate_half:
            rot = RotateHalfConfig(theta_base=s.rope_theta, rotate_dim=s.rotary_dim)
        hq, hkv = self.n_heads_q_local, self.n_heads_kv_local
        self.attn = heuristics.instantiate_attention(DSSelfAttentionConfig(
            max_tokens=mt, n_heads_q=hq, n_heads_kv=hkv, head_size=s.head_size,
            max_sequences=sm.max_ragged_sequence_count,
            scale_factor=s.attn_scale if s.attn_scale is not None else s.head_size ** -0.5, input_dtype=dt,
            output_dtype=dt, masking_type=MaskingType.causal, positional_embedding_type=s.positional,
            positional_embedding_config=rot), ec)
        self.embed = heuristics.instantiate_embed(DSEmbeddingsConfig(
            max_tokens=mt, residual_dtype=dt, embedding_dim=s.model_dim, positional_embedding=s.learned_positions,
            positional_offset=s.positional_offset), ec)
        lin = lambda i, o, act=ActivationType.IDENTITY: heuristics.instantiate_linear(   # noqa: E731
            DSLinearConfig(max_tokens=mt, in_channels=i, out_channels=o, activation=act, input_dtype=dt,
                           output_dtype=dt), ec)
        qkv_out = (hq + 2 * hkv) * s.head_size
        self.qkv = lin(s.model_dim, qkv_out)
        self.attn_out = lin(hq * s.head_size, s.model_dim)
        if s.n_experts > 0:
            self.moe = heuristics.instantiate_moe(DSMoEConfig(
                max_tokens=mt, model_dim=s.model_dim,
                intermediate_features=self._local_dim(s.moe_intermediate_dim or s.intermediate_dim),
                n_experts=s.n_experts, top_k=s.top_k, input_dtype=dt, output_dtype=dt, activation=s.mlp_activation,
                normalize_scores=s.normalize_expert_scores), ec)
            if s.shared_expert_dim > 0:
                self.shared_in = lin(s.model_dim, self._local_dim(s.shared_expert_dim), s.mlp_activation)
                self.shared_out = lin(self._local_dim(s.shared_expert_dim), s.model_dim)
        else:
            self.mlp_in = lin(s.model_dim, self._local_dim(s.intermediate_dim), s.mlp_activation)
            self.mlp_out = lin(self._local_dim(s.intermediate_dim), s.model_dim)
        self.norm = heuristics.instantiate_pre_norm(DSNormConfig(
            max_tokens=mt, type=s.norm_type, channels=s.model_dim, residual_dtype=dt, input_dtype=dt, output_dtype=dt,
            eps=s.norm_eps), ec)
        self.unembed = heuristics.instantiate_unembed(DSUnembedConfig(
            max_tokens=mt, dtype=dt, norm_type=s.norm_type, model_dim=s.model_dim,
            max_sequences=sm.max_ragged_sequence_count,
            vocab_size=sharding.sharded_unembed_dim(s.vocab_size, self._tp_rank, self._tp_size), eps=s.norm_eps), ec)
        # let the linear implementation pick its storage (e.g. FP6) for every 2-D projection weight (once: the engine
        # builds the model twice — a sizing probe and the real one — over the same weight dict)
        if self.weights.get("_transformed"):
            return
        self.weights["_transformed"] = True
        for lw in self.weights["layers"]:
            for k, mod in (("wqkv", self.qkv), ("wo", self.attn_out), ("w_in", getattr(self, "mlp_in", None)),
                           ("w_out", getattr(self, "mlp_out", None))):
                if mod is not None and lw.get(k) is not None:
                    lw[k] = mod.transform_param(lw[k])

    # ------------------------------------------------------------------------------------------ KV planning
    def kv_cache_config(self, block_size: int) -> KVCacheConfig:
        s = self.spec
        return KVCacheConfig(block_size=block_size, cache_shape=(s.num_layers, self.n_heads_kv_local, s.head_size),
                             cache_dtype=s.dtype)

    def get_kv_requirements(self, sequence, max_new_tokens: int, max_new_blocks: int):
        bs = self.state_manager.kv_block_size
        total = sequence.seen_tokens + max_new_tokens
        need = ceil_div(total, bs) - sequence.cur_allocated_blocks
        if need > max_new_blocks:
            fit = (sequence.cur_allocated_blocks + max_new_blocks) * bs - sequence.seen_tokens
            return max(0, fit), max_new_blocks
        return max_new_tokens, max(0, need)

    def get_remaining_block_capacity(self, sequence) -> int:
        bs = self.state_manager.kv_block_size
        return sequence.cur_allocated_blocks * bs - sequence.seen_tokens

    def maybe_allocate_kv(self, sequence, n_new_tokens: int) -> None:
        _, need = self.get_kv_requirements(sequence, n_new_tokens, self.state_manager.free_blocks)
        if need > 0:
            sequence.extend_kv_cache(self.state_manager.allocate_blocks(need))

    def maybe_free_kv(self, sequence) -> None:
        """Sliding-window models could return blocks here; dense causal attention keeps everything."""

    # ------------------------------------------------------------------------------------------ forward
    def _mlp(self, lw: LayerWeights, x: torch.Tensor) -> torch.Tensor:
        if self.spec.n_experts > 0:
            y = self.moe(x, lw.moe_gate, lw.moe_w1, lw.moe_w2)
            if self.spec.shared_expert_dim > 0:
                sh = self.shared_out(self.shared_in(x, lw.shared_w_in), lw.shared_w_out)
                if lw.shared_gate is not None:
                    sh = sh * torch.sigmoid(torch.nn.functional.linear(x, lw.shared_gate))
                y = y + sh
            return self._reduce(y)
        return self._reduce(self.mlp_out(self.mlp_in(x, lw.w_in, lw.b_in), lw.w_out, lw.b_out))

    def _attention(self, li: int, lw: LayerWeights, x: torch.Tensor, batch) -> torch.Tensor:
        qkv = self.qkv(x, lw.wqkv, lw.bqkv)
        ctx = self.attn(qkv, self.state_manager.get_cache(li), batch)
        return self._reduce(self.attn_out(ctx, lw.wo, lw.bo))

    @torch.no_grad()
    def forward(self, batch) -> torch.Tensor:
        W, s = self.weights, self.spec
        residual = self.embed(batch, W["embed"], W.get("pos_embed"))
        hidden = None
        for li, lw in enumerate(W["layers"]):
            residual, x = self.norm(residual, hidden, lw.ln1_w, lw.ln1_b)
            if s.parallel_attn_mlp:
                xm = x
                if s.separate_mlp_norm:
                    xm = self.norm(residual, None, lw.ln2_w, lw.ln2_b)[1]
                hidden = self._attention(li, lw, x, batch) + self._mlp(lw, xm)
            else:
                attn = self._attention(li, lw, x, batch)
                residual, x = self.norm(residual, attn, lw.ln2_w, lw.ln2_b)
                hidden = self._mlp(lw, x)
        residual = residual + hidden if hidden is not None else residual
        return self._gather_logits(self.unembed(residual, W["unembed"], batch, W.get("unembed_b"), W["final_norm_w"],
                                                W.get("final_norm_b")))
