"""Concrete module implementations on the framework's sm_100a kernels (reference
``inference/v2/modules/implementations/{attention,embedding,linear,moe,pre_norm,post_norm,unembed}/*``).

=========================  =======================================================================================
 ``dense_blocked_attention``  ``sxb_kv_rotary_blocked`` (rotary + paged-cache append) + ``sxb_paged_attention``
 ``ragged_embedding``         ``sxb_ragged_embed`` (token + optional positional table with offset)
 ``blas_fp_linear``           tcgen05 GEMM + fused bias/activation or gated activation
 ``quantized_wf6af16_linear`` weights stored FP6 (group-scaled), dequantised to bf16 per call, tcgen05 GEMM
 ``cutlass_multi_gemm_moe``   top-k gating / scatter / per-expert GEMMs over contiguous row ranges / weighted gather
 ``cuda_pre_ln / cuda_pre_rms / cuda_post_ln``   fused residual-add + norm kernels
 ``ragged_unembed``           (norm +) last-token gather + vocabulary GEMM
=========================  =======================================================================================
"""
from __future__ import annotations

import math
from typing import Any, Dict, Optional, Tuple

import torch

from .....ops import functional as F
from .....ops import moe_ops
from ...inference_parameter import InferenceParameter
from ...inference_utils import ActivationType, DtypeEnum, NormTypeEnum, is_gated
from ...kernels import ragged_ops as R
from ..configs import (DSEmbeddingsConfig, DSLinearConfig, DSMoEConfig, DSNormConfig, DSSelfAttentionConfig,
                       DSUnembedConfig, MaskingType, PositionalEmbeddingType)
from ..interfaces import (DSEmbeddingBase, DSEmbeddingRegistry, DSLinearBase, DSLinearRegistry, DSMoEBase,
                          DSMoERegistry, DSPostNormBase, DSPostNormRegistry, DSPreNormBase, DSPreNormRegistry,
                          DSSelfAttentionBase, DSSelfAttentionRegistry, DSUnembedBase, DSUnembedRegistry, _param)


def _dt(x) -> DtypeEnum:
    """Config models may store enum members by value: normalise."""
    return x if isinstance(x, DtypeEnum) else DtypeEnum(x)


def _act_name(act: ActivationType) -> str:
    return {ActivationType.GELU: "gelu", ActivationType.RELU: "relu", ActivationType.SILU: "silu",
            ActivationType.GEGLU: "gelu", ActivationType.ReGLU: "relu", ActivationType.SiGLU: "silu",
            ActivationType.IDENTITY: "identity"}[ActivationType(act)]


def _apply_activation(y: torch.Tensor, bias: Optional[torch.Tensor], act: ActivationType) -> torch.Tensor:
    act = ActivationType(act)
    if is_gated(act):
        # fused weights are [gate | up] halves (HF gate_proj / up_proj order): act(gate) * up
        if bias is not None:
            y = y + bias
        return F.gated_act(y, _act_name(act))
    if act == ActivationType.IDENTITY:
        return y if bias is None else y + bias
    return F.bias_act(y, bias, _act_name(act))


# ============================================================================================ attention
@DSSelfAttentionRegistry.register_module
class DSDenseBlockedAttention(DSSelfAttentionBase):

    @staticmethod
    def name() -> str:
        return "dense_blocked_attention"

    @staticmethod
    def supports_config(config: DSSelfAttentionConfig) -> bool:
        if MaskingType(config.masking_type) != MaskingType.causal:
            return False
        return PositionalEmbeddingType(config.positional_embedding_type) in (PositionalEmbeddingType.none,
                                                                             PositionalEmbeddingType.rotate_half)

    def __init__(self, config: DSSelfAttentionConfig, implementation_config: Dict[str, Any]) -> None:
        super().__init__(config, implementation_config)
        rot = config.positional_embedding_config
        if PositionalEmbeddingType(config.positional_embedding_type) == PositionalEmbeddingType.rotate_half:
            self._rot_dim = config.head_size if rot is None or rot.rotate_dim is None else rot.rotate_dim
            self._theta = 10000.0 if rot is None else rot.theta_base
        else:
            self._rot_dim, self._theta = 0, 10000.0

    @property
    def kv_block_size(self) -> int:
        return 64

    @property
    def q_block_size(self) -> int:
        return 128

    def forward(self, q_k_v: torch.Tensor, kv_cache: torch.Tensor, batch, inv_freqs=None) -> torch.Tensor:
        c = self._config
        t2s, pos, tables = batch.token_to_seq(), batch.token_positions(), batch.block_tables()
        R.kv_rotary_blocked(q_k_v, kv_cache, t2s, pos, tables, c.n_heads_q, c.n_heads_kv, c.head_size, self._rot_dim,
                            self._theta)
        q = q_k_v.view(q_k_v.shape[0], c.n_heads_q + 2 * c.n_heads_kv, c.head_size)[:, :c.n_heads_q]
        out = R.ragged_attention(q, kv_cache, batch, scale=c.scale_factor)
        return out.reshape(out.shape[0], c.n_heads_q * c.head_size)


# ============================================================================================ embedding
@DSEmbeddingRegistry.register_module
class DSRaggedEmbedding(DSEmbeddingBase):

    @staticmethod
    def name() -> str:
        return "ragged_embedding"

    @staticmethod
    def supports_config(config: DSEmbeddingsConfig) -> bool:
        return not config.use_token_type and config.output_normalization is None

    def __init__(self, config, implementation_config):
        super().__init__(config, implementation_config)
        self._output = None

    @property
    def output(self) -> torch.Tensor:
        return self._output

    def forward(self, ragged_batch, word_embeddings, position_embeddings=None, token_type_ids=None,
                token_type_embeddings=None) -> torch.Tensor:
        ids = ragged_batch.input_ids()
        if self._config.positional_embedding:
            self._output = R.ragged_embed(ids, word_embeddings, ragged_batch.token_positions(), position_embeddings,
                                          self._config.positional_offset)
        else:
            self._output = R.ragged_embed(ids, word_embeddings)
        return self._output


# ============================================================================================ linear
@DSLinearRegistry.register_module
class BlasFPLinear(DSLinearBase):

    @staticmethod
    def name() -> str:
        return "blas_fp_linear"

    @staticmethod
    def supports_config(config: DSLinearConfig) -> bool:
        return _dt(config.input_dtype) == _dt(config.output_dtype) and _dt(config.input_dtype) in (
            DtypeEnum.fp16, DtypeEnum.bf16, DtypeEnum.fp32)

    def __init__(self, config, implementation_config):
        super().__init__(config, implementation_config)
        self._output = None

    def transform_param(self, param: torch.Tensor) -> InferenceParameter:
        return _param(param.to(_dt(self._config.input_dtype).torch))

    def forward(self, hidden_states, w, b=None) -> torch.Tensor:
        self._output = _apply_activation(F.linear(hidden_states, w), b, self._config.activation)
        return self._output

    @property
    def output(self) -> torch.Tensor:
        return self._output


@DSLinearRegistry.register_module
class QuantizedWf6Af16Linear(DSLinearBase):
    """FP6 (e3m2, group-scaled) weights, 16-bit activations — the reference's ``cuda_linear`` W6A16 path
    (``kernels/core_ops/cuda_linear``).  Weights cost 6 bits + scales in HBM; each call dequantises the layer into a
    bf16 scratch and runs the tcgen05 GEMM (decode GEMMs are weight-bandwidth bound: 2.6x less weight traffic)."""

    GROUP = 512

    @staticmethod
    def name() -> str:
        return "quantized_wf6af16_linear"

    @staticmethod
    def supports_config(config: DSLinearConfig) -> bool:
        return _dt(config.input_dtype) == _dt(config.output_dtype) and _dt(config.input_dtype) in (
            DtypeEnum.fp16, DtypeEnum.bf16) and config.in_channels % 64 == 0 and config.out_channels % 256 == 0

    def __init__(self, config, implementation_config):
        super().__init__(config, implementation_config)
        self._output = None
        from .....ops.fp_quantizer import FP_Quantize
        self._q = FP_Quantize(group_size=self.GROUP)

    def transform_param(self, param: torch.Tensor) -> InferenceParameter:
        if param.dim() == 1:
            return _param(param.to(_dt(self._config.input_dtype).torch))
        w = param.to(torch.bfloat16).contiguous()
        if w.numel() % self.GROUP != 0:
            return _param(w)
        q = self._q.quantize(w, q_bits=6, q_mantisa_bits=2, return_meta_tensor=False)
        return InferenceParameter.initialize(q, shape_marker=torch.tensor(list(w.shape)))

    def forward(self, hidden_states, w, b=None) -> torch.Tensor:
        if w.dtype == torch.uint8:
            shape = tuple(int(s) for s in w.shape_marker.tolist())
            dense = self._q.dequantize(w, fp_out=torch.empty(shape, dtype=torch.bfloat16, device=w.device), q_bits=6,
                                       q_mantisa_bits=2).view(shape)
        else:
            dense = w
        self._output = _apply_activation(F.linear(hidden_states, dense.to(hidden_states.dtype)), b,
                                         self._config.activation)
        return self._output

    @property
    def output(self) -> torch.Tensor:
        return self._output


# ============================================================================================ MoE
@DSMoERegistry.register_module
class DSMultiGemmMoE(DSMoEBase):

    @staticmethod
    def name() -> str:
        return "cutlass_multi_gemm_moe"

    @staticmethod
    def supports_config(config: DSMoEConfig) -> bool:
        return _dt(config.input_dtype) == _dt(config.output_dtype) and config.top_k in (1, 2, 4, 8)

    def __init__(self, config, implementation_config):
        super().__init__(config, implementation_config)
        self._output = None

    def transform_gate_param(self, param):
        return _param(param.to(_dt(self._config.input_dtype).torch))

    def transform_moe_mlp_1_param(self, param):
        return _param(param.to(_dt(self._config.input_dtype).torch))      # [E, (1|2) * F, H]

    def transform_moe_mlp_2_param(self, param):
        return _param(param.to(_dt(self._config.input_dtype).torch))      # [E, H, F]

    def forward(self, hidden_states, gate_w, mlp_1_w, mlp_2_w, mlp_1_b=None, mlp_2_b=None) -> torch.Tensor:
        c = self._config
        logits = torch.nn.functional.linear(hidden_states, gate_w)
        scores, assign, offsets, counts = moe_ops.top_k_gating(logits, c.top_k, renormalize=c.normalize_scores)
        from .....moe import fused_ep
        if fused_ep.grouped_inference_ok(hidden_states, mlp_1_w, mlp_2_w, mlp_1_b, mlp_2_b):
            # one grouped tcgen05 GEMM per projection over expert-contiguous rows, no host synchronisation
            self._output = fused_ep.grouped_experts_inference(
                hidden_states, scores, assign, offsets, counts, mlp_1_w, mlp_2_w,
                lambda t: _apply_activation(t, None, c.activation), c.top_k)
            return self._output
        moe_in, cumsum, slots = moe_ops.moe_scatter(hidden_states, assign, offsets, counts)
        out = torch.empty_like(moe_in)
        lo = 0
        for e, hi in enumerate(cumsum.tolist()):
            if hi > lo:
                h = _apply_activation(F.linear(moe_in[lo:hi], mlp_1_w[e]),
                                      None if mlp_1_b is None else mlp_1_b[e], c.activation)
                y = F.linear(h, mlp_2_w[e])
                out[lo:hi] = y if mlp_2_b is None else y + mlp_2_b[e]
            lo = hi
        self._output = moe_ops.moe_gather(out, scores, slots)
        return self._output

    @property
    def output(self) -> torch.Tensor:
        return self._output


# ============================================================================================ norms
def _norm(cfg: DSNormConfig, x, gamma, beta, residual=None):
    if NormTypeEnum(cfg.type) == NormTypeEnum.RMSNorm:
        return F.rms_norm(x, gamma, cfg.eps, residual)
    return F.layer_norm(x, gamma, beta, cfg.eps, residual)


@DSPreNormRegistry.register_module
class DSPreLNCUDAModule(DSPreNormBase):

    @staticmethod
    def name() -> str:
        return "cuda_pre_ln"

    @staticmethod
    def supports_config(config: DSNormConfig) -> bool:
        return NormTypeEnum(config.type) == NormTypeEnum.LayerNorm

    def forward(self, residual, hidden_states, gamma, beta=None) -> Tuple[torch.Tensor, torch.Tensor]:
        if hidden_states is None:
            return residual, _norm(self._config, residual, gamma, beta)[0]
        y, res = _norm(self._config, hidden_states, gamma, beta, residual)
        return res, y


@DSPreNormRegistry.register_module
class DSPreRMSCUDAModule(DSPreLNCUDAModule):

    @staticmethod
    def name() -> str:
        return "cuda_pre_rms"

    @staticmethod
    def supports_config(config: DSNormConfig) -> bool:
        return NormTypeEnum(config.type) == NormTypeEnum.RMSNorm


@DSPostNormRegistry.register_module
class DSPostLNCUDAModule(DSPostNormBase):

    @staticmethod
    def name() -> str:
        return "cuda_post_ln"

    @staticmethod
    def supports_config(config: DSNormConfig) -> bool:
        return NormTypeEnum(config.type) == NormTypeEnum.LayerNorm

    def forward(self, residual, hidden_states, gamma, beta=None) -> torch.Tensor:
        return _norm(self._config, hidden_states, gamma, beta, residual)[0]


# ============================================================================================ unembed
@DSUnembedRegistry.register_module
class DSRaggedUnembed(DSUnembedBase):

    @staticmethod
    def name() -> str:
        return "ragged_unembed"

    @staticmethod
    def supports_config(config: DSUnembedConfig) -> bool:
        return True

    def __init__(self, config, implementation_config):
        super().__init__(config, implementation_config)
        self._output = None

    @property
    def output(self) -> torch.Tensor:
        return self._output

    def forward(self, hidden_states, vocab_embedding, ragged_metadata, bias=None, gamma=None, beta=None):
        c = self._config
        last = R.logits_gather(hidden_states, ragged_metadata.last_token_index())
        if c.norm_type is not None and gamma is not None:
            if NormTypeEnum(c.norm_type) == NormTypeEnum.RMSNorm:
                last = F.rms_norm(last, gamma, c.eps)[0]
            else:
                last = F.layer_norm(last, gamma, beta, c.eps)[0]
        logits = F.linear(last, vocab_embedding)
        self._output = logits if bias is None else logits + bias
        return self._output
