"""Module configs (reference ``inference/v2/modules/configs/*``)."""
from enum import Enum
from typing import Dict, Optional

from ...inference_utils import ActivationType, DtypeEnum, NormTypeEnum
from ..ds_module import DSModuleConfig
from .....runtime.config_utils import DeepSpeedConfigModel


class PositionalEmbeddingType(Enum):
    none = "none"
    rotate_half = "rotate_half"
    rotate_every_other = "rotate_every_other"
    alibi = "alibi"


class RotateHalfConfig(DeepSpeedConfigModel):
    use_trained_freqs: bool = False
    theta_base: float = 10_000.0
    rotate_dim: Optional[int] = None      # None: rotate the whole head; < head_size: partial rotary (Phi, GPT-NeoX)


class MaskingType(Enum):
    none = "none"
    causal = "causal"
    local = "local"
    symmetric = "symmetric"
    asymmetric = "asymmetric"


class DSSelfAttentionConfig(DSModuleConfig):
    n_heads_q: int
    n_heads_kv: int
    head_size: int
    max_sequences: int
    scale_factor: float = 1.0
    input_dtype: DtypeEnum = DtypeEnum.bf16
    output_dtype: DtypeEnum = DtypeEnum.bf16
    masking_type: MaskingType = MaskingType.causal
    masking_args: Dict = {}
    positional_embedding_type: PositionalEmbeddingType = PositionalEmbeddingType.none
    positional_embedding_config: Optional[RotateHalfConfig] = None


class DSEmbeddingsConfig(DSModuleConfig):
    residual_dtype: DtypeEnum = DtypeEnum.bf16
    embedding_dim: int
    positional_embedding: bool = False
    positional_offset: int = 0            # OPT stores position p at row p + 2
    use_token_type: bool = False
    output_normalization: Optional[NormTypeEnum] = None


class DSLinearConfig(DSModuleConfig):
    in_channels: int
    out_channels: int                     # after gating: a gated layer's weight is [2 * out_channels, in_channels]
    activation: ActivationType = ActivationType.IDENTITY
    input_dtype: DtypeEnum = DtypeEnum.bf16
    output_dtype: DtypeEnum = DtypeEnum.bf16


class DSMoEConfig(DSModuleConfig):
    model_dim: int
    intermediate_features: int
    n_experts: int
    top_k: int = 1
    input_dtype: DtypeEnum = DtypeEnum.bf16
    output_dtype: DtypeEnum = DtypeEnum.bf16
    activation: ActivationType = ActivationType.IDENTITY
    normalize_scores: bool = False


class DSNormConfig(DSModuleConfig):
    type: NormTypeEnum
    channels: int
    residual_dtype: DtypeEnum = DtypeEnum.bf16
    input_dtype: DtypeEnum = DtypeEnum.bf16
    output_dtype: DtypeEnum = DtypeEnum.bf16
    eps: float = 1e-5


class DSUnembedConfig(DSModuleConfig):
    dtype: DtypeEnum = DtypeEnum.bf16
    norm_type: Optional[NormTypeEnum] = None
    model_dim: int
    max_sequences: int
    vocab_size: int
    eps: float = 1e-5
