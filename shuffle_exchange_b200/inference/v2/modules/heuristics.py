"""Pick a module implementation for a config (reference ``inference/v2/modules/heuristics.py:35-195``)."""
from __future__ import annotations

from ..config_v2 import RaggedInferenceEngineConfig
from ..inference_utils import NormTypeEnum
from . import implementations  # noqa: F401  (registers the implementations)
from .configs import (DSEmbeddingsConfig, DSLinearConfig, DSMoEConfig, DSNormConfig, DSSelfAttentionConfig,
                      DSUnembedConfig)
from .interfaces import (DSEmbeddingBase, DSEmbeddingRegistry, DSLinearBase, DSLinearRegistry, DSMoEBase,
                         DSMoERegistry, DSPostNormBase, DSPostNormRegistry, DSPreNormBase, DSPreNormRegistry,
                         DSSelfAttentionBase, DSSelfAttentionRegistry, DSUnembedBase, DSUnembedRegistry)
from .module_registry import ConfigBundle


def instantiate_attention(attention_config: DSSelfAttentionConfig,
                          engine_config: RaggedInferenceEngineConfig) -> DSSelfAttentionBase:
    return DSSelfAttentionRegistry.instantiate_config(ConfigBundle(name="dense_blocked_attention",
                                                                   config=attention_config))


def instantiate_embed(embed_config: DSEmbeddingsConfig, engine_config: RaggedInferenceEngineConfig) -> DSEmbeddingBase:
    return DSEmbeddingRegistry.instantiate_config(ConfigBundle(name="ragged_embedding", config=embed_config))


def instantiate_linear(linear_config: DSLinearConfig, engine_config: RaggedInferenceEngineConfig) -> DSLinearBase:
    mode = getattr(getattr(engine_config, "quantization", None), "quantization_mode", None)
    if mode is None:
        name = "blas_fp_linear"
    elif mode == "wf6af16":
        name = "quantized_wf6af16_linear"
        impl = DSLinearRegistry.registry[name]
        if not impl.supports_config(linear_config):
            name = "blas_fp_linear"       # shapes the FP6 layout cannot hold stay in 16-bit
    else:
        raise ValueError(f"Unsupported quantization mode: {mode}")
    return DSLinearRegistry.instantiate_config(ConfigBundle(name=name, config=linear_config))


def instantiate_moe(moe_config: DSMoEConfig, engine_config: RaggedInferenceEngineConfig) -> DSMoEBase:
    return DSMoERegistry.instantiate_config(ConfigBundle(name="cutlass_multi_gemm_moe", config=moe_config,
                                                         implementation_config={"weight_dtype": moe_config.input_dtype}))


def instantiate_post_norm(norm_config: DSNormConfig, engine_config: RaggedInferenceEngineConfig) -> DSPostNormBase:
    return DSPostNormRegistry.instantiate_config(ConfigBundle(name="cuda_post_ln", config=norm_config))


def instantiate_pre_norm(norm_config: DSNormConfig, engine_config: RaggedInferenceEngineConfig) -> DSPreNormBase:
    name = "cuda_pre_rms" if NormTypeEnum(norm_config.type) == NormTypeEnum.RMSNorm else "cuda_pre_ln"
    return DSPreNormRegistry.instantiate_config(ConfigBundle(name=name, config=norm_config))


def instantiate_unembed(unembed_config: DSUnembedConfig, engine_config: RaggedInferenceEngineConfig) -> DSUnembedBase:
    return DSUnembedRegistry.instantiate_config(ConfigBundle(name="ragged_unembed", config=unembed_config))
