"""Engine construction helpers (reference: ``inference/v2/engine_factory.py:33-145``)."""
from __future__ import annotations

import logging
import os
import pickle

import torch

from .checkpoint import HuggingFaceCheckpointEngine, InMemoryModelEngine
from .config_v2 import RaggedInferenceEngineConfig
from .engine_v2 import InferenceEngineV2
from .logging import inference_logger
from .model_implementations.policies import POLICIES


def build_engine_from_model(causal_lm, engine_config: RaggedInferenceEngineConfig = None, moe_layers=None, top_k: int = 2,
                            device=None) -> InferenceEngineV2:
    """Serve one of the framework's own training models (``models.llama`` / ``models.mixtral`` layouts) directly."""
    from .model_implementations.llama_v2 import RaggedLlamaModel
    return InferenceEngineV2(lambda sm: RaggedLlamaModel(causal_lm, sm, moe_layers=moe_layers, top_k=top_k),
                             engine_config, device=device)


def _policy_for(model_config, checkpoint_engine):
    mt = getattr(model_config, "model_type", "").lower()
    if mt not in POLICIES:
        raise ValueError(f"Unsupported model type {mt}; supported: {sorted(POLICIES)}")
    if mt == "opt" and not getattr(model_config, "do_layer_norm_before", True):
        raise ValueError("Detected OPT-350m model. This model is not currently supported.")
    return POLICIES[mt](model_config, checkpoint_engine=checkpoint_engine)


def build_engine_from_hf_model(hf_model: torch.nn.Module, engine_config: RaggedInferenceEngineConfig = None,
                               device=None) -> InferenceEngineV2:
    """Wrap an instantiated HuggingFace ``*ForCausalLM`` (parameters are read through ``InMemoryModelEngine``)."""
    engine_config = engine_config or RaggedInferenceEngineConfig()
    ce = InMemoryModelEngine(hf_model)
    policy = _policy_for(hf_model.config, ce)
    policy._engine_config = engine_config
    first = next(iter(hf_model.parameters()), None)
    if first is not None and first.is_floating_point():
        policy.dtype_override = first.dtype
    return InferenceEngineV2(policy, engine_config, device=device)


def build_engine_from_ds_checkpoint(path: str, engine_config: RaggedInferenceEngineConfig,
                                    debug_level: int = logging.INFO) -> InferenceEngineV2:
    """Directory written by ``InferenceEngineV2.serialize``: ``ds_model_config.pkl`` plus, per TP rank,
    ``params_rank_R_of_N.pt`` (one flat byte buffer) and ``metadata_rank_R_of_N.json`` (reference
    ``engine_factory.py:33-66``)."""
    inference_logger(level=debug_level)
    with open(os.path.join(path, "ds_model_config.pkl"), "rb") as f:
        model_config = pickle.load(f)
    mt = getattr(model_config, "model_type", "").lower()
    if mt not in POLICIES:
        raise ValueError(f"Unsupported model type {mt}; supported: {sorted(POLICIES)}")
    policy = POLICIES[mt](model_config, checkpoint_engine=None, inf_checkpoint_path=path)
    policy._engine_config = engine_config
    return InferenceEngineV2(policy, engine_config)


def build_hf_engine(path: str, engine_config: RaggedInferenceEngineConfig = None,
                    debug_level: int = logging.INFO) -> InferenceEngineV2:
    """Engine for a local HuggingFace checkpoint directory (or a serialized v2 checkpoint) of a supported
    architecture: llama, mistral, mixtral, falcon, opt, phi, phi3, qwen, qwen2, qwen2_moe."""
    engine_config = engine_config or RaggedInferenceEngineConfig()
    if os.path.exists(os.path.join(path, "ds_model_config.pkl")):
        return build_engine_from_ds_checkpoint(path, engine_config, debug_level=debug_level)
    inference_logger(level=debug_level)
    ce = HuggingFaceCheckpointEngine(path)
    policy = _policy_for(ce.model_config, ce)
    policy._engine_config = engine_config
    return InferenceEngineV2(policy, engine_config)
