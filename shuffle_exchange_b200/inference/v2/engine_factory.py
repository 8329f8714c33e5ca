"""Engine construction helpers (reference: ``inference/v2/engine_factory.py:33-145``)."""
from __future__ import annotations

import torch

from .config_v2 import RaggedInferenceEngineConfig
from .engine_v2 import InferenceEngineV2

_SUPPORTED = ("llama", "mistral", "mixtral", "qwen2", "qwen2_moe", "phi3")


def build_engine_from_model(causal_lm, engine_config: RaggedInferenceEngineConfig = None, moe_layers=None, top_k: int = 2,
                            device=None) -> InferenceEngineV2:
    from .model_implementations.llama_v2 import RaggedLlamaModel
    return InferenceEngineV2(lambda sm: RaggedLlamaModel(causal_lm, sm, moe_layers=moe_layers, top_k=top_k),
                             engine_config, device=device)


def build_hf_engine(path: str, engine_config: RaggedInferenceEngineConfig = None, debug_level: int = 0) -> InferenceEngineV2:
    """Load a HuggingFace checkpoint directory of a supported architecture and wrap it in an engine."""
    import json
    import os
    from ...models.llama import LlamaConfig, LlamaForCausalLM
    with open(os.path.join(path, "config.json")) as f:
        hf = json.load(f)
    mt = hf.get("model_type", "").lower()
    if mt not in _SUPPORTED:
        raise ValueError(f"Unsupported model type {mt}; supported: {_SUPPORTED}")
    if mt in ("mixtral", "qwen2_moe"):
        raise NotImplementedError("MoE checkpoints: build the model with models.mixtral and use build_engine_from_model")
    cfg = LlamaConfig(vocab_size=hf["vocab_size"], hidden_size=hf["hidden_size"], intermediate_size=hf["intermediate_size"],
                      num_hidden_layers=hf["num_hidden_layers"], num_attention_heads=hf["num_attention_heads"],
                      num_key_value_heads=hf.get("num_key_value_heads", hf["num_attention_heads"]),
                      max_position_embeddings=hf.get("max_position_embeddings", 8192),
                      rope_theta=hf.get("rope_theta", 10000.0), rms_norm_eps=hf.get("rms_norm_eps", 1e-5),
                      tie_word_embeddings=hf.get("tie_word_embeddings", False), activation_checkpointing=False)
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    model = LlamaForCausalLM(cfg, device=dev, dtype=torch.bfloat16)
    sd = {}
    for fn in sorted(os.listdir(path)):
        if fn.endswith(".safetensors"):
            from safetensors.torch import load_file
            sd.update(load_file(os.path.join(path, fn)))
        elif fn.endswith(".bin") and fn.startswith("pytorch_model"):
            sd.update(torch.load(os.path.join(path, fn), map_location="cpu", weights_only=True))
    model.load_hf_state_dict(sd)
    return build_engine_from_model(model.eval(), engine_config)
