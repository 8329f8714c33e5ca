"""Model implementations of the v2 engine (reference ``inference/v2/model_implementations``)."""
from .inference_transformer_base import ArchSpec, DSInferenceModelBase, DSTransformerModelBase, LayerWeights
from .inference_policy_base import ContainerMap, InferenceV2Policy
from .policies import (POLICIES, FalconPolicy, Llama2Policy, MistralPolicy, MixtralPolicy, OPTPolicy, Phi3Policy,
                       PhiPolicy, Qwen2MoePolicy, Qwen2Policy, QwenPolicy)
from .llama_v2 import RaggedLlamaModel
