"""Per-architecture injection policies and containers (reference: ``module_inject/containers/__init__.py``)."""
from .base import (AttentionTensors, BaseTransformerContainer, DSPolicy, HFLayerAdapter, MLPTensors, NormTensors,
                   TransformerPolicy)
from .bert import DS_BERTContainer, HFBertLayerPolicy
from .bloom import BLOOMLayerPolicy, DS_BloomContainer
from .clip import DS_CLIPContainer, HFCLIPLayerPolicy
from .distil_bert import DS_DistilBERTContainer, HFDistilBertLayerPolicy
from .gpt2 import DS_GPT2Container, HFGPT2LayerPolicy
from .gptj import DS_GPTJContainer, HFGPTJLayerPolicy
from .gptneo import DS_GPTNEOContainer, HFGPTNEOLayerPolicy
from .gptneox import DS_GPTNEOXContainer, GPTNEOXLayerPolicy
from .internlm import DS_InternLMContainer, InternLMLayerPolicy
from .llama import DS_LLAMA2Container, DS_LLAMAContainer, LLAMA2LayerPolicy, LLAMALayerPolicy
from .megatron_gpt import DS_MegatronGPTContainer, MegatronLayerPolicy
from .megatron_gpt_moe import BaseTransformerMoEContainer, DS_MegatronGPTMoEContainer, MegatronMoELayerPolicy
from .opt import DS_OPTContainer, HFOPTLayerPolicy
from .unet import UNetPolicy
from .vae import VAEPolicy

# policy -> container (reference ``replace_module.py:policy_to_ds_container``)
POLICY_TO_CONTAINER = {
    HFGPT2LayerPolicy: DS_GPT2Container, HFBertLayerPolicy: DS_BERTContainer, HFOPTLayerPolicy: DS_OPTContainer,
    LLAMALayerPolicy: DS_LLAMAContainer, LLAMA2LayerPolicy: DS_LLAMA2Container, InternLMLayerPolicy: DS_InternLMContainer,
    HFGPTJLayerPolicy: DS_GPTJContainer, HFGPTNEOLayerPolicy: DS_GPTNEOContainer, GPTNEOXLayerPolicy: DS_GPTNEOXContainer,
    BLOOMLayerPolicy: DS_BloomContainer, HFDistilBertLayerPolicy: DS_DistilBERTContainer,
    HFCLIPLayerPolicy: DS_CLIPContainer, MegatronLayerPolicy: DS_MegatronGPTContainer,
    MegatronMoELayerPolicy: DS_MegatronGPTMoEContainer,
}
