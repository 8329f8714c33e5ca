"""Architecture policy registry (reference: ``module_inject/replace_policy.py``): the ordered list
``replace_transformer_layer`` tries on every module, plus the diffusers generic policies."""
from __future__ import annotations

from .containers import (BLOOMLayerPolicy, DSPolicy, GPTNEOXLayerPolicy, HFBertLayerPolicy, HFCLIPLayerPolicy,
                         HFDistilBertLayerPolicy, HFGPT2LayerPolicy, HFGPTJLayerPolicy, HFGPTNEOLayerPolicy,
                         HFOPTLayerPolicy, InternLMLayerPolicy, LLAMA2LayerPolicy, LLAMALayerPolicy, MegatronLayerPolicy,
                         MegatronMoELayerPolicy, TransformerPolicy, UNetPolicy, VAEPolicy)

# transformer-based policies
replace_policies = [HFBertLayerPolicy, HFGPTNEOLayerPolicy, GPTNEOXLayerPolicy, HFGPTJLayerPolicy, MegatronLayerPolicy,
                    HFGPT2LayerPolicy, BLOOMLayerPolicy, HFOPTLayerPolicy, HFCLIPLayerPolicy, HFDistilBertLayerPolicy,
                    LLAMALayerPolicy, LLAMA2LayerPolicy, InternLMLayerPolicy, MegatronMoELayerPolicy]

# non-transformer-based policies
generic_policies = [UNetPolicy, VAEPolicy]
