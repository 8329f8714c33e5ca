"""Architecture policies (reference: ``module_inject/replace_policy.py`` + ``containers/*``).  The reference keeps one
container per architecture to map HF parameter names onto its fused inference kernels; here the policies only carry the
information AutoTP cannot derive structurally (fused-QKV layout, which linears reduce) for the same architecture list."""
from __future__ import annotations


class DSPolicy:
    _orig_layer_class_name = None
    row_parallel = ()
    fused_qkv = ()

    def match(self, module) -> bool:
        return self._orig_layer_class_name is not None and type(module).__name__ == self._orig_layer_class_name


def _mk(name, cls_name, row, fused=()):
    return type(name, (DSPolicy,), {"_orig_layer_class_name": cls_name, "row_parallel": tuple(row), "fused_qkv": tuple(fused)})


HFBertLayerPolicy = _mk("HFBertLayerPolicy", "BertLayer", ("attention.output.dense", "output.dense"))
HFGPT2LayerPolicy = _mk("HFGPT2LayerPolicy", "GPT2Block", ("attn.c_proj", "mlp.c_proj"), ("attn.c_attn",))
HFGPTJLayerPolicy = _mk("HFGPTJLayerPolicy", "GPTJBlock", ("attn.out_proj", "mlp.fc_out"))
HFGPTNEOLayerPolicy = _mk("HFGPTNEOLayerPolicy", "GPTNeoBlock", ("attn.attention.out_proj", "mlp.c_proj"))
GPTNEOXLayerPolicy = _mk("GPTNEOXLayerPolicy", "GPTNeoXLayer", ("attention.dense", "mlp.dense_4h_to_h"), ("attention.query_key_value",))
BLOOMLayerPolicy = _mk("BLOOMLayerPolicy", "BloomBlock", ("self_attention.dense", "mlp.dense_4h_to_h"), ("self_attention.query_key_value",))
HFOPTLayerPolicy = _mk("HFOPTLayerPolicy", "OPTDecoderLayer", ("self_attn.out_proj", "fc2"))
LLAMALayerPolicy = _mk("LLAMALayerPolicy", "LlamaDecoderLayer", ("self_attn.o_proj", "mlp.down_proj"))
LLAMA2LayerPolicy = LLAMALayerPolicy
InternLMLayerPolicy = _mk("InternLMLayerPolicy", "InternLMDecoderLayer", ("self_attn.o_proj", "mlp.down_proj"))
MegatronLayerPolicy = _mk("MegatronLayerPolicy", "ParallelTransformerLayer", ("attention.dense", "mlp.dense_4h_to_h"), ("attention.query_key_value",))
HFDistilBertLayerPolicy = _mk("HFDistilBertLayerPolicy", "TransformerBlock", ("attention.out_lin", "ffn.lin2"))
HFCLIPLayerPolicy = _mk("HFCLIPLayerPolicy", "CLIPEncoderLayer", ("self_attn.out_proj", "mlp.fc2"))
UNetPolicy = _mk("UNetPolicy", "UNet2DConditionModel", ())
VAEPolicy = _mk("VAEPolicy", "AutoencoderKL", ())

replace_policies = [HFBertLayerPolicy, HFGPTNEOLayerPolicy, GPTNEOXLayerPolicy, HFGPTJLayerPolicy, MegatronLayerPolicy,
                    HFGPT2LayerPolicy, BLOOMLayerPolicy, HFOPTLayerPolicy, HFCLIPLayerPolicy, HFDistilBertLayerPolicy,
                    LLAMALayerPolicy, LLAMA2LayerPolicy, InternLMLayerPolicy]
generic_policies = [UNetPolicy, VAEPolicy]
