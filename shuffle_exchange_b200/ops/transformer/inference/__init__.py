"""v1 inference transformer ops (reference ``deepspeed/ops/transformer/inference``)."""
from .config import DeepSpeedInferenceConfig
from .bias_add import nhwc_bias_add
from .ds_attention import BloomSelfAttention, DeepSpeedSelfAttention
from .ds_mlp import DeepSpeedMLP
from .moe_inference import DeepSpeedMoEInference, DeepSpeedMoEInferenceConfig
from .diffusers_attention import DeepSpeedDiffusersAttention
from .diffusers_transformer_block import DeepSpeedDiffusersTransformerBlock
from .diffusers_2d_transformer import Diffusers2DTransformerConfig
