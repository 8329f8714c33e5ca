from .auto_tp import AutoTP, ReplaceWithTensorSlicing, tp_model_init
from .layers import (ColumnParallel, EmbeddingLayer, GatherReplacedLayerParams, LinearAllreduce, LinearLayer,
                     LmHeadLinearAllreduce, Normalize, RMSNormalize, RowParallel, TensorParallel_Layer, set_autotp_mode)
from .replace_module import generic_injection, replace_module, replace_transformer_layer, revert_transformer_layer
from .replace_policy import DSPolicy, replace_policies
from .inject import module_inject
from .module_quantize import quantize_transformer_layer
from .load_checkpoint import load_model_with_checkpoint
from .utils import policy_to_ds_container
