"""Reusable parameter kinds (reference: ``inference/v2/model_implementations/common_parameters``).  ``finalize``
produces the fused layouts of :class:`DSTransformerModelBase`: ``wqkv`` = q|k|v rows, gated ``w_in`` = gate|up rows,
expert tensors stacked on a leading expert dimension.  Tensor-parallel slicing happens afterwards on the assembled
dictionary (``sharding.shard_weights``), not per parameter."""
from .attn_output_parameters import AttentionOutputParameter
from .embedding_parameters import EmbeddingParameter
from .invfreq_parameters import InvFreqParameter
from .mlp_parameters import GatedMLPParameter, MLP1Parameter, MLP2Parameter
from .moe_parameters import (MoEGatingWeightParameter, UnfusedMoEGatedMLPParameter, UnfusedMoEMLP1Parameter,
                             UnfusedMoEMLP2Parameter)
from .norm_parameters import NormParameter
from .qkv_parameters import (FusedQKVParameter, GQAMegatronQKVParameter, MegatronQKVParameter, UnfusedQKVParameter)
from .unembed_parameters import UnembedParameter
