"""``DeepSpeedBERTInference`` — architecture-named alias of the v1 inference layer (reference ``transformers/ds_bert.py``); the per-model
differences (masking, rotary, ALiBi, norm type) are all carried by ``DeepSpeedInferenceConfig``."""
from .ds_transformer import DeepSpeedTransformerInference


class DeepSpeedBERTInference(DeepSpeedTransformerInference):

    def __init__(self, config, mp_group=None, quantize_scales=None, quantize_groups=1, merge_count=1,
                 mlp_extra_grouping=False):
        super().__init__(config, mp_group, quantize_scales, quantize_groups, merge_count, mlp_extra_grouping)
