"""ZeRO-3 ``Init`` context used while quantised replacements are constructed (reference
``inference/quantization/quantization_context.py:10-13``): new parameters created inside are partitioned on creation."""
from ...runtime.zero import partition_parameters


class QuantizationContext(partition_parameters.Init):

    def __init__(self, config_dict_or_path, param_swapper=None) -> None:
        super().__init__(config_dict_or_path=config_dict_or_path, param_swapper=param_swapper)
