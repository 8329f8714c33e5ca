"""FastGen (inference v2): ragged continuous batching over a paged KV cache (reference: ``deepspeed/inference/v2``)."""
from .config_v2 import DeepSpeedTPConfig, QuantizationConfig, RaggedInferenceEngineConfig  # noqa: F401
from .engine_factory import (build_engine_from_ds_checkpoint, build_engine_from_hf_model,  # noqa: F401
                             build_engine_from_model, build_hf_engine)
from .engine_v2 import InferenceEngineV2  # noqa: F401
from .scheduling_utils import SchedulingError, SchedulingResult  # noqa: F401
