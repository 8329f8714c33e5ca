"""``"activation_checkpointing"`` section (reference: ``runtime/activation_checkpointing/config.py``):

"activation_checkpointing": {
  "partition_activations": false, "contiguous_memory_optimization": false, "cpu_checkpointing": false,
  "number_checkpoints": null, "synchronize_checkpoint_boundary": false, "profile": false
}
"""
from ..config import ActivationCheckpointingConfig

ACTIVATION_CHKPT = "activation_checkpointing"
ACT_CHKPT_PARTITION_ACTIVATIONS = "partition_activations"
ACT_CHKPT_NUMBER_CHECKPOINTS = "number_checkpoints"
ACT_CHKPT_CONTIGUOUS_MEMORY_OPTIMIZATION = "contiguous_memory_optimization"
ACT_CHKPT_SYNCHRONIZE_CHECKPOINT_BOUNDARY = "synchronize_checkpoint_boundary"
ACT_CHKPT_PROFILE = "profile"
ACT_CHKPT_CPU_CHECKPOINTING = "cpu_checkpointing"
ACT_CHKPT_DEFAULT = {k: v.default for k, v in ActivationCheckpointingConfig.model_fields.items()}


class DeepSpeedActivationCheckpointingConfig(ActivationCheckpointingConfig):
    """Constructed from the WHOLE config dictionary, like the reference class."""

    def __init__(self, param_dict=None, **kw):
        section = dict((param_dict or {}).get(ACTIVATION_CHKPT, {})) if param_dict is not None else {}
        section.update(kw)
        super().__init__(**section)
