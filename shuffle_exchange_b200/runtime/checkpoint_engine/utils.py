"""Engine selection (reference: deepspeed/runtime/checkpoint_engine/utils.py:15-38)."""
from ...utils.logging import logger
from .torch_checkpoint_engine import TorchCheckpointEngine


def create_checkpoint_engine(config_params, groups, zero_stage, has_moe_layers, optimizer):
    wc = getattr(config_params, "checkpoint_writer_config", None)
    if wc is not None:
        try:
            if wc.decoupled:
                from .decoupled_checkpoint_engine import DecoupledCheckpointEngine
                return DecoupledCheckpointEngine(config_params)
            if wc.type in ("fast", "mock"):
                from .fast_checkpoint_engine import FastCheckpointEngine
                return FastCheckpointEngine(config_params)
        except Exception as e:  # pragma: no cover
            logger.warning(f"falling back to the torch checkpoint engine: {e}")
    if getattr(config_params, "nebula_config", None) and config_params.nebula_config.get("enabled", False):
        try:
            from .nebula_checkpoint_engine import NebulaCheckpointEngine
            return NebulaCheckpointEngine(config_params=config_params.nebula_config)
        except ImportError as err:
            logger.error(f"No torch_nebula was found! Will fall back to torch.save. Details: {err}")
    return TorchCheckpointEngine(config_params)
