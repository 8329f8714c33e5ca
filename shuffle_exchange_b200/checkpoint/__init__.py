from .deepspeed_checkpoint import DeepSpeedCheckpoint
from .reshape_utils import get_files, get_files_with_prefix, get_zero_files, merge_state, partition_data
from .universal_checkpoint import enable_universal_checkpoint, load_hp_checkpoint_state
from .utils import (clone_tensors_for_torch_save, get_layer_ckpt_name_for_rank, get_model_ckpt_name_for_rank,
                    get_zero_ckpt_name_for_rank)
from .reshape_3d_utils import get_model_3d_descriptor, model_3d_desc
from .reshape_meg_2d import get_mpu_ranks, meg_2d_parallel_map, reshape_meg_2d_parallel
from .zero_checkpoint import ZeROCheckpoint
