"""Names and defaults of the top-level config keys (reference: ``deepspeed/runtime/constants.py``).

The names are the import surface user code and integrations rely on (``from deepspeed.runtime.constants import
TRAIN_BATCH_SIZE`` ...), so they are kept; each key is written next to its default as one ``NAME, NAME_DEFAULT = key,
default`` pair, and :mod:`.config` builds its scalar table from these pairs (``tests/test_config.py`` checks that every
pair listed in ``SCALAR_KEYS`` is honoured by :class:`DeepSpeedConfig`).  Structured sections (``fp16``, ``bf16``,
``zero_optimization``, ...) are pydantic models in :mod:`.config`; for those only the section name and the field names
are given here.
"""
from .model_checkpointing.constants import ValidationMode  # noqa: F401  (re-exported like the reference)

# ------------------------------------------------------------------------------------------ routes
ROUTE_TRAIN, ROUTE_EVAL, ROUTE_PREDICT, ROUTE_ENCODE = "train", "eval", "predict", "encode"

# ----------------------------------------------------------------------------------- batch triad
TRAIN_BATCH_SIZE, TRAIN_BATCH_SIZE_DEFAULT = "train_batch_size", None
TRAIN_MICRO_BATCH_SIZE_PER_GPU, TRAIN_MICRO_BATCH_SIZE_PER_GPU_DEFAULT = "train_micro_batch_size_per_gpu", None
GRADIENT_ACCUMULATION_STEPS, GRADIENT_ACCUMULATION_STEPS_DEFAULT = "gradient_accumulation_steps", None

# --------------------------------------------------------------------- optimizer / scheduler blocks
OPTIMIZER, OPTIMIZER_TYPE_DEFAULT, OPTIMIZER_PARAMS = "optimizer", None, "params"
SCHEDULER, SCHEDULER_TYPE_DEFAULT, SCHEDULER_PARAMS = "scheduler", None, "params"
TYPE = "type"
LEGACY_FUSION, LEGACY_FUSION_DEFAULT = "legacy_fusion", False
MAX_GRAD_NORM = "max_grad_norm"
ZERO_ALLOW_UNTESTED_OPTIMIZER, ZERO_ALLOW_UNTESTED_OPTIMIZER_DEFAULT = "zero_allow_untested_optimizer", False
ZERO_FORCE_DS_CPU_OPTIMIZER, ZERO_FORCE_DS_CPU_OPTIMIZER_DEFAULT = "zero_force_ds_cpu_optimizer", True

# ------------------------------------------------------------------------------ top-level scalars
STEPS_PER_PRINT, STEPS_PER_PRINT_DEFAULT = "steps_per_print", None
SPARSE_GRADIENTS, SPARSE_GRADIENTS_DEFAULT = "sparse_gradients", False
GRADIENT_CLIPPING, GRADIENT_CLIPPING_DEFAULT = "gradient_clipping", 0.0
GRAPH_HARVESTING, GRAPH_HARVESTING_DEFAULT = "graph_harvesting", False
COMMUNICATION_DATA_TYPE, COMMUNICATION_DATA_TYPE_DEFAULT = "communication_data_type", None
SEQ_PARALLEL_COMMUNICATION_DATA_TYPE, SEQ_PARALLEL_COMMUNICATION_DATA_TYPE_DEFAULT = \
    "seq_parallel_communication_data_type", "fp32"
PRESCALE_GRADIENTS, PRESCALE_GRADIENTS_DEFAULT = "prescale_gradients", False
GRADIENT_PREDIVIDE_FACTOR, GRADIENT_PREDIVIDE_FACTOR_DEFAULT = "gradient_predivide_factor", 1.0
DISABLE_ALLGATHER, DISABLE_ALLGATHER_DEFAULT = "disable_allgather", False
DUMP_STATE, DUMP_STATE_DEFAULT = "dump_state", False
VOCABULARY_SIZE, VOCABULARY_SIZE_DEFAULT = "vocabulary_size", None
WALL_CLOCK_BREAKDOWN, WALL_CLOCK_BREAKDOWN_DEFAULT = "wall_clock_breakdown", False
MEMORY_BREAKDOWN, MEMORY_BREAKDOWN_DEFAULT = "memory_breakdown", False
DATALOADER_DROP_LAST, DATALOADER_DROP_LAST_DEFAULT = "dataloader_drop_last", False
USE_DATA_BEFORE_EXPERT_PARALLEL, USE_DATA_BEFORE_EXPERT_PARALLEL_DEFAULT = "use_data_before_expert_parallelism", False
TENSOR_CORE_ALIGN_SIZE = 8

# attribute of DeepSpeedConfig -> (key, default): the scalar table of runtime/config.py.  Where this framework's
# effective default differs from the reference's "unset" marker the effective value is given (steps_per_print 10 and
# vocabulary_size 1201 are what the reference's getters return for an absent key).
SCALAR_KEYS = {
    "steps_per_print": (STEPS_PER_PRINT, 10),
    "dump_state": (DUMP_STATE, DUMP_STATE_DEFAULT),
    "disable_allgather": (DISABLE_ALLGATHER, DISABLE_ALLGATHER_DEFAULT),
    "prescale_gradients": (PRESCALE_GRADIENTS, PRESCALE_GRADIENTS_DEFAULT),
    "gradient_predivide_factor": (GRADIENT_PREDIVIDE_FACTOR, GRADIENT_PREDIVIDE_FACTOR_DEFAULT),
    "sparse_gradients_enabled": (SPARSE_GRADIENTS, SPARSE_GRADIENTS_DEFAULT),
    "gradient_clipping": (GRADIENT_CLIPPING, GRADIENT_CLIPPING_DEFAULT),
    "graph_harvesting": (GRAPH_HARVESTING, GRAPH_HARVESTING_DEFAULT),
    "zero_allow_untested_optimizer": (ZERO_ALLOW_UNTESTED_OPTIMIZER, ZERO_ALLOW_UNTESTED_OPTIMIZER_DEFAULT),
    "zero_force_ds_cpu_optimizer": (ZERO_FORCE_DS_CPU_OPTIMIZER, ZERO_FORCE_DS_CPU_OPTIMIZER_DEFAULT),
    "wall_clock_breakdown": (WALL_CLOCK_BREAKDOWN, WALL_CLOCK_BREAKDOWN_DEFAULT),
    "memory_breakdown": (MEMORY_BREAKDOWN, MEMORY_BREAKDOWN_DEFAULT),
    "use_data_before_expert_parallel_": (USE_DATA_BEFORE_EXPERT_PARALLEL, USE_DATA_BEFORE_EXPERT_PARALLEL_DEFAULT),
    "dataloader_drop_last": (DATALOADER_DROP_LAST, DATALOADER_DROP_LAST_DEFAULT),
    "vocabulary_size": (VOCABULARY_SIZE, 1201),
    "seq_parallel_communication_data_type": (SEQ_PARALLEL_COMMUNICATION_DATA_TYPE,
                                             SEQ_PARALLEL_COMMUNICATION_DATA_TYPE_DEFAULT),
}

# ------------------------------------------------------------------------------- precision blocks
BFLOAT16, BFLOAT16_OLD = "bf16", "bfloat16"          # the old spelling is still accepted
BFLOAT16_ENABLED, BFLOAT16_ENABLED_DEFAULT = "enabled", False
CHECK_OVERFLOW, BFLOAT16_CHECK_OVERFLOW_DEFAULT = "check_overflow", False
BFLOAT16_IMMEDIATE_GRAD_UPDATE, BFLOAT16_IMMEDIATE_GRAD_UPDATE_DEFAULT = "immediate_grad_update", True

FP16 = "fp16"
FP16_ENABLED, FP16_ENABLED_DEFAULT = "enabled", False
FP16_LOSS_SCALE, FP16_LOSS_SCALE_DEFAULT = "loss_scale", 0          # 0 selects dynamic loss scaling
FP16_AUTO_CAST, FP16_AUTO_CAST_DEFAULT = "auto_cast", False
FP16_INITIAL_SCALE_POWER, FP16_INITIAL_SCALE_POWER_DEFAULT = "initial_scale_power", 16
FP16_LOSS_SCALE_WINDOW, FP16_LOSS_SCALE_WINDOW_DEFAULT = "loss_scale_window", 1000
FP16_HYSTERESIS, FP16_HYSTERESIS_DEFAULT = "hysteresis", 2
FP16_CONSECUTIVE_HYSTERESIS, FP16_CONSECUTIVE_HYSTERESIS_DEFAULT = "consecutive_hysteresis", False
FP16_MIN_LOSS_SCALE, FP16_MIN_LOSS_SCALE_DEFAULT = "min_loss_scale", 1
FP16_MASTER_WEIGHTS_AND_GRADS, FP16_MASTER_WEIGHTS_AND_GRADS_DEFAULT = "fp16_master_weights_and_grads", False

AMP = "amp"
AMP_ENABLED, AMP_ENABLED_DEFAULT = "enabled", False

TORCH_AUTOCAST = "torch_autocast"
TORCH_AUTOCAST_ENABLED, TORCH_AUTOCAST_ENABLED_DEFAULT = "enabled", False
TORCH_AUTOCAST_DTYPE = "dtype"
TORCH_AUTOCAST_LOWER_PRECISION_SAFE_MODULES = "lower_precision_safe_modules"

DATA_TYPES = "data_types"
GRAD_ACCUM_DTYPE, GRAD_ACCUM_DTYPE_DEFAULT = "grad_accum_dtype", None

# ------------------------------------------------------------------------------- sparse attention
SPARSE_ATTENTION = "sparse_attention"
SPARSE_DENSE_MODE, SPARSE_FIXED_MODE, SPARSE_VARIABLE_MODE = "dense", "fixed", "variable"
SPARSE_BIGBIRD_MODE, SPARSE_BSLONGFORMER_MODE = "bigbird", "bslongformer"
SPARSE_MODE, SPARSE_MODE_DEFAULT = "mode", SPARSE_FIXED_MODE
SPARSE_BLOCK, SPARSE_BLOCK_DEFAULT = "block", 16
SPARSE_DIFFERENT_LAYOUT_PER_HEAD, SPARSE_DIFFERENT_LAYOUT_PER_HEAD_DEFAULT = "different_layout_per_head", False
SPARSE_NUM_LOCAL_BLOCKS, SPARSE_NUM_LOCAL_BLOCKS_DEFAULT = "num_local_blocks", 4
SPARSE_NUM_GLOBAL_BLOCKS, SPARSE_NUM_GLOBAL_BLOCKS_DEFAULT = "num_global_blocks", 1
SPARSE_ATTENTION_TYPE, SPARSE_ATTENTION_TYPE_DEFAULT = "attention", "bidirectional"
SPARSE_HORIZONTAL_GLOBAL_ATTENTION, SPARSE_HORIZONTAL_GLOBAL_ATTENTION_DEFAULT = "horizontal_global_attention", False
SPARSE_NUM_DIFFERENT_GLOBAL_PATTERNS, SPARSE_NUM_DIFFERENT_GLOBAL_PATTERNS_DEFAULT = \
    "num_different_global_patterns", 1
SPARSE_NUM_RANDOM_BLOCKS, SPARSE_NUM_RANDOM_BLOCKS_DEFAULT = "num_random_blocks", 0
SPARSE_LOCAL_WINDOW_BLOCKS, SPARSE_LOCAL_WINDOW_BLOCKS_DEFAULT = "local_window_blocks", [4]
SPARSE_GLOBAL_BLOCK_INDICES, SPARSE_GLOBAL_BLOCK_INDICES_DEFAULT = "global_block_indices", [0]
SPARSE_GLOBAL_BLOCK_END_INDICES, SPARSE_GLOBAL_BLOCK_END_INDICES_DEFAULT = "global_block_end_indices", None
SPARSE_NUM_SLIDING_WINDOW_BLOCKS, SPARSE_NUM_SLIDING_WINDOW_BLOCKS_DEFAULT = "num_sliding_window_blocks", 3

# ------------------------------------------------------- eigenvalue / progressive layer drop blocks
EIGENVALUE = "eigenvalue"
EIGENVALUE_ENABLED, EIGENVALUE_ENABLED_DEFAULT = "enabled", False
EIGENVALUE_VERBOSE, EIGENVALUE_VERBOSE_DEFAULT = "verbose", False
EIGENVALUE_MAX_ITER, EIGENVALUE_MAX_ITER_DEFAULT = "max_iter", 100
EIGENVALUE_TOL, EIGENVALUE_TOL_DEFAULT = "tol", 1e-2
EIGENVALUE_STABILITY, EIGENVALUE_STABILITY_DEFAULT = "stability", 1e-6
EIGENVALUE_GAS_BOUNDARY_RESOLUTION, EIGENVALUE_GAS_BOUNDARY_RESOLUTION_DEFAULT = "gas_boundary_resolution", 1
EIGENVALUE_LAYER_NAME, EIGENVALUE_LAYER_NAME_DEFAULT = "layer_name", "bert.encoder.layer"
EIGENVALUE_LAYER_NUM, EIGENVALUE_LAYER_NUM_DEFAULT = "layer_num", 0

PROGRESSIVE_LAYER_DROP = "progressive_layer_drop"
PLD_ENABLED, PLD_ENABLED_DEFAULT = "enabled", False
PLD_THETA, PLD_THETA_DEFAULT = "theta", 1.0
PLD_GAMMA, PLD_GAMMA_DEFAULT = "gamma", 0.001

# ------------------------------------------------------------------------------- checkpoint block
CHECKPOINT = "checkpoint"
CHECKPOINT_TAG_VALIDATION = "tag_validation"
CHECKPOINT_TAG_VALIDATION_DEFAULT = ValidationMode.WARN
CHECKPOINT_TAG_VALIDATION_MODES = [ValidationMode.WARN, ValidationMode.IGNORE, ValidationMode.FAIL]
LOAD_UNIVERSAL_CHECKPOINT, LOAD_UNIVERSAL_CHECKPOINT_DEFAULT = "load_universal", False
USE_NODE_LOCAL_STORAGE_CHECKPOINT, USE_NODE_LOCAL_STORAGE_CHECKPOINT_DEFAULT = "use_node_local_storage", False
CHECKPOINT_PARALLEL_WRITE = "parallel_write"
CHECKPOINT_PARALLEL_WRITE_PIPELINE_STAGE, CHECKPOINT_PARALLEL_WRITE_PIPELINE_STAGE_DEFAULT = "pipeline_stage", False

# ---------------------------------------------------------------- attribute names and bookkeeping
PIPE_REPLICATED = "ds_pipe_replicated"        # marks parameters replicated across pipeline stages
DATA_PARALLEL_GROUP = "data_parallel_group"
GLOBAL_RANK = "global_rank"
