"""Autotuning constants (reference: ``deepspeed/autotuning/constants.py``)."""
import os

DEFAULT_TEMPLATE_PATH_ZERO_0 = os.path.join(os.path.dirname(os.path.realpath(__file__)), "config_templates", "template_zero0.json")
DEFAULT_TEMPLATE_PATH_ZERO_1 = os.path.join(os.path.dirname(os.path.realpath(__file__)), "config_templates", "template_zero1.json")
DEFAULT_TEMPLATE_PATH_ZERO_2 = os.path.join(os.path.dirname(os.path.realpath(__file__)), "config_templates", "template_zero2.json")
DEFAULT_TEMPLATE_PATH_ZERO_3 = os.path.join(os.path.dirname(os.path.realpath(__file__)), "config_templates", "template_zero3.json")
METRIC_PERCENT_DIFF_CONST = 0.05
DS_CONFIG = "ds_config"
BUFSIZE = 1
AUTOTUNING = "autotuning"
AUTOTUNING_ENABLED, AUTOTUNING_ENABLED_DEFAULT = "enabled", False
AUTOTUNING_FAST, AUTOTUNING_FAST_DEFAULT = "fast", True
AUTOTUNING_RESULTS_DIR, AUTOTUNING_RESULTS_DIR_DEFAULT = "results_dir", "autotuning_results"
AUTOTUNING_EXPS_DIR, AUTOTUNING_EXPS_DIR_DEFAULT = "exps_dir", "autotuning_exps"
AUTOTUNING_OVERWRITE, AUTOTUNING_OVERWRITE_DEFAULT = "overwrite", True
AUTOTUNING_START_PROFILE_STEP, AUTOTUNING_START_PROFILE_STEP_DEFAULT = "start_profile_step", 3
AUTOTUNING_END_PROFILE_STEP, AUTOTUNING_END_PROFILE_STEP_DEFAULT = "end_profile_step", 5
AUTOTUNING_METRIC_PATH, AUTOTUNING_METRIC_PATH_DEFAULT = "metric_path", None
AUTOTUNING_TUNER_TYPE = "tuner_type"
AUTOTUNING_TUNER_GRIDSEARCH, AUTOTUNING_TUNER_RANDOM, AUTOTUNING_TUNER_MODELBASED = "gridsearch", "random", "model_based"
AUTOTUNING_TUNER_TYPE_DEFAULT = AUTOTUNING_TUNER_GRIDSEARCH
AUTOTUNING_TUNER_EARLY_STOPPING, AUTOTUNING_TUNER_EARLY_STOPPING_DEFAULT = "tuner_early_stopping", 5
AUTOTUNING_TUNER_NUM_TRIALS, AUTOTUNING_TUNER_NUM_TRIALS_DEFAULT = "tuner_num_trials", 50
AUTOTUNING_ARG_MAPPINGS, AUTOTUNING_ARG_MAPPINGS_DEFAULT = "arg_mappings", None
AUTOTUNING_MAX_TRAIN_BATCH_SIZE, AUTOTUNING_MAX_TRAIN_BATCH_SIZE_DEFAULT = "max_train_batch_size", None
AUTOTUNING_MIN_TRAIN_BATCH_SIZE, AUTOTUNING_MIN_TRAIN_BATCH_SIZE_DEFAULT = "min_train_batch_size", 1
AUTOTUNING_MAX_TRAIN_MICRO_BATCH_SIZE_PER_GPU, AUTOTUNING_MAX_TRAIN_MICRO_BATCH_SIZE_PER_GPU_DEFAULT = "max_train_micro_batch_size_per_gpu", 1024
AUTOTUNING_MIN_TRAIN_MICRO_BATCH_SIZE_PER_GPU, AUTOTUNING_MIN_TRAIN_MICRO_BATCH_SIZE_PER_GPU_DEFAULT = "min_train_micro_batch_size_per_gpu", 1
AUTOTUNING_NUM_TUNING_MICRO_BATCH_SIZES, AUTOTUNING_NUM_TUNING_MICRO_BATCH_SIZES_DEFAULT = "num_tuning_micro_batch_sizes", 3
AUTOTUNING_MP_SIZE, AUTOTUNING_MP_SIZE_DEFAULT = "mp_size", 1
AUTOTUNING_METRIC = "metric"
AUTOTUNING_METRIC_LATENCY, AUTOTUNING_METRIC_THROUGHPUT, AUTOTUNING_METRIC_FLOPS, AUTOTUNING_METRIC_FORWARD, \
    AUTOTUNING_METRIC_BACKWRAD, AUTOTUNING_METRIC_STEPS = "latency", "throughput", "flops", "forward", "backward", "steps"
AUTOTUNING_METRIC_DEFAULT = AUTOTUNING_METRIC_THROUGHPUT
MODEL_INFO = "model_info"
MODEL_INFO_PROFILE, MODEL_INFO_PROFILE_DEFAULT = "profile", False
MODEL_INFO_NUM_PARAMS, MODEL_INFO_NUM_PARAMS_DEFAULT = "num_params", None
MODEL_INFO_HIDDEN_SIZE, MODEL_INFO_NUM_LAYERS = "hidden_size", "num_layers"
AUTOTUNING_MODEL_INFO_PATH, AUTOTUNING_MODEL_INFO_PATH_DEFAULT = "model_info_path", None
GRADIENT_ACCUMULATION_STEPS = "gradient_accumulation_steps"
TRAIN_MICRO_BATCH_SIZE_PER_GPU = "train_micro_batch_size_per_gpu"
TRAIN_BATCH_SIZE = "train_batch_size"
OFFLOAD_OPTIMIZER, OFFLOAD_PARAM, ZERO_OPTIMIZATION = "offload_optimizer", "offload_param", "zero_optimization"
ZERO_OPTIMIZATION_STAGE, ZERO_OPTIMIZATION_STAGE_DEFAULT = "stage", 0
DEFAULT_HF_CONFIG = {"train_batch_size": "auto", "train_micro_batch_size_per_gpu": "auto", "gradient_accumulation_steps": "auto"}
DEFAULT_MIN_MEM_CONFIG = {"train_micro_batch_size_per_gpu": 1, "zero_optimization": {"stage": 3}, "memory_break_down": False}
DEFAULT_TUNING_SPACE_ZERO_0 = {"zero_optimization": {"stage": 0}}
DEFAULT_TUNING_SPACE_ZERO_1 = {"zero_optimization": {"stage": 1, "reduce_bucket_size": [5e7, 5e8, 1e9], "allgather_bucket_size": [5e7, 5e8, 1e9]}}
DEFAULT_TUNING_SPACE_ZERO_2 = {"zero_optimization": {"stage": 2, "overlap_comm": [True, False], "reduce_scatter": [False, True],
                                                     "reduce_bucket_size": [5e7, 5e8, 1e9], "allgather_bucket_size": [5e7, 5e8, 1e9],
                                                     "contiguous_gradients": [False, True]}}
DEFAULT_TUNING_SPACE_ZERO_3 = {"zero_optimization": {"stage": 3, "overlap_comm": [True, False], "fused_collectives": [True, False],
                                                     "stage3_prefetch_bucket_size": [5e7, 5e8, 1e9]}}
GLOBAL_TUNING_SPACE = "global"
TUNING_MICRO_BATCH_SIZE_PREFIX = "z"
