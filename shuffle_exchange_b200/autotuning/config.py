"""``autotuning`` config block (reference: ``deepspeed/autotuning/config.py:11-96``)."""
from .constants import *  # noqa: F401,F403


class DeepSpeedAutotuningConfig:
    def __init__(self, param_dict):
        block = param_dict.get(AUTOTUNING, {}) if param_dict else {}
        g = block.get
        self.enabled = g(AUTOTUNING_ENABLED, AUTOTUNING_ENABLED_DEFAULT)
        self.fast = g(AUTOTUNING_FAST, AUTOTUNING_FAST_DEFAULT)
        self.results_dir = g(AUTOTUNING_RESULTS_DIR, AUTOTUNING_RESULTS_DIR_DEFAULT)
        assert self.results_dir, "results_dir cannot be empty"
        self.exps_dir = g(AUTOTUNING_EXPS_DIR, AUTOTUNING_EXPS_DIR_DEFAULT)
        assert self.exps_dir, "exps_dir cannot be empty"
        self.overwrite = g(AUTOTUNING_OVERWRITE, AUTOTUNING_OVERWRITE_DEFAULT)
        self.start_profile_step = g(AUTOTUNING_START_PROFILE_STEP, AUTOTUNING_START_PROFILE_STEP_DEFAULT)
        self.end_profile_step = g(AUTOTUNING_END_PROFILE_STEP, AUTOTUNING_END_PROFILE_STEP_DEFAULT)
        self.metric = g(AUTOTUNING_METRIC, AUTOTUNING_METRIC_DEFAULT)
        self.metric_path = g(AUTOTUNING_METRIC_PATH, AUTOTUNING_METRIC_PATH_DEFAULT)
        self.tuner_type = g(AUTOTUNING_TUNER_TYPE, AUTOTUNING_TUNER_TYPE_DEFAULT)
        self.tuner_early_stopping = g(AUTOTUNING_TUNER_EARLY_STOPPING, AUTOTUNING_TUNER_EARLY_STOPPING_DEFAULT)
        self.tuner_num_trials = g(AUTOTUNING_TUNER_NUM_TRIALS, AUTOTUNING_TUNER_NUM_TRIALS_DEFAULT)
        self.arg_mappings = g(AUTOTUNING_ARG_MAPPINGS, AUTOTUNING_ARG_MAPPINGS_DEFAULT)
        self.model_info = get_model_info_config(block)
        self.model_info_path = g(AUTOTUNING_MODEL_INFO_PATH, AUTOTUNING_MODEL_INFO_PATH_DEFAULT)
        self.mp_size = g(AUTOTUNING_MP_SIZE, AUTOTUNING_MP_SIZE_DEFAULT)
        self.max_train_batch_size = g(AUTOTUNING_MAX_TRAIN_BATCH_SIZE, AUTOTUNING_MAX_TRAIN_BATCH_SIZE_DEFAULT)
        self.min_train_batch_size = g(AUTOTUNING_MIN_TRAIN_BATCH_SIZE, AUTOTUNING_MIN_TRAIN_BATCH_SIZE_DEFAULT)
        self.max_train_micro_batch_size_per_gpu = g(AUTOTUNING_MAX_TRAIN_MICRO_BATCH_SIZE_PER_GPU,
                                                    AUTOTUNING_MAX_TRAIN_MICRO_BATCH_SIZE_PER_GPU_DEFAULT)
        self.min_train_micro_batch_size_per_gpu = g(AUTOTUNING_MIN_TRAIN_MICRO_BATCH_SIZE_PER_GPU,
                                                    AUTOTUNING_MIN_TRAIN_MICRO_BATCH_SIZE_PER_GPU_DEFAULT)
        self.num_tuning_micro_batch_sizes = g(AUTOTUNING_NUM_TUNING_MICRO_BATCH_SIZES,
                                              AUTOTUNING_NUM_TUNING_MICRO_BATCH_SIZES_DEFAULT)


def get_model_info_config(param_dict):
    if MODEL_INFO in param_dict and param_dict[MODEL_INFO] is not None:
        info = {MODEL_INFO_PROFILE: MODEL_INFO_PROFILE_DEFAULT, MODEL_INFO_NUM_PARAMS: MODEL_INFO_NUM_PARAMS_DEFAULT}
        info.update(param_dict[MODEL_INFO])
        return info
    return None


def get_default_model_info_config():
    return {MODEL_INFO_PROFILE: MODEL_INFO_PROFILE_DEFAULT, MODEL_INFO_NUM_PARAMS: MODEL_INFO_NUM_PARAMS_DEFAULT}
