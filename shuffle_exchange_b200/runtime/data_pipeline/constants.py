"""JSON key names of the ``data_efficiency`` config block (data sampling / curriculum learning / dynamic
batching / random layer-token drop).  Key strings match the reference's
(deepspeed/runtime/data_pipeline/constants.py) because user configs carry them."""

# -- top level ---------------------------------------------------------------------------------------
DATA_EFFICIENCY, DATA_SAMPLING, DATA_ROUTING = "data_efficiency", "data_sampling", "data_routing"
DATA_EFFICIENCY_ENABLED = DATA_SAMPLING_ENABLED = DATA_ROUTING_ENABLED = "enabled"
DATA_EFFICIENCY_ENABLED_DEFAULT = DATA_SAMPLING_ENABLED_DEFAULT = DATA_ROUTING_ENABLED_DEFAULT = False
DATA_EFFICIENCY_SEED, DATA_EFFICIENCY_SEED_DEFAULT = "seed", 1234
DATA_SAMPLING_NUM_EPOCHS, DATA_SAMPLING_NUM_EPOCHS_DEFAULT = "num_epochs", 1000
DATA_SAMPLING_NUM_WORKERS, DATA_SAMPLING_NUM_WORKERS_DEFAULT = "num_workers", 0
DATA_SAMPLING_PIN_MEMORY, DATA_SAMPLING_PIN_MEMORY_DEFAULT = "pin_memory", False

# -- curriculum learning ---------------------------------------------------------------------------------
CURRICULUM_LEARNING = CURRICULUM_LEARNING_LEGACY = "curriculum_learning"
CURRICULUM_LEARNING_ENABLED = CURRICULUM_ENABLED_LEGACY = "enabled"
CURRICULUM_LEARNING_ENABLED_DEFAULT = CURRICULUM_ENABLED_DEFAULT_LEGACY = False
(CURRICULUM_LEARNING_CLUSTER_PATH, CURRICULUM_LEARNING_METRICS, CURRICULUM_LEARNING_SAMPLE_PATH,
 CURRICULUM_LEARNING_METRIC_PATH, CURRICULUM_LEARNING_CLUSTERING_TYPE, CURRICULUM_LEARNING_SINGLE_CLUSTER,
 CURRICULUM_LEARNING_CLUSTER_PREFIX, CURRICULUM_LEARNING_DIFFICULTY_TYPE) = (
    "data_cluster_path", "curriculum_metrics", "index_to_sample_path", "index_to_metric_path", "clustering_type",
    "single_cluster", "cluster", "difficulty_type")
CURRICULUM_LEARNING_VALUE_BASED, CURRICULUM_LEARNING_PERCENTILE_BASED = "value", "percentile"
CURRICULUM_LEARNING_MIN_DIFFICULTY, CURRICULUM_LEARNING_MAX_DIFFICULTY = "min_difficulty", "max_difficulty"
CURRICULUM_LEARNING_SCHEDULE_TYPE, CURRICULUM_LEARNING_SCHEDULE_CONFIG = "schedule_type", "schedule_config"
(CURRICULUM_LEARNING_SCHEDULE_DIFFICULTY, CURRICULUM_LEARNING_SCHEDULE_MAX_STEP,
 CURRICULUM_LEARNING_SCHEDULE_TOTAL_STEP, CURRICULUM_LEARNING_SCHEDULE_DIFFICULTY_STEP,
 CURRICULUM_LEARNING_SCHEDULE_ROOT_DEGREE) = ("difficulty", "max_step", "total_curriculum_step", "difficulty_step",
                                              "root_degree")
(CURRICULUM_LEARNING_SCHEDULE_FIXED_DISCRETE, CURRICULUM_LEARNING_SCHEDULE_FIXED_ROOT,
 CURRICULUM_LEARNING_SCHEDULE_FIXED_LINEAR, CURRICULUM_LEARNING_SCHEDULE_CUSTOM) = ("fixed_discrete", "fixed_root",
                                                                                    "fixed_linear", "custom")
# sampler state-dict keys
(CURRICULUM_LEARNING_CURRENT_DIFFICULTY, CURRICULUM_LEARNING_BATCH, CURRICULUM_LEARNING_CONSUMED_SAMPLES,
 CURRICULUM_LEARNING_STEP, CURRICULUM_LEARNING_CURRENT_DIFFICULTIES, CURRICULUM_LEARNING_DATA_CLUSTER_PATHS,
 CURRICULUM_LEARNING_DATA_CLUSTER_CURRENT_POSITION, CURRICULUM_LEARNING_NP_RNG_STATE) = (
    "current_difficulty", "batch", "consumed_samples", "curriculum_step", "current_difficulties", "data_cluster_paths",
    "data_cluster_current_position", "np_rng_state")

# -- dynamic batching -------------------------------------------------------------------------------------
DYNAMIC_BATCHING, DYNAMIC_BATCHING_ENABLED, DYNAMIC_BATCHING_ENABLED_DEFAULT = "dynamic_batching", "enabled", False
DYNAMIC_BATCHING_METRICS_PATH = "metrics_path"
DYNAMIC_BATCHING_LR_SCALING_METHOD, DYNAMIC_BATCHING_LR_SCALING_METHOD_DEFAULT = "lr_scaling_method", "linear"
DYNAMIC_BATCHING_MIN_BATCH_SIZE, DYNAMIC_BATCHING_MIN_BATCH_SIZE_DEFAULT = "min_batch_size", 1
DYNAMIC_BATCHING_MAX_BATCH_SIZE, DYNAMIC_BATCHING_MAX_BATCH_SIZE_DEFAULT = "max_batch_size", None
DYNAMIC_BATCHING_SEQUENCE_PICKING_ORDER, DYNAMIC_BATCHING_SEQUENCE_PICKING_ORDER_DEFAULT = \
    "sequence_picking_order", "dataloader"
DYNAMIC_BATCHING_MAX_TOKENS, DYNAMIC_BATCHING_VERBOSE = "max_tokens", "verbose"

# -- random layer-token drop ---------------------------------------------------------------------------
RANDOM_LTD, RANDOM_LTD_ENABLED, RANDOM_LTD_ENABLED_DEFAULT = "random_ltd", "enabled", False
(RANDOM_LTD_MODEL_MASK_NAME, RANDOM_LTD_MODEL_TYPE, RANDOM_LTD_MICRO_BATCH_SIZE, RANDOM_LTD_GLOBAL_BATCH_SIZE,
 RANDOM_LTD_SAMPLE_INDEX, RANDOM_LTD_ATTENTION_MASK, RANDOM_LTD_HIDDEN_STATE_ORDER, RANDOM_LTD_LAYER_NUM,
 RANDOM_LTD_LAYER_ID, RANDOM_LTD_TOTAL_LAYER_NUM, RANDOM_LTD_CONSUMED_LAYER_TOKENS) = (
    "model_mask_name", "model_type", "micro_batch_size", "global_batch_size", "sample_idx", "attention_mask",
    "hidden_state_order", "random_ltd_layer_num", "random_ltd_layer_id", "total_layer_num", "consumed_layer_tokens")
(RANDOM_LTD_SCHEDULER, RANDOM_LTD_MAX_VALUE, RANDOM_LTD_MIN_VALUE, RANDOM_LTD_CURRENT_VALUE,
 RANDOM_LTD_SCHEDULE_CONFIG, RANDOM_LTD_INCREASE_STEP, RANDOM_LTD_REQUIRE_STEP, RANDOM_LTD_SCHEDULER_TYPE,
 RANDOM_LTD_CURR_STEP) = ("random_ltd_schedule", "max_value", "min_value", "current_value", "schedule_config",
                          "seq_per_step", "require_steps", "schedule_type", "current_steps")
RANDOM_LTD_LAYER_TOKEN_LR_SCHEDULE, RANDOM_LTD_LAYER_TOKEN_LR_ENABLED = "layer_token_lr_schedule", "enabled"
RANDOM_LTD_LAYER_TOKEN_LR_ENABLED_DEFAULT = False
RANDOM_LTD_TOTAL_LAYER_TOKENS, RANDOM_LTD_WARMUP_TYPE, RANDOM_LTD_WARMUP_LAYER_TOKENS = \
    "total_layer_tokens", "warmup_type", "warmup_layer_tokens"
