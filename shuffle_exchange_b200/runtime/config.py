"""Top-level JSON/HJSON config (reference: deepspeed/runtime/config.py:648-997 ``DeepSpeedConfig``;
key inventory in SURVEY.md §5.6).

Accepted inputs: a path to a JSON/HJSON file, a dict, or a base64-encoded JSON string.  Duplicate
keys are rejected.  The batch triad ``train_batch_size = micro_batch x grad_accum x dp_world`` is
resolved from any one or two of its members (:meth:`_resolve_batch_triad`).  Scalar sections are
declared in one table (``_SCALARS``) instead of one getter function per key.
"""
from __future__ import annotations

import base64
import copy
import json
import os
from enum import Enum
from typing import Any, Dict, List, Optional, Union

import hjson
import torch
from pydantic import Field

from ..comm import comm as dist
from ..comm.config import DeepSpeedCommsConfig
from ..utils.logging import logger
from . import constants as C
from .config_utils import (DeepSpeedConfigModel, ScientificNotationEncoder, dict_raise_error_on_duplicate_keys,
                           get_scalar_param)
from .zero.config import ZeroStageEnum, get_zero_config

ADAGRAD_OPTIMIZER = "adagrad"
ADAM_OPTIMIZER = "adam"
ADAMW_OPTIMIZER = "adamw"
LAMB_OPTIMIZER = "lamb"
ONEBIT_ADAM_OPTIMIZER = "onebitadam"
ZERO_ONE_ADAM_OPTIMIZER = "zerooneadam"
ONEBIT_LAMB_OPTIMIZER = "onebitlamb"
MUADAM_OPTIMIZER = "muadam"
MUADAMW_OPTIMIZER = "muadamw"
MUSGD_OPTIMIZER = "musgd"
LION_OPTIMIZER = "lion"
SGD_OPTIMIZER = "sgd"
DEEPSPEED_OPTIMIZERS = [ADAGRAD_OPTIMIZER, ADAM_OPTIMIZER, ADAMW_OPTIMIZER, LAMB_OPTIMIZER, ONEBIT_ADAM_OPTIMIZER,
                        ONEBIT_LAMB_OPTIMIZER, ZERO_ONE_ADAM_OPTIMIZER, MUADAM_OPTIMIZER, MUADAMW_OPTIMIZER,
                        MUSGD_OPTIMIZER, LION_OPTIMIZER, SGD_OPTIMIZER]
TORCH_ADAM_PARAM = "torch_adam"
ADAM_W_MODE = "adam_w_mode"
ADAM_W_MODE_DEFAULT = True


class DeepSpeedConfigError(Exception):
    pass


# ------------------------------------------------------------------------- structured blocks
class FP16Config(DeepSpeedConfigModel):
    enabled: bool = False
    auto_cast: bool = False
    loss_scale: float = Field(0, ge=0)            # 0 = dynamic
    initial_scale_power: int = Field(16, ge=0)
    loss_scale_window: int = Field(1000, ge=0)
    hysteresis: int = Field(2, ge=0)
    consecutive_hysteresis: bool = False
    min_loss_scale: float = Field(1, ge=0)
    fp16_master_weights_and_grads: bool = False

    def initial_dynamic_scale(self):
        return 2 ** self.initial_scale_power

    def dynamic_loss_scale_args(self):
        return {"init_scale": 2 ** self.initial_scale_power, "scale_window": self.loss_scale_window,
                "delayed_shift": self.hysteresis, "consecutive_hysteresis": self.consecutive_hysteresis,
                "min_scale": self.min_loss_scale}


class BF16Config(DeepSpeedConfigModel):
    enabled: bool = False
    immediate_grad_update: bool = True
    check_grad_overflow: bool = False


class TorchAutocastConfig(DeepSpeedConfigModel):
    enabled: bool = False
    dtype: str = "bfloat16"
    lower_precision_safe_modules: Optional[List[str]] = None


class ActivationCheckpointingConfig(DeepSpeedConfigModel):
    partition_activations: bool = False
    contiguous_memory_optimization: bool = False
    cpu_checkpointing: bool = False
    number_checkpoints: Optional[int] = None
    synchronize_checkpoint_boundary: bool = False
    profile: bool = False


class FlopsProfilerConfig(DeepSpeedConfigModel):
    enabled: bool = False
    recompute_fwd_factor: float = 0.0
    profile_step: int = Field(1, ge=1)
    module_depth: int = -1
    top_modules: int = Field(1, ge=1)
    detailed: bool = True
    output_file: Optional[str] = None


class TensorBoardConfig(DeepSpeedConfigModel):
    enabled: bool = False
    output_path: str = ""
    job_name: str = "DeepSpeedJobName"


class WandbConfig(DeepSpeedConfigModel):
    enabled: bool = False
    group: Optional[str] = None
    team: Optional[str] = None
    project: str = "deepspeed"


class CSVConfig(DeepSpeedConfigModel):
    enabled: bool = False
    output_path: str = ""
    job_name: str = "DeepSpeedJobName"


class CometConfig(DeepSpeedConfigModel):
    enabled: bool = False
    samples_log_interval: int = 100
    project: Optional[str] = None
    workspace: Optional[str] = None
    api_key: Optional[str] = None
    experiment_name: Optional[str] = None
    experiment_key: Optional[str] = None
    online: Optional[bool] = None
    mode: Optional[str] = None


class MonitorConfig(DeepSpeedConfigModel):
    tensorboard: TensorBoardConfig = Field(default_factory=TensorBoardConfig)
    comet: CometConfig = Field(default_factory=CometConfig)
    wandb: WandbConfig = Field(default_factory=WandbConfig)
    csv_monitor: CSVConfig = Field(default_factory=CSVConfig)

    @property
    def enabled(self):
        return self.tensorboard.enabled or self.wandb.enabled or self.csv_monitor.enabled or self.comet.enabled


class AioConfig(DeepSpeedConfigModel):
    block_size: int = 1048576
    queue_depth: int = 8
    intra_op_parallelism: int = 1
    single_submit: bool = False
    overlap_events: bool = True
    use_gds: bool = False


class PipelineConfig(DeepSpeedConfigModel):
    stages: Union[str, int] = "auto"
    partition: str = "best"
    seed_layers: bool = False
    activation_checkpoint_interval: int = 0
    pipe_partitioned: bool = True
    grad_partitioned: bool = True
    use_reentrant: bool = True


class CheckpointWriterConfig(DeepSpeedConfigModel):
    type: str = "python"             # mock | python | fast
    io_buffer_size: int = 1 << 30
    io_buffer_double: bool = True
    show_statistics: bool = False
    data_parallel: Optional[str] = None   # socket | machine | replica
    decoupled: bool = False
    io_multiplier: int = 1


class CheckpointConfig(DeepSpeedConfigModel):
    tag_validation: str = "Warn"     # Ignore | Warn | Fail
    load_universal: bool = False
    use_node_local_storage: bool = False
    parallel_write: Dict[str, Any] = Field(default_factory=dict)
    writer: Optional[CheckpointWriterConfig] = None


class DataTypesConfig(DeepSpeedConfigModel):
    grad_accum_dtype: Optional[str] = None


class TensorParallelConfig(DeepSpeedConfigModel):
    autotp_size: int = 0
    tp_size: int = 1
    tp_grain_size: int = 1
    tp_overlap_comm: bool = False
    mpu: Any = None
    tp_group: Any = None
    injection_policy: Any = None
    keep_module_on_host: bool = False
    replace_with_kernel_inject: bool = False


class CompileConfig(DeepSpeedConfigModel):
    deepcompile: bool = False
    free_activation: bool = False
    offload_activation: bool = False
    offload_opt_states: bool = False
    double_buffer: bool = True
    symmetric_memory: bool = False
    debug_log: bool = False
    offload_parameters: bool = False
    sync_before_reduce: bool = False
    sync_after_reduce: bool = False
    sync_before_allgather: bool = False
    sync_after_allgather: bool = False
    keep_int_input_tensors: bool = True
    keep_all_input_tensors: bool = False


class HybridEngineConfig(DeepSpeedConfigModel):
    enabled: bool = False
    max_out_tokens: int = 512
    inference_tp_size: int = 1
    release_inference_cache: bool = False
    pin_parameters: bool = True
    tp_gather_partition_size: int = 8


class EigenvalueConfig(DeepSpeedConfigModel):
    enabled: bool = False
    verbose: bool = False
    max_iter: int = 100
    tol: float = 1e-2
    stability: float = 1e-6
    gas_boundary_resolution: int = 1
    layer_name: str = "bert.encoder.layer"
    layer_num: int = 0


class PLDConfig(DeepSpeedConfigModel):
    enabled: bool = False
    theta: float = 1.0
    gamma: float = 0.001


class TimersConfig(DeepSpeedConfigModel):
    class Throughput(DeepSpeedConfigModel):
        enabled: bool = True
        synchronized: bool = True
    throughput: Throughput = Field(default_factory=Throughput)


# scalar top-level keys: attribute name -> (json key, default)
_SCALARS = C.SCALAR_KEYS          # attribute -> (key, default), declared next to the key names

_COMM_DTYPES = {None: None, "fp32": torch.float32, "fp16": torch.float16, "bf16": torch.bfloat16,
                "float32": torch.float32, "float16": torch.float16, "bfloat16": torch.bfloat16}


def _load_config_source(config) -> dict:
    if isinstance(config, dict):
        return copy.deepcopy(config)
    if isinstance(config, (str, os.PathLike)):
        if os.path.exists(config):
            with open(config, "r") as f:
                return hjson.load(f, object_pairs_hook=dict_raise_error_on_duplicate_keys)
        try:
            decoded = base64.urlsafe_b64decode(str(config)).decode("utf-8")
            return hjson.loads(decoded, object_pairs_hook=dict_raise_error_on_duplicate_keys)
        except Exception as e:
            raise DeepSpeedConfigError(
                f"Expected a string path to an existing deepspeed config, or a dictionary or a valid base64. "
                f"Received: {config}") from e
    raise DeepSpeedConfigError(f"Unsupported config type {type(config)}")


class DeepSpeedConfig:
    def __init__(self, config: Union[str, dict], mpu=None, mesh_device=None):
        self._param_dict = _load_config_source(config)
        try:
            self.global_rank = dist.get_rank()
            if mpu is not None:
                self.world_size = mpu.get_data_parallel_world_size()
            elif mesh_device is not None:
                self.world_size = dist.get_world_size(mesh_device.get_group(mesh_dim="data_parallel"))
            else:
                sp = self._param_dict.get("sequence_parallel_size", 1)
                self.world_size = max(1, dist.get_world_size() // max(1, sp))
        except Exception:
            self.global_rank, self.world_size = 0, 1
        tp = self._param_dict.get("tensor_parallel", {}).get("autotp_size", 0)
        if tp and tp > 1 and mpu is None:
            self.world_size = max(1, self.world_size // tp)
        logger.debug(f"Config mesh_device {mesh_device} world_size = {self.world_size}")

        self.elasticity_enabled = self._param_dict.get("elasticity", {}).get("enabled", False)
        if self.elasticity_enabled:
            self._apply_elasticity()

        self._initialize_params(copy.copy(self._param_dict))
        self._resolve_batch_triad()
        self._do_sanity_check()

    # ------------------------------------------------------------------------------ elasticity
    def _apply_elasticity(self):
        from ..elasticity import compute_elastic_config, elasticity_enabled, ensure_immutable_elastic_config
        from ..elasticity.config import ElasticityConfig, ElasticityConfigError
        from ..elasticity.constants import ELASTICITY, IGNORE_NON_ELASTIC_BATCH_INFO, IGNORE_NON_ELASTIC_BATCH_INFO_DEFAULT, MODEL_PARALLEL_SIZE, NUM_GPUS_PER_NODE
        from ..version import __version__
        pd = self._param_dict
        assert elasticity_enabled(pd)
        final_batch, valid_gpus, micro = compute_elastic_config(ds_config=pd, target_deepspeed_version=__version__,
                                                                world_size=self.world_size)
        ecfg = pd[ELASTICITY]
        ensure_immutable_elastic_config(runtime_elastic_config_dict=ecfg)
        self.elastic_model_parallel_size = ecfg.get(MODEL_PARALLEL_SIZE, 1)
        self.num_gpus_per_node = ecfg.get(NUM_GPUS_PER_NODE, 1)
        if not ecfg.get(IGNORE_NON_ELASTIC_BATCH_INFO, IGNORE_NON_ELASTIC_BATCH_INFO_DEFAULT):
            present = [k for k in (C.TRAIN_BATCH_SIZE, C.TRAIN_MICRO_BATCH_SIZE_PER_GPU, C.GRADIENT_ACCUMULATION_STEPS)
                       if k in pd]
            if present:
                raise ElasticityConfigError(
                    f"One or more batch related parameters were found in your ds_config ({present}). These parameters "
                    f"*will not be used* since elastic training is enabled, which takes control of these parameters. "
                    f"If you want to suppress this error set '{IGNORE_NON_ELASTIC_BATCH_INFO}': true in elasticity.")
        gas = final_batch // (micro * self.world_size)
        pd[C.TRAIN_BATCH_SIZE] = final_batch
        pd[C.TRAIN_MICRO_BATCH_SIZE_PER_GPU] = micro
        pd[C.GRADIENT_ACCUMULATION_STEPS] = gas
        logger.info(f"[Elasticity] valid GPU counts: {valid_gpus}; batch={final_batch} micro={micro} gas={gas}")

    # ------------------------------------------------------------------------------ parameters
    def _initialize_params(self, pd: dict):
        self.train_batch_size = pd.get(C.TRAIN_BATCH_SIZE)
        self.train_micro_batch_size_per_gpu = pd.get(C.TRAIN_MICRO_BATCH_SIZE_PER_GPU)
        self.gradient_accumulation_steps = pd.get(C.GRADIENT_ACCUMULATION_STEPS)
        for attr, (key, default) in _SCALARS.items():
            setattr(self, attr, get_scalar_param(pd, key, default))

        cdt = pd.get(C.COMMUNICATION_DATA_TYPE, C.COMMUNICATION_DATA_TYPE_DEFAULT)
        if cdt not in _COMM_DTYPES:
            raise ValueError(f"Invalid communication_data_type. Supported data types: ['fp16', 'bf16', 'fp32']. Got: {cdt}")
        self.communication_data_type = _COMM_DTYPES[cdt]
        self.seq_parallel_communication_data_type = _COMM_DTYPES.get(self.seq_parallel_communication_data_type,
                                                                     torch.float32)

        self.zero_config = get_zero_config(pd)
        self.mics_shard_size = self.zero_config.mics_shard_size
        self.mics_hierarchial_params_gather = self.zero_config.mics_hierarchical_params_gather
        self.zero_optimization_stage = int(self.zero_config.stage)
        self.zero_enabled = self.zero_optimization_stage > 0

        self.activation_checkpointing_config = ActivationCheckpointingConfig(**pd.get("activation_checkpointing", {}))
        self.comms_config = DeepSpeedCommsConfig(pd)
        mon = {k: pd[k] for k in ("tensorboard", "wandb", "csv_monitor", "comet") if k in pd}
        self.monitor_config = MonitorConfig(**mon)

        self.fp16_config = FP16Config(**pd.get(C.FP16, {}))
        bf = pd.get(C.BFLOAT16, pd.get(C.BFLOAT16_OLD, {}))
        self.bf16_config = BF16Config(**bf)
        assert not (self.fp16_config.enabled and self.bf16_config.enabled), "bfloat16 and fp16 modes cannot be simultaneously enabled"
        self.fp16_enabled = self.fp16_config.enabled
        self.fp16_auto_cast = self.fp16_config.auto_cast
        self.bfloat16_enabled = self.bf16_config.enabled
        self.bfloat16_immediate_grad_update = self.bf16_config.immediate_grad_update
        self.bfloat16_check_grad_overflow = self.bf16_config.check_grad_overflow
        self.fp16_master_weights_and_gradients = self.fp16_config.fp16_master_weights_and_grads
        self.loss_scale = self.fp16_config.loss_scale
        self.initial_dynamic_scale = self.fp16_config.initial_dynamic_scale()
        self.dynamic_loss_scale_args = self.fp16_config.dynamic_loss_scale_args()
        self.amp_enabled = pd.get(C.AMP, {}).get(C.AMP_ENABLED, C.AMP_ENABLED_DEFAULT)
        self.amp_params = {k: v for k, v in pd.get(C.AMP, {}).items() if k != C.AMP_ENABLED}
        self.torch_autocast_config = TorchAutocastConfig(**pd.get(C.TORCH_AUTOCAST, {}))
        self.torch_autocast_enabled = self.torch_autocast_config.enabled
        self.torch_autocast_dtype = getattr(torch, self.torch_autocast_config.dtype, torch.bfloat16)
        self.torch_autocast_lower_precision_safe_modules = self.torch_autocast_config.lower_precision_safe_modules

        self.compression_config = pd.get("compression_training", {})
        self.weight_quantization_config = pd.get("weight_quantization")

        opt = pd.get(C.OPTIMIZER)
        self.optimizer_name = opt.get(C.TYPE) if opt else C.OPTIMIZER_TYPE_DEFAULT
        if self.optimizer_name is not None and self.optimizer_name.lower() in DEEPSPEED_OPTIMIZERS:
            self.optimizer_name = self.optimizer_name.lower()
        self.optimizer_params = (opt or {}).get(C.OPTIMIZER_PARAMS)
        self.optimizer_legacy_fusion = (opt or {}).get(C.LEGACY_FUSION, C.LEGACY_FUSION_DEFAULT)
        sched = pd.get(C.SCHEDULER)
        self.scheduler_name = sched.get(C.TYPE) if sched else C.SCHEDULER_TYPE_DEFAULT
        self.scheduler_params = (sched or {}).get(C.SCHEDULER_PARAMS)

        self.flops_profiler_config = FlopsProfilerConfig(**pd.get("flops_profiler", {}))
        self.timers_config = TimersConfig(**pd.get("timers", {}))
        self.autotuning_config = pd.get("autotuning", {})
        self.eigenvalue_config = EigenvalueConfig(**pd.get(C.EIGENVALUE, {}))
        for k in ("enabled", "verbose", "max_iter", "tol", "stability", "gas_boundary_resolution", "layer_name",
                  "layer_num"):
            setattr(self, f"eigenvalue_{k}", getattr(self.eigenvalue_config, k))
        self.hybrid_engine = HybridEngineConfig(**pd.get("hybrid_engine", {}))
        self.sparse_attention = pd.get(C.SPARSE_ATTENTION)
        self.pipeline = PipelineConfig(**pd.get("pipeline", {})).model_dump()
        self.pld_config = PLDConfig(**pd.get(C.PROGRESSIVE_LAYER_DROP, {}))
        self.pld_enabled = self.pld_config.enabled
        self.pld_params = {C.PLD_THETA: self.pld_config.theta, C.PLD_GAMMA: self.pld_config.gamma} if self.pld_enabled else False
        self.curriculum_enabled_legacy = pd.get("curriculum_learning", {}).get("enabled", False)
        self.curriculum_params_legacy = pd.get("curriculum_learning", {}) if self.curriculum_enabled_legacy else False
        self.data_efficiency_config = pd.get("data_efficiency", {})
        self.data_efficiency_enabled = self.data_efficiency_config.get("enabled", False)

        self.checkpoint_config = CheckpointConfig(**pd.get(C.CHECKPOINT, {}))
        tv = self.checkpoint_config.tag_validation.upper()
        if tv not in C.CHECKPOINT_TAG_VALIDATION_MODES:
            raise DeepSpeedConfigError(f"Checkpoint config contains invalid tag_validation value of {tv}, expecting one of ['IGNORE', 'WARN', 'FAIL']")
        self.checkpoint_tag_validation_enabled = tv != C.ValidationMode.IGNORE
        self.checkpoint_tag_validation_fail = tv == C.ValidationMode.FAIL
        self.load_universal_checkpoint = self.checkpoint_config.load_universal
        self.use_node_local_storage = self.checkpoint_config.use_node_local_storage
        self.checkpoint_parallel_write_pipeline = bool(self.checkpoint_config.parallel_write.get(C.CHECKPOINT_PARALLEL_WRITE_PIPELINE_STAGE,
                                                                                     C.CHECKPOINT_PARALLEL_WRITE_PIPELINE_STAGE_DEFAULT))
        self.checkpoint_writer_config = self.checkpoint_config.writer

        self.data_types_config = DataTypesConfig(**pd.get(C.DATA_TYPES, {}))
        self.grad_accum_dtype = self.data_types_config.grad_accum_dtype
        self.aio_config = AioConfig(**pd.get("aio", {})).model_dump()
        self.nebula_config = pd.get("nebula", {})
        self.compile_config = CompileConfig(**pd.get("compile", {}))
        self.tensor_parallel_config = TensorParallelConfig(**pd.get("tensor_parallel", {}))
        self.sequence_parallel_size = pd.get("sequence_parallel_size", 1)
        self.elasticity_config = pd.get("elasticity", {})

    # ---------------------------------------------------------------------------- batch triad
    def _resolve_batch_triad(self):
        tb, mb, gas, ws = (self.train_batch_size, self.train_micro_batch_size_per_gpu,
                           self.gradient_accumulation_steps, self.world_size)
        if tb is not None and mb is not None and gas is not None:
            pass
        elif tb is not None and mb is not None:
            gas = tb // (mb * ws)
        elif tb is not None and gas is not None:
            mb = tb // (ws * gas)
        elif mb is not None and gas is not None:
            tb = mb * gas * ws
        elif tb is not None:
            gas = 1
            mb = tb // ws
        elif mb is not None:
            gas = 1
            tb = mb * ws
        else:
            raise DeepSpeedConfigError("Either train_batch_size or train_micro_batch_size_per_gpu needs to be provided")
        self.train_batch_size, self.train_micro_batch_size_per_gpu, self.gradient_accumulation_steps = tb, mb, gas
        assert tb > 0, f"Train batch size: {tb} has to be greater than 0"
        assert mb > 0, f"Micro batch size per gpu: {mb} has to be greater than 0"
        assert gas > 0, f"Gradient accumulation steps: {gas} has to be greater than 0"
        assert tb == mb * gas * ws, (
            f"Check batch related parameters. train_batch_size is not equal to micro_batch_per_gpu * gradient_acc_step "
            f"* world_size {tb} != {mb} * {gas} * {ws}")

    _configure_train_batch_size = _resolve_batch_triad

    # ------------------------------------------------------------------------------- sanity
    def _do_sanity_check(self):
        assert self.zero_optimization_stage <= int(ZeroStageEnum.max_stage), \
            f"DeepSpeedConfig: Maximum supported ZeRO stage is {int(ZeroStageEnum.max_stage)}"
        if self.fp16_master_weights_and_gradients:
            assert self.zero_enabled and self.zero_optimization_stage in (1, 2, 3), \
                "fp16_master_weights_and_grads is only supported with ZeRO"
        vs = self.vocabulary_size
        if vs and vs % C.TENSOR_CORE_ALIGN_SIZE != 0:
            logger.debug(f"DeepSpeedConfig: vocabulary size {vs} is not aligned to {C.TENSOR_CORE_ALIGN_SIZE}, "
                         "may impact tensor core utilization.")
        if self.optimizer_params is not None and C.MAX_GRAD_NORM in self.optimizer_params:
            if self.fp16_enabled or self.bfloat16_enabled:
                logger.warning("DeepSpeedConfig: In mixed-precision mode, max_grad_norm in the optimizer params is "
                               "superseded by the top-level gradient_clipping key; setting it to 0")
                self.optimizer_params["max_grad_norm"] = 0.0

    def print_user_config(self):
        logger.info("  json = {}".format(json.dumps(self._param_dict, sort_keys=True, indent=4,
                                                    cls=ScientificNotationEncoder, separators=(",", ":"))))

    def print(self, name):
        logger.info(f"{name}:")
        for arg in sorted(vars(self)):
            if arg != "_param_dict":
                logger.info(f"  {arg} {'.' * (29 - len(arg))} {getattr(self, arg)}")
        self.print_user_config()


# ================================================================================================================
# Module-level getters under the reference's names (``runtime/config.py:111-640``): each reads ONE setting straight
# from a raw config dictionary, applying the same default the engine uses.  Kept for scripts that import them.
class DtypeEnum(Enum):
    fp16 = (torch.float16, "torch.float16", "fp16", "float16", "half")
    fp32 = (torch.float32, "torch.float32", "fp32", "float32", "float")
    bf16 = (torch.bfloat16, "torch.bfloat16", "bf16", "bfloat16", "bfloat")
    int8 = (torch.int8, "torch.int8", "int8")

    def __new__(cls, *values):
        obj = object.__new__(cls)
        obj._value_ = values[0]
        for other in values[1:]:
            cls._value2member_map_[other] = obj
        obj._all_values = values
        return obj

    def __repr__(self):
        return f"<{type(self).__name__}.{self.name}: {', '.join(map(repr, self._all_values))}>"


def _sec(d, key):
    v = (d or {}).get(key, {})
    return v if isinstance(v, dict) else {}


def get_pld_enabled(d): return bool(_sec(d, "progressive_layer_drop").get("enabled", False))
def get_pld_params(d):
    if "progressive_layer_drop" not in (d or {}):
        return False
    p = dict(_sec(d, "progressive_layer_drop")); p.pop("enabled", None); return p
def get_amp_enabled(d): return bool(_sec(d, "amp").get("enabled", False))
def get_amp_params(d):
    if "amp" not in (d or {}):
        return False
    p = dict(_sec(d, "amp")); p.pop("enabled", None); return p
def get_torch_autocast_enabled(d): return bool(_sec(d, "torch_autocast").get("enabled", False))
def get_torch_autocast_dtype(d):
    return DtypeEnum(_sec(d, "torch_autocast").get("dtype", "bfloat16")).value
def get_lower_precision_safe_modules(d): return _sec(d, "torch_autocast").get("lower_precision_safe_modules")
def get_gradient_accumulation_steps(d): return (d or {}).get("gradient_accumulation_steps", None)
def get_sparse_gradients_enabled(d): return (d or {}).get("sparse_gradients", False)
def get_communication_data_type(d, key="communication_data_type"):
    v = (d or {}).get(key, None)
    if v is None:
        return None
    if v not in _COMM_DTYPES:
        raise ValueError(f"Invalid communication_data_type. Supported data types: ['fp16', 'bf16', 'fp32']. Got: {v}")
    return _COMM_DTYPES[v]
def get_prescale_gradients(d): return (d or {}).get("prescale_gradients", False)
def get_gradient_predivide_factor(d): return (d or {}).get("gradient_predivide_factor", 1.0)
def get_steps_per_print(d): return (d or {}).get("steps_per_print", 10)
def get_disable_allgather(d): return (d or {}).get("disable_allgather", False)
def get_dump_state(d): return (d or {}).get("dump_state", False)
def get_gradient_clipping(d): return (d or {}).get("gradient_clipping", 0.0)
def get_graph_harvesting(d): return (d or {}).get("graph_harvesting", False)
def get_train_batch_size(d): return (d or {}).get("train_batch_size", None)
def get_train_micro_batch_size_per_gpu(d): return (d or {}).get("train_micro_batch_size_per_gpu", None)
def get_wall_clock_breakdown(d): return (d or {}).get("wall_clock_breakdown", False)
def get_memory_breakdown(d): return (d or {}).get("memory_breakdown", False)
def get_dataloader_drop_last(d): return (d or {}).get("dataloader_drop_last", False)
def get_zero_allow_untested_optimizer(d): return (d or {}).get("zero_allow_untested_optimizer", False)
def get_zero_force_ds_cpu_optimizer(d): return (d or {}).get("zero_force_ds_cpu_optimizer", True)


def get_optimizer_name(d):
    o = _sec(d, "optimizer")
    return o.get("type") if "type" in o else None
def get_optimizer_params(d):
    return _sec(d, "optimizer").get("params") if get_optimizer_name(d) is not None else None
def get_optimizer_gradient_clipping(d):
    p = get_optimizer_params(d)
    return p.get("max_grad_norm") if p else None
def get_optimizer_legacy_fusion(d): return _sec(d, "optimizer").get("legacy_fusion", False)
def get_scheduler_name(d):
    s = _sec(d, "scheduler")
    return s.get("type") if "type" in s else None
def get_scheduler_params(d):
    return _sec(d, "scheduler").get("params") if get_scheduler_name(d) is not None else None


def get_pipeline_config(d):
    """``pipeline`` section with defaults filled."""
    return PipelineConfig(**_sec(d, "pipeline")).model_dump()
def get_hybrid_engine_config(d): return HybridEngineConfig(**_sec(d, "hybrid_engine"))
def get_expert_data_topo_config(d): return (d or {}).get("use_data_before_expert_parallelism", False)
def get_checkpoint_params(d): return _sec(d, "checkpoint")
def get_data_types_params(d): return _sec(d, "data_types")
def get_checkpoint_tag_validation_mode(checkpoint_params):
    mode = str(checkpoint_params.get("tag_validation", "Warn")).upper()
    if mode not in ("IGNORE", "WARN", "FAIL"):
        raise DeepSpeedConfigError(f"Checkpoint config contains invalid tag_validation value of {mode}, expecting one of "
                                   f"['IGNORE', 'WARN', 'FAIL']")
    return mode
def get_checkpoint_parallel_write_pipeline(checkpoint_params):
    v = checkpoint_params.get("parallel_write", {}).get("pipeline_stage", False)
    if v not in (True, False):
        raise DeepSpeedConfigError(f"checkpoint::parallel_write::pipeline_stage value of '{v}' is invalid, "
                                   f"expecting: true or false")
    return v


def get_eigenvalue_config(d):
    if get_eigenvalue_enabled(d):
        return (True, get_eigenvalue_verbose(d), get_eigenvalue_max_iter(d), get_eigenvalue_tol(d),
                get_eigenvalue_stability(d), get_eigenvalue_gas_boundary_resolution(d), get_eigenvalue_layer_name(d),
                get_eigenvalue_layer_num(d))
    e = EigenvalueConfig()
    return (False, e.verbose, e.max_iter, e.tol, e.stability, e.gas_boundary_resolution, e.layer_name, e.layer_num)
def get_eigenvalue_enabled(d): return bool(_sec(d, "eigenvalue").get("enabled", False))
def _eig(d, k): return _sec(d, "eigenvalue").get(k, getattr(EigenvalueConfig(), k))
def get_eigenvalue_verbose(d): return _eig(d, "verbose")
def get_eigenvalue_max_iter(d): return _eig(d, "max_iter")
def get_eigenvalue_tol(d): return _eig(d, "tol")
def get_eigenvalue_stability(d): return _eig(d, "stability")
def get_eigenvalue_gas_boundary_resolution(d): return _eig(d, "gas_boundary_resolution")
def get_eigenvalue_layer_name(d): return _eig(d, "layer_name")
def get_eigenvalue_layer_num(d): return _eig(d, "layer_num")


# ---- "sparse_attention" section: mode + that mode's layout parameters with defaults
_SPARSE_COMMON = {"block": 16, "different_layout_per_head": False}
_SPARSE_MODES = {
    "dense": {},
    "fixed": {"num_local_blocks": 4, "num_global_blocks": 1, "attention": "bidirectional",
              "horizontal_global_attention": False, "num_different_global_patterns": 1},
    "variable": {"num_random_blocks": 0, "local_window_blocks": [4], "global_block_indices": [0],
                 "global_block_end_indices": None, "attention": "bidirectional", "horizontal_global_attention": False},
    "bigbird": {"num_random_blocks": 1, "num_sliding_window_blocks": 3, "num_global_blocks": 1},
    "bslongformer": {"num_sliding_window_blocks": 3, "global_block_indices": [0], "global_block_end_indices": None},
}


def get_sparse_attention_mode(sparsity): return sparsity.get("mode", "fixed")
def get_sparse_attention_type(d): return _sec(d, "sparse_attention").get("attention", "bidirectional")
def _sparse_cfg(sparsity, mode):
    out = {"mode": mode, **{k: sparsity.get(k, v) for k, v in _SPARSE_COMMON.items()}}
    out.update({k: sparsity.get(k, v) for k, v in _SPARSE_MODES[mode].items()})
    return out
def get_sparse_dense_config(sparsity): return _sparse_cfg(sparsity, "dense")
def get_sparse_fixed_config(sparsity): return _sparse_cfg(sparsity, "fixed")
def get_sparse_variable_config(sparsity): return _sparse_cfg(sparsity, "variable")
def get_sparse_bigbird_config(sparsity): return _sparse_cfg(sparsity, "bigbird")
def get_sparse_bslongformer_config(sparsity): return _sparse_cfg(sparsity, "bslongformer")
def get_sparse_attention(d):
    if "sparse_attention" not in (d or {}):
        return None
    sparsity = d["sparse_attention"]
    mode = get_sparse_attention_mode(sparsity)
    if mode not in _SPARSE_MODES:
        raise NotImplementedError(f"Given sparsity mode, {mode}, has not been implemented yet!")
    return _sparse_cfg(sparsity, mode)


class DeepSpeedConfigWriter:
    """Build / load / save a config dictionary (reference ``runtime/config.py:643-660``)."""

    def __init__(self, data=None):
        self.data = data if data is not None else {}

    def add_config(self, key, value):
        self.data[key] = value

    def load_config(self, filename):
        with open(filename, "r") as f:
            self.data = json.load(f, object_pairs_hook=dict_raise_error_on_duplicate_keys)

    def write_config(self, filename):
        with open(filename, "w") as outfile:
            json.dump(self.data, outfile)
