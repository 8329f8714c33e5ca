"""shuffle_exchange_b200 — a B200-native (sm_100a) large-model training engine with the capabilities,
``initialize()`` / ``engine.step()`` API, config surface and checkpoint layout of the
vituslcx/Shuffle-exchange DeepSpeed fork.  Usage mirrors the reference::

    import shuffle_exchange_b200 as deepspeed
    engine, optimizer, loader, sched = deepspeed.initialize(model=model, config=cfg,
                                                            model_parameters=model.parameters(),
                                                            method="shuffle", rings=4, shuffle_step=50, slice_count=2)
    loss = engine(batch); engine.backward(loss); engine.step()
    engine.optimizer.shuffle_exchange()          # the fork's per-batch call

Public surface (reference: deepspeed/__init__.py:69-374): ``initialize``, ``init_inference``,
``add_config_arguments``, ``default_inference_config``, ``init_distributed``, ``zero`` (``Init``,
``GatheredParameters``, ...), ``PipelineModule`` / ``LayerSpec`` / ``TiedLayerSpec``, ``checkpointing``,
``moe``, ``sequence``, ``DeepSpeedTransformerLayer`` / ``DeepSpeedTransformerConfig``, ``OnDevice``,
``log_dist``, ``logger``, ``__version__``.
"""
from __future__ import annotations

import argparse
import os
import sys
from typing import Optional, Union

import torch
from torch.optim import Optimizer
from torch.optim.lr_scheduler import _LRScheduler

from .version import __git_branch__, __git_hash__, __version__
from . import comm  # noqa: F401
from .comm import init_distributed  # noqa: F401
from .device import get_accelerator  # noqa: F401
from .utils.logging import log_dist, logger  # noqa: F401
from .runtime.config import DeepSpeedConfig, DeepSpeedConfigError  # noqa: F401
from .runtime.activation_checkpointing import checkpointing  # noqa: F401
from .runtime import zero  # noqa: F401
from .runtime.engine import DeepSpeedEngine, DeepSpeedOptimizerCallable, DeepSpeedSchedulerCallable  # noqa: F401
from .runtime.lr_schedules import add_tuning_arguments  # noqa: F401
from .utils import groups  # noqa: F401

dist = None

__version_major__, __version_minor__, __version_patch__ = (int(x) for x in __version__.split(".")[:3])


def _lazy(name):
    import importlib
    return importlib.import_module(name, __name__)


def __getattr__(name):
    """Heavy / optional sub-packages are imported on first use."""
    table = {
        "PipelineModule": (".runtime.pipe.module", "PipelineModule"),
        "LayerSpec": (".runtime.pipe.module", "LayerSpec"),
        "TiedLayerSpec": (".runtime.pipe.module", "TiedLayerSpec"),
        "PipelineEngine": (".runtime.pipe.engine", "PipelineEngine"),
        "DeepSpeedHybridEngine": (".runtime.hybrid_engine", "DeepSpeedHybridEngine"),
        "InferenceEngine": (".inference.engine", "InferenceEngine"),
        "DeepSpeedInferenceConfig": (".inference.config", "DeepSpeedInferenceConfig"),
        "DeepSpeedTransformerLayer": (".ops.transformer.transformer", "DeepSpeedTransformerLayer"),
        "DeepSpeedTransformerConfig": (".ops.transformer.transformer", "DeepSpeedTransformerConfig"),
        "OnDevice": (".utils.init_on_device", "OnDevice"),
        "DeepSpeedOptimizer": (".runtime.base_optimizer", "DeepSpeedOptimizer"),
        "ZeROOptimizer": (".runtime.base_optimizer", "ZeROOptimizer"),
        "is_compile_supported": (".runtime.compiler", "is_compile_supported"),
    }
    if name in table:
        mod, attr = table[name]
        return getattr(_lazy(mod), attr)
    if name in ("moe", "sequence", "pipe", "ops", "module_inject", "inference", "linear", "profiling", "elasticity",
                "autotuning", "compression", "checkpoint", "monitor", "launcher", "models", "parallel", "nvme", "io"):
        return _lazy("." + name)
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")


def initialize(args=None, model: torch.nn.Module = None,
               optimizer: Optional[Union[Optimizer, DeepSpeedOptimizerCallable]] = None,
               model_parameters=None, training_data=None,
               lr_scheduler: Optional[Union[_LRScheduler, DeepSpeedSchedulerCallable]] = None,
               distributed_port: int = 29500, mpu=None, dist_init_required: Optional[bool] = None, collate_fn=None,
               config=None, mesh_param=None, config_params=None,
               shuffle_step=50, rings=8, method="RR", slice_count=2):
    """Build the engine.  Returns ``(engine, optimizer, training_dataloader, lr_scheduler)``.

    The last four keyword arguments are the fork's decentralised-averaging knobs (reference:
    deepspeed/__init__.py:82-85); they may equivalently be given in the config under
    ``zero_optimization.replica_averaging``.
    """
    log_dist(f"sxb200 info: version={__version__}, git-hash={__git_hash__}, git-branch={__git_branch__}", ranks=[0])
    zero.partition_parameters.shutdown_init_context()
    assert model is not None, "deepspeed.initialize requires a model"
    global dist
    from . import comm as dist  # noqa: F811
    dist.init_distributed(dist_backend=get_accelerator().communication_backend_name(),
                          distributed_port=distributed_port, dist_init_required=dist_init_required)
    if config is None:
        config = config_params
    mesh_device = None
    if mesh_param:
        logger.info(f"mesh_param to Initialize mesh device: {mesh_param}")
        mesh_device = dist.initialize_mesh_device(mesh_param, ("data_parallel", "sequence_parallel"))
    elif config is not None:
        probe = config
        if isinstance(config, (str, os.PathLike)) or isinstance(config, dict):
            try:
                from .runtime.config import _load_config_source
                probe = _load_config_source(config)
            except Exception:
                probe = {}
        sp = probe.get("sequence_parallel_size", 1) if isinstance(probe, dict) else 1
        if sp > 1 and mpu is None:
            groups._create_sequence_parallel_groups(sp)
    if hasattr(args, "deepscale_config") and args.deepscale_config is not None:
        logger.warning("************ --deepscale_config is deprecated, please use --deepspeed_config ************")
        if hasattr(args, "deepspeed_config"):
            assert args.deepspeed_config is None, "Not sure how to proceed, we were given both a deepscale_config and deepspeed_config"
        args.deepspeed_config = args.deepscale_config
        args.deepscale_config = None
    if hasattr(args, "deepspeed_config") and args.deepspeed_config is not None:
        assert config is None, "Not sure how to proceed, we were given deepspeed configs in the deepspeed arguments and deepspeed.initialize() function call"
        config = args.deepspeed_config
    assert config is not None, "DeepSpeed requires --deepspeed_config to specify configuration file"

    from .runtime.pipe.module import PipelineModule
    if not isinstance(model, PipelineModule):
        config_class = DeepSpeedConfig(config, mpu, mesh_device=mesh_device)
        tp_size = config_class.tensor_parallel_config.autotp_size
        if tp_size and tp_size > 1 and mpu is None:
            from .module_inject.layers import configure_tensor_parallel_runtime
            from .runtime.tensor_parallel import TpTrainingManager
            groups._init_tp_mesh_device(tp_size)
            configure_tensor_parallel_runtime(config_class.tensor_parallel_config)
            TpTrainingManager(model=model, tp_size=tp_size, dtype=torch.bfloat16 if config_class.bfloat16_enabled else torch.float16)
        if config_class.hybrid_engine.enabled:
            from .runtime.hybrid_engine import DeepSpeedHybridEngine
            engine_cls = DeepSpeedHybridEngine
        else:
            engine_cls = DeepSpeedEngine
        engine = engine_cls(args=args, model=model, optimizer=optimizer, model_parameters=model_parameters,
                            training_data=training_data, lr_scheduler=lr_scheduler, mpu=mpu,
                            dist_init_required=dist_init_required, collate_fn=collate_fn, config=config,
                            mesh_device=mesh_device, config_class=config_class,
                            rings=rings, shuffle_step=shuffle_step, method=method, slice_count=slice_count)
    else:
        assert mpu is None, "mpu must be None with pipeline parallelism"
        mpu = model.mpu()
        config_class = DeepSpeedConfig(config, mpu)
        from .runtime.pipe.engine import PipelineEngine
        engine = PipelineEngine(args=args, model=model, optimizer=optimizer, model_parameters=model_parameters,
                                training_data=training_data, lr_scheduler=lr_scheduler, mpu=mpu,
                                dist_init_required=dist_init_required, collate_fn=collate_fn, config=config,
                                config_class=config_class)
    zero.partition_parameters.restore_init_context()
    return engine, engine.optimizer, engine.training_dataloader, engine.lr_scheduler


def _add_core_arguments(parser):
    group = parser.add_argument_group("DeepSpeed", "DeepSpeed configurations")
    group.add_argument("--deepspeed", default=False, action="store_true",
                       help="Enable DeepSpeed (helper flag for user code, no impact on DeepSpeed backend)")
    group.add_argument("--deepspeed_config", default=None, type=str, help="DeepSpeed json configuration file.")
    group.add_argument("--deepscale", default=False, action="store_true",
                       help="Deprecated enable DeepSpeed (helper flag for user code, no impact on DeepSpeed backend)")
    group.add_argument("--deepscale_config", default=None, type=str, help="Deprecated DeepSpeed json configuration file.")
    return parser


def add_config_arguments(parser):
    """Add ``--deepspeed`` / ``--deepspeed_config`` (and deprecated aliases) to an argparse parser."""
    return _add_core_arguments(parser)


def default_inference_config():
    from .inference.config import DeepSpeedInferenceConfig
    return DeepSpeedInferenceConfig().model_dump()


def init_inference(model, config=None, **kwargs):
    """Wrap ``model`` for inference (reference: deepspeed/__init__.py:299-374): ``config`` (path or dict)
    and kwargs are merged; conflicting duplicates raise."""
    from .inference.config import DeepSpeedInferenceConfig
    from .inference.engine import InferenceEngine
    log_dist(f"sxb200 info: version={__version__}, git-hash={__git_hash__}, git-branch={__git_branch__}", ranks=[0])
    if config is None:
        config = {}
    if isinstance(config, str):
        import json
        with open(config, "r") as f:
            config_dict = json.load(f)
    elif isinstance(config, dict):
        config_dict = config
    else:
        raise ValueError(f"'config' argument expected string or dictionary, got {type(config)}")
    overlap = set(config_dict.keys()).intersection(kwargs.keys())
    for key in overlap:
        if config_dict[key] != kwargs[key]:
            raise ValueError(f"Conflicting argument '{key}' in 'config':{config_dict[key]} and kwargs:{kwargs[key]}")
    config_dict.update(kwargs)
    return InferenceEngine(model, config=DeepSpeedInferenceConfig(**config_dict))
