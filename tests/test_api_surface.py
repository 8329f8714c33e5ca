"""The compatibility checklist of SURVEY.md §9: names a user of the reference imports or calls must exist here with the
same spelling (signatures of the entry points are checked where the survey pins them)."""
import inspect

import shuffle_exchange_b200 as sxb


def test_module_level_names():
    for name in ("initialize", "init_inference", "add_config_arguments", "default_inference_config", "init_distributed",
                 "zero", "checkpointing", "PipelineModule", "LayerSpec", "TiedLayerSpec", "DeepSpeedTransformerLayer",
                 "DeepSpeedTransformerConfig", "OnDevice", "log_dist", "logger", "__version__", "__git_hash__"):
        assert hasattr(sxb, name), name
    sig = inspect.signature(sxb.initialize).parameters
    for kw in ("args", "model", "optimizer", "model_parameters", "training_data", "lr_scheduler", "distributed_port",
               "mpu", "dist_init_required", "collate_fn", "config", "mesh_param", "config_params", "shuffle_step",
               "rings", "method", "slice_count"):
        assert kw in sig, kw
    for name in ("Init", "GatheredParameters", "register_external_parameter", "unregister_external_parameter",
                 "TiledLinear", "MiCS_Init", "estimate_zero2_model_states_mem_needs_all_live",
                 "estimate_zero2_model_states_mem_needs_all_cold", "estimate_zero3_model_states_mem_needs_all_live",
                 "estimate_zero3_model_states_mem_needs_all_cold"):
        assert hasattr(sxb.zero, name), name
    for name in ("checkpoint", "configure", "get_cuda_rng_tracker", "model_parallel_cuda_manual_seed", "reset",
                 "is_configured"):
        assert hasattr(sxb.checkpointing, name), name
    from shuffle_exchange_b200.moe.layer import MoE  # noqa: F401
    from shuffle_exchange_b200.sequence.layer import DistributedAttention  # noqa: F401
    from shuffle_exchange_b200.pipe import LayerSpec, PipelineModule, TiedLayerSpec  # noqa: F401


ENGINE_METHODS = """forward backward step train eval zero_grad no_sync is_gradient_accumulation_boundary
set_gradient_accumulation_boundary was_step_applied get_lr get_mom get_type get_global_grad_norm get_batch_info
set_train_batch_size set_train_micro_batch_size train_batch_size train_micro_batch_size_per_gpu
gradient_accumulation_steps deepspeed_io save_checkpoint load_checkpoint save_16bit_model save_fp16_model
module_state_dict load_module_state_dict empty_partition_cache compile offload_states reload_states destroy
allreduce_gradients""".split()


def test_engine_methods_and_signatures():
    from shuffle_exchange_b200.runtime.engine import DeepSpeedEngine
    for m in ENGINE_METHODS:
        assert callable(getattr(DeepSpeedEngine, m, None)), m
    p = inspect.signature(DeepSpeedEngine.backward).parameters
    assert "retain_graph" in p and "scale_wrt_gas" in p
    p = inspect.signature(DeepSpeedEngine.save_checkpoint).parameters
    assert list(p)[1:] == ["save_dir", "tag", "client_state", "save_latest", "exclude_frozen_parameters"]
    p = inspect.signature(DeepSpeedEngine.load_checkpoint).parameters
    assert list(p)[1:] == ["load_dir", "tag", "load_module_strict", "load_optimizer_states",
                           "load_lr_scheduler_states", "load_module_only", "custom_load_fn"]


def test_pipeline_engine_and_zero12_optimizer_surface():
    from shuffle_exchange_b200.runtime.pipe.engine import PipelineEngine
    for m in ("train_batch", "eval_batch", "set_dataloader", "set_dataiterator", "set_batch_fn", "is_first_stage",
              "is_last_stage", "get_additional_losses", "forward", "backward", "step"):
        assert callable(getattr(PipelineEngine, m, None)), m
    p = inspect.signature(PipelineEngine.eval_batch).parameters
    for kw in ("data_iter", "return_logits", "compute_loss", "reduce_output", "bcast_loss", "num_micro_batches"):
        assert kw in p, kw
    from shuffle_exchange_b200.runtime.zero.stage_1_and_2 import DeepSpeedZeroOptimizer
    for m in ("backward", "step", "zero_grad", "state_dict", "load_state_dict", "override_loss_scale", "set_lr", "get_lr",
              "update_lp_params", "refresh_fp32_params", "reduce_gradients",
              "overlapping_partition_gradients_reduce_epilogue", "destroy", "shuffle_exchange", "synchronization",
              "reset_rings"):
        assert callable(getattr(DeepSpeedZeroOptimizer, m, None)), m
    for a in ("param_groups", "state", "loss_scale", "cur_scale"):
        assert hasattr(DeepSpeedZeroOptimizer, a) or a in DeepSpeedZeroOptimizer.__init__.__code__.co_names, a
