import base64
import json

import pytest
import torch

from shuffle_exchange_b200.runtime.config import DeepSpeedConfig, DeepSpeedConfigError
from shuffle_exchange_b200.runtime.zero.config import get_zero_config


def test_batch_triad_inference():
    c = DeepSpeedConfig({"train_batch_size": 32, "train_micro_batch_size_per_gpu": 4})
    assert c.gradient_accumulation_steps == 8
    c = DeepSpeedConfig({"train_micro_batch_size_per_gpu": 4, "gradient_accumulation_steps": 2})
    assert c.train_batch_size == 8
    c = DeepSpeedConfig({"train_batch_size": 16})
    assert c.train_micro_batch_size_per_gpu == 16 and c.gradient_accumulation_steps == 1
    with pytest.raises(AssertionError):
        DeepSpeedConfig({"train_batch_size": 10, "train_micro_batch_size_per_gpu": 4, "gradient_accumulation_steps": 2})
    with pytest.raises(DeepSpeedConfigError):
        DeepSpeedConfig({"steps_per_print": 1})


def test_duplicate_keys_and_sources(tmp_path):
    p = tmp_path / "c.json"
    p.write_text('{"train_batch_size": 4, "train_batch_size": 8}')
    with pytest.raises(ValueError):
        DeepSpeedConfig(str(p))
    p.write_text('{"train_batch_size": 4, /* hjson comment */ "bf16": {"enabled": true}}')
    c = DeepSpeedConfig(str(p))
    assert c.bfloat16_enabled and c.train_batch_size == 4
    b64 = base64.urlsafe_b64encode(json.dumps({"train_batch_size": 2}).encode()).decode()
    assert DeepSpeedConfig(b64).train_batch_size == 2


def test_zero_config_aliases_and_deprecation():
    z = get_zero_config({"zero_optimization": {"stage": 3, "stage3_prefetch_bucket_size": 123, "cpu_offload": True}})
    assert int(z.stage) == 3 and z.prefetch_bucket_size == 123 and z.overlap_comm is True
    assert str(z.offload_optimizer.device).endswith("cpu")
    z = get_zero_config({"zero_optimization": {"stage": 1}})
    assert z.overlap_comm is False and z.reduce_bucket_size == 500000000
    z = get_zero_config({"zero_optimization": {"stage": 2, "replica_averaging": {"method": "shuffle", "rings": 4}}})
    assert z.replica_averaging.method == "shuffle" and z.replica_averaging.shuffle_step == 50
    with pytest.raises(Exception):
        get_zero_config({"zero_optimization": {"stage": 2, "no_such_key": 1}})


def test_precision_and_misc_sections():
    c = DeepSpeedConfig({"train_batch_size": 1, "fp16": {"enabled": True, "initial_scale_power": 8, "hysteresis": 3},
                         "gradient_clipping": 1.5, "communication_data_type": "bf16",
                         "optimizer": {"type": "AdamW", "params": {"lr": 1e-3}},
                         "scheduler": {"type": "WarmupLR", "params": {"warmup_num_steps": 10}},
                         "checkpoint": {"tag_validation": "Fail"}, "wall_clock_breakdown": True})
    assert c.fp16_enabled and c.initial_dynamic_scale == 256 and c.dynamic_loss_scale_args["delayed_shift"] == 3
    assert c.communication_data_type == torch.bfloat16 and c.optimizer_name == "adamw"
    assert c.checkpoint_tag_validation_fail and c.wall_clock_breakdown
    with pytest.raises(AssertionError):
        DeepSpeedConfig({"train_batch_size": 1, "fp16": {"enabled": True}, "bf16": {"enabled": True}})


def test_every_declared_scalar_key_is_read_with_its_default():
    """runtime/constants.py declares key names next to their defaults; the config object must read exactly those."""
    from shuffle_exchange_b200.runtime import constants as C
    base = {C.TRAIN_BATCH_SIZE: 4}
    c = DeepSpeedConfig(base)
    probe = {bool: lambda d: not d, int: lambda d: d + 7, float: lambda d: d + 0.5, str: lambda d: "bf16"}
    for attr, (key, default) in C.SCALAR_KEYS.items():
        if attr != "seq_parallel_communication_data_type":          # converted to a torch dtype after reading
            assert getattr(c, attr) == default, (attr, getattr(c, attr), default)
        value = probe[type(default)](default)
        c2 = DeepSpeedConfig({**base, key: value})
        got = getattr(c2, attr)
        assert got == (torch.bfloat16 if attr == "seq_parallel_communication_data_type" else value), (attr, got)
    # section names: old and new spelling of the bf16 block, optimizer / scheduler sub-keys
    c3 = DeepSpeedConfig({**base, C.BFLOAT16_OLD: {C.BFLOAT16_ENABLED: True},
                          C.OPTIMIZER: {C.TYPE: "AdamW", C.OPTIMIZER_PARAMS: {"lr": 1e-3}, C.LEGACY_FUSION: True},
                          C.SCHEDULER: {C.TYPE: "WarmupLR", C.SCHEDULER_PARAMS: {"warmup_num_steps": 3}},
                          C.CHECKPOINT: {C.CHECKPOINT_TAG_VALIDATION: "fail", C.LOAD_UNIVERSAL_CHECKPOINT: True}})
    assert c3.bfloat16_enabled and c3.optimizer_name == "adamw" and c3.optimizer_legacy_fusion
    assert c3.scheduler_name == "WarmupLR" and c3.scheduler_params == {"warmup_num_steps": 3}
    assert c3.checkpoint_tag_validation_fail and c3.load_universal_checkpoint
