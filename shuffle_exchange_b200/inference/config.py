"""Inference (v1) configuration (reference: ``deepspeed/inference/config.py:15-304``).  Field names and aliases kept;
Triton switches are accepted and ignored (this framework has no Triton path — the CUDA kernels are always on)."""
from __future__ import annotations

from enum import Enum
from typing import Any, Dict, Optional, Union

import torch
from pydantic import Field, field_validator, model_validator

from ..runtime.config_utils import DeepSpeedConfigModel
from ..runtime.zero.config import DeepSpeedZeroConfig


class DtypeEnum(Enum):
    fp16 = (torch.float16, "torch.float16", "fp16", "float16", "half")
    fp32 = (torch.float32, "torch.float32", "fp32", "float32", "float")
    bf16 = (torch.bfloat16, "torch.bfloat16", "bf16", "bfloat16", "bfloat")
    int8 = (torch.int8, "torch.int8", "int8")

    @classmethod
    def from_any(cls, v):
        for member in cls:
            if v in member.value:
                return member
        raise ValueError(f"{v!r} is not a valid dtype")

    from_str = from_any          # reference spelling (strings only there; torch dtypes are accepted too)


class MoETypeEnum(str, Enum):
    residual = "residual"
    standard = "standard"


class DeepSpeedTPConfig(DeepSpeedConfigModel):
    enabled: bool = True
    tp_size: int = 1
    tp_grain_size: int = 64
    mpu: Any = None
    tp_group: Any = None


class DeepSpeedMoEConfig(DeepSpeedConfigModel):
    enabled: bool = True
    ep_size: int = 1
    moe_experts: list = Field([1], alias="num_experts")
    type: MoETypeEnum = MoETypeEnum.standard
    ep_mp_group: Any = None
    ep_group: Any = Field(None, alias="expert_group")


class QuantTypeEnum(str, Enum):
    asym = "asymmetric"
    sym = "symmetric"


class BaseQuantConfig(DeepSpeedConfigModel):
    enabled: bool = True
    num_bits: int = 8
    q_type: QuantTypeEnum = QuantTypeEnum.sym
    q_groups: int = 1


class WeightQuantConfig(BaseQuantConfig):
    enabled: bool = True
    quantized_initialization: Dict = {}
    post_init_quant: Dict = {}


class ActivationQuantConfig(BaseQuantConfig):
    enabled: bool = True


class QKVQuantConfig(DeepSpeedConfigModel):
    enabled: bool = True


class QuantizationConfig(DeepSpeedConfigModel):
    enabled: bool = True
    activation: ActivationQuantConfig = ActivationQuantConfig()
    weight: WeightQuantConfig = WeightQuantConfig()
    qkv: QKVQuantConfig = QKVQuantConfig()


class InferenceCheckpointConfig(DeepSpeedConfigModel):
    checkpoint_dir: Optional[str] = None
    save_mp_checkpoint_path: Optional[str] = None
    base_dir: Optional[str] = None


class DeepSpeedInferenceConfig(DeepSpeedConfigModel):
    replace_with_kernel_inject: bool = Field(False, alias="kernel_inject")
    dtype: Any = torch.float16
    tensor_parallel: DeepSpeedTPConfig = Field({}, alias="tp")
    enable_cuda_graph: bool = False
    use_triton: bool = False
    triton_autotune: bool = False
    zero: DeepSpeedZeroConfig = {}
    triangular_masking: bool = Field(True, alias="tm")
    moe: Union[bool, DeepSpeedMoEConfig] = {}
    keep_module_on_host: bool = False
    quant: QuantizationConfig = {}
    checkpoint: Optional[Union[str, Dict]] = None
    base_dir: str = ""
    set_empty_params: bool = False
    save_mp_checkpoint_path: Optional[str] = None
    checkpoint_config: InferenceCheckpointConfig = Field({}, alias="ckpt_config")
    return_tuple: bool = True
    training_mp_size: int = 1
    replace_method: str = Field("auto", json_schema_extra={"deprecated": True})
    injection_policy: Optional[Dict] = Field(None, alias="injection_dict")
    injection_policy_tuple: Optional[tuple] = None
    config: Optional[Dict] = Field(None, alias="args")
    max_out_tokens: int = Field(1024, alias="max_tokens")
    min_out_tokens: int = Field(1, alias="min_tokens")
    transposed_mode: bool = Field(False, alias="transposed_mode")
    mp_size: int = Field(1, json_schema_extra={"deprecated": True, "new_param": "tensor_parallel.tp_size"})
    mpu: Any = Field(None, json_schema_extra={"deprecated": True, "new_param": "tensor_parallel.mpu"})
    ep_size: int = Field(1, json_schema_extra={"deprecated": True, "new_param": "moe.ep_size"})
    ep_group: Any = Field(None, alias="expert_group", json_schema_extra={"deprecated": True, "new_param": "moe.ep_group"})
    ep_mp_group: Any = Field(None, alias="expert_mp_group", json_schema_extra={"deprecated": True, "new_param": "moe.ep_mp_group"})
    moe_experts: list = Field([1], json_schema_extra={"deprecated": True, "new_param": "moe.moe_experts"})
    moe_type: MoETypeEnum = Field(MoETypeEnum.standard, json_schema_extra={"deprecated": True, "new_param": "moe.type"})

    @field_validator("dtype", mode="before")
    @classmethod
    def validate_dtype(cls, v):
        if isinstance(v, DtypeEnum):
            return v.value[0]
        if not isinstance(v, (str, torch.dtype)):
            raise TypeError(f"Invalid type for dtype: {type(v)}")
        return DtypeEnum.from_any(v).value[0]

    @field_validator("moe", mode="before")
    @classmethod
    def moe_backward_compat(cls, v):
        """Old configs say ``moe: true/false``; that is the ``enabled`` switch of the MoE block."""
        return {"enabled": v} if isinstance(v, bool) else v

    @field_validator("use_triton")
    @classmethod
    def has_triton(cls, v):
        if v:
            raise ValueError("use_triton: this framework has no Triton kernel path (its inference kernels are the "
                             "sm_100a CUDA kernels of csrc/); leave the switch off")
        return v

    @model_validator(mode="after")
    def _legacy(self):
        if self.mp_size != 1 and self.tensor_parallel.tp_size == 1:
            self.tensor_parallel.tp_size = self.mp_size
        if self.mpu is not None and self.tensor_parallel.mpu is None:
            self.tensor_parallel.mpu = self.mpu
        return self
