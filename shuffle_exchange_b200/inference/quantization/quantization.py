"""``_init_group_wise_weight_quantization``: replace the layers named by the ``weight_quantization.post_init_quant``
config with quantised versions (reference ``deepspeed/inference/quantization/quantization.py:20-111``).

    "weight_quantization": {"post_init_quant": {
        "fc":  {"num_bits": 4, "group_size": 64, "group_dim": 1, "symmetric": false},
        "embed": {"num_bits": 8, "group_size": 64, "group_dim": 1, "symmetric": false}}}

Every key is matched as a substring of the module path; a module may match at most one key.  Small weights are
converted first and the replaced module is dropped immediately so the peak stays near one extra layer."""
from __future__ import annotations

import contextlib
import gc
from collections import deque
from typing import Dict

import torch
from torch import nn

from ...utils.logging import logger
from . import layers
from .layers import QUANTIZATION_LAYER_MAPPINGS
from .utils import get_AsyncPartitionedParameterSwapper, recursive_setattr


def _init_group_wise_weight_quantization(model: nn.Module, ds_config: Dict) -> nn.Module:
    assert "weight_quantization" in ds_config, "Please provide quantization config in ds_config"
    qcfg = ds_config["weight_quantization"]["post_init_quant"]
    zero = ds_config.get("zero_optimization", {})
    z3 = zero.get("stage") == 3
    offloading = "offload_param" in zero
    layers.is_zero3_enabled = z3
    if z3:
        from .quantization_context import QuantizationContext
        ctx = QuantizationContext(config_dict_or_path=ds_config,
                                  param_swapper=get_AsyncPartitionedParameterSwapper(model))
    else:
        ctx = contextlib.nullcontext()
    matched: Dict[str, list] = {}
    count = 0
    with ctx:
        todo = [(n, m) for n, m in model.named_modules() if type(m) in QUANTIZATION_LAYER_MAPPINGS]
        if not offloading:
            def size(nm):
                w = nm[1].weight
                return w.ds_tensor.numel() if z3 and hasattr(w, "ds_tensor") else w.numel()
            todo.sort(key=size)
        queue = deque(todo)
        while queue:
            name, module = queue.popleft()
            keys = [k for k in qcfg if k in name]
            assert len(keys) <= 1, f"{name} matched multiple quantization key word {keys}"
            if not keys:
                continue
            key = keys[0]
            gathered = z3 and hasattr(module.weight, "all_gather")
            if gathered:
                module.weight.all_gather()
            assert module.weight.dtype in (torch.float16, torch.bfloat16), "Model weight is expected in half."
            new_module = QUANTIZATION_LAYER_MAPPINGS[type(module)](qcfg[key], module)
            if gathered:
                module.weight.partition()
            recursive_setattr(model, name, new_module)
            matched.setdefault(key, []).append(name)
            count += 1
            del module
            gc.collect()
    layers.quantized_weight_registry.clear()
    logger.info(f"Group-wise weight quantization summary: convert {count} node(s) to quantized implementation")
    logger.info("\n" + "".join(f"Key: {k}, matched modules:\n" + "".join(f"\t{n}\n" for n in names)
                               for k, names in matched.items()))
    return model
