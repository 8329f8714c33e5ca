"""Module surgery for compression (reference: ``deepspeed/compression/helper.py:17-322``)."""
from __future__ import annotations

import torch

from ..utils.logging import logger
from torch import nn

from .basic_layer import BNLayer_Compress, Conv2dLayer_Compress, Embedding_Compress, LinearLayer_Compress
from .constants import *  # noqa: F401,F403


def recursive_getattr(model, module_name):
    out = model
    for part in module_name.split("."):
        out = getattr(out, part)
    return out


def recursive_setattr(model, module_name, module):
    parts = module_name.split(".")
    holder = model
    for part in parts[:-1]:
        holder = getattr(holder, part)
    setattr(holder, parts[-1], module)


def _clone_into(old, new):
    new.weight.data = old.weight.data
    if getattr(old, "bias", None) is not None:
        new.bias.data = old.bias.data
    return new


def module_replacement(model, module_name, compression_technique=None, mpu=None):
    """Swap ``module_name`` for its compression-aware class (once) and switch on the requested techniques."""
    old = recursive_getattr(model, module_name)
    if isinstance(old, (LinearLayer_Compress, Conv2dLayer_Compress, Embedding_Compress, BNLayer_Compress)):
        new = old
    elif isinstance(old, nn.Linear):
        new = _clone_into(old, LinearLayer_Compress(old.in_features, old.out_features, bias=old.bias is not None)
                          .to(device=old.weight.device, dtype=old.weight.dtype))
    elif isinstance(old, nn.Conv2d):
        new = _clone_into(old, Conv2dLayer_Compress(old.in_channels, old.out_channels, old.kernel_size, old.stride, old.padding,
                                                    old.dilation, old.groups, old.bias is not None, old.padding_mode)
                          .to(device=old.weight.device, dtype=old.weight.dtype))
    elif isinstance(old, nn.BatchNorm2d):
        new = BNLayer_Compress(old.num_features, old.eps, old.momentum, old.affine, old.track_running_stats) \
            .to(old.weight.device, old.weight.dtype)
        new.load_state_dict(old.state_dict())
    elif isinstance(old, nn.Embedding):
        new = Embedding_Compress(old.num_embeddings, old.embedding_dim, old.padding_idx, old.max_norm, old.norm_type,
                                 old.scale_grad_by_freq, old.sparse).to(device=old.weight.device, dtype=old.weight.dtype)
        new.weight.data = old.weight.data
    else:
        new = None
    if compression_technique is not None and new is not None:
        for name, p in compression_technique.items():
            if name == SPARSE_PRUNING and p[SPARSE_PRUNING_ENABLED]:
                new.enable_sparse_pruning(p[SPARSE_PRUNING_DENSE_RATIO], p[SPARSE_PRUNING_METHOD])
            elif name == ROW_PRUNING and p[ROW_PRUNING_ENABLED]:
                new.enable_row_pruning(p[ROW_PRUNING_DENSE_RATIO], p[ROW_PRUNING_METHOD])
            elif name == HEAD_PRUNING and p[HEAD_PRUNING_ENABLED]:
                new.enable_head_pruning(p[HEAD_PRUNING_DENSE_RATIO], p[HEAD_PRUNING_METHOD], p[HEAD_PRUNING_NUM_HEADS])
            elif name == ACTIVATION_QUANTIZATION and p[ACTIVATION_QUANTIZATION_ENABLED]:
                new.enable_activation_quantization(p[ACTIVATION_QUANTIZE_BITS], p[ACTIVATION_QUANTIZE_TYPE], p[ACTIVATION_QUANTIZE_RANGE])
            elif name == WEIGHT_QUANTIZATION and p[WEIGHT_QUANTIZE_ENABLED]:
                new.enable_weight_quantization(p[WEIGHT_QUANTIZE_START_BITS], p[WEIGHT_QUANTIZE_TARGET_BITS],
                                               p[WEIGHT_QUANTIZATION_PERIOD], p[WEIGHT_QUANTIZE_IN_FORWARD_ENABLED],
                                               p[WEIGHT_QUANTIZE_TYPE], p[WEIGHT_QUANTIZE_GROUPS])
            elif name == CHANNEL_PRUNING and p[CHANNEL_PRUNING_ENABLED]:
                new.enable_channel_pruning(p[CHANNEL_PRUNING_DENSE_RATIO], p[CHANNEL_PRUNING_METHOD])
    if new is not None and new is not old:
        recursive_setattr(model, module_name, new)
    return model


def is_module_compressible(module, mpu=None):
    return isinstance(module, (nn.Linear, nn.Embedding, nn.BatchNorm2d, nn.Conv2d))


def compression_preparation(model, compression_technique_list, mpu):
    """First turn every compressible module into its compression class, then enable the per-group techniques."""
    for name, m in list(model.named_modules()):
        if is_module_compressible(m, mpu):
            module_replacement(model, name, mpu=mpu)
    for module_names, _related, technique in compression_technique_list:
        for name in module_names:
            module_replacement(model, name, technique)
    return model


def fix_compression(model, module_name, compression_technique, mask=None, dim_reduction=False):
    """Make the masks / fake quantisation of ``module_name`` permanent.  Returns the mask a related module must apply."""
    module = recursive_getattr(model, module_name)
    for name, p in compression_technique.items():
        if name == WEIGHT_QUANTIZATION and p[WEIGHT_QUANTIZE_IN_FORWARD_ENABLED] and p[WEIGHT_QUANTIZE_ENABLED]:
            return module.fix_weight_quantization()
        if name == SPARSE_PRUNING and p[SPARSE_PRUNING_ENABLED]:
            return module.fix_sparse_pruning_helper()
        if name == ROW_PRUNING and (p[ROW_PRUNING_ENABLED] or mask is not None):
            return module.fix_row_col_pruning_helper(mask, dim_reduction=dim_reduction)
        if name == HEAD_PRUNING and (p[HEAD_PRUNING_ENABLED] or mask is not None):
            return module.fix_head_pruning_helper(mask, p[HEAD_PRUNING_NUM_HEADS], dim_reduction=dim_reduction)
        if name == CHANNEL_PRUNING and (p[CHANNEL_PRUNING_ENABLED] or mask is not None):
            return module.fix_channel_pruning_helper(mask, dim_reduction=dim_reduction)
    return None


def convert_conv1d_to_linear(model, convert_type):
    """GPT-2 style ``Conv1D`` -> ``nn.Linear`` (transposed weight) so the compression layers apply."""
    for name, m in list(model.named_modules()):
        if isinstance(m, convert_type):
            lin = nn.Linear(m.weight.data.size(0), m.weight.data.size(1), bias=m.bias is not None)
            lin.weight.data = m.weight.data.t().contiguous()
            if m.bias is not None:
                lin.bias.data = m.bias.data
            recursive_setattr(model, name, lin)
    return model


# ---- snip_momentum sparse pruning through Intel neural-compressor (reference ``compression/helper.py:274-322``)
def generate_pruners(config, model):
    """Create neural-compressor pruners for ``model`` from its pruning config list.  neural-compressor is optional: it
    is only needed for the ``snip_momentum`` method, every other pruning method is built in."""
    try:
        from neural_compressor.compression.pruner import utils as nc_utils
        from neural_compressor.compression.pruner.pruners import get_pruner
    except ImportError as e:
        raise ImportError("snip_momentum pruning needs neural-compressor >= 2.1.1 (`pip install neural-compressor`)") from e
    assert isinstance(model, torch.nn.Module)
    pruners_info = nc_utils.process_config(config)
    pruners = []
    for info in pruners_info:
        modules = nc_utils.parse_to_prune(info, model)
        if not modules:
            logger.warning("one pruner hooks no layers, please have a check")
        pruners.append(get_pruner(info, modules))
        info["modules"] = list(modules.keys())
        info["len_of_modules"] = len(info["modules"])
        logger.info(info)
    return pruners


def register_on_step_begin(model):
    """Run every pruner's ``on_step_begin`` right before each forward (global step counter kept by the pruner)."""
    def hook(module, input):
        for pruner in getattr(module, "pruners", ()):
            pruner.on_step_begin(0)
    return model.register_forward_pre_hook(hook)


def rewrite_optimizer_step(opt):
    """Bracket ``opt.step`` with the pruners' ``on_before_optimizer_step`` / ``on_after_optimizer_step`` (the mask is
    re-applied after the update so pruned weights stay zero)."""
    import types

    def new_step(self, closure=None):
        for pruner in getattr(self, "pruners", ()):
            pruner.on_before_optimizer_step()
        res = self.orig_step(closure) if closure is not None else self.orig_step()
        for pruner in getattr(self, "pruners", ()):
            pruner.on_after_optimizer_step()
        return res

    opt.orig_step = opt.step
    opt.step = types.MethodType(new_step, opt)
    return opt
