"""Compression-aware layers: drop-in ``nn.Linear`` / ``nn.Conv2d`` / ``nn.Embedding`` / ``nn.BatchNorm2d`` that can, per
schedule, fake-quantise weights and activations (QAT) and apply sparse / row / head / channel pruning masks; masks can
later be made permanent (``fix_*_helper`` with ``dim_reduction``) to shrink the layer.

Reference: ``deepspeed/compression/basic_layer.py:17-840``.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F
from torch import nn
from torch.nn import init

from .utils import AsymQuantizer, BinaryQuantizer, SymQuantizer, TernaryQuantizer, TopKBinarizer


class QuantAct(nn.Module):
    """Activation fake-quantiser.  ``dynamic``: per-call range (optionally per token); ``static``: exponential moving
    average of min/max collected in training, frozen in eval."""

    def __init__(self, act_range_momentum=0.95, quant_mode="symmetric"):
        super().__init__()
        self.act_range_momentum = act_range_momentum
        self.quant_mode = quant_mode
        self.act_function = SymQuantizer.apply if quant_mode == "symmetric" else AsymQuantizer.apply
        self.register_buffer("x_min_max", torch.zeros(2))

    def forward(self, x, num_bits, *args):
        if self.training:
            lo, hi = x.detach().min(), x.detach().max()
            if self.x_min_max[0] == self.x_min_max[1]:
                self.x_min_max[0], self.x_min_max[1] = lo, hi
            m = self.act_range_momentum
            self.x_min_max[0] = self.x_min_max[0] * m + lo * (1 - m)
            self.x_min_max[1] = self.x_min_max[1] * m + hi * (1 - m)
        return self.act_function(x, num_bits, self.x_min_max[0], self.x_min_max[1])


class Embedding_Compress(nn.Embedding):
    def __init__(self, *kargs):
        super().__init__(*kargs)
        self.weight.start_bits = self.weight.target_bits = None
        self.weight.q_period = 0
        self.weight_quantization_enabled_in_forward = False
        self.weight_quantization_enabled = False

    def extra_repr(self):
        return "num_embeddings={}, embedding_dim={}, weight_quantization={}".format(self.num_embeddings, self.embedding_dim,
                                                                                    self.weight.target_bits)

    def enable_weight_quantization(self, start_bits, target_bits, quantization_period, weight_quantization_enabled_in_forward,
                                   quantization_type, num_groups):
        self.weight.start_bits, self.weight.target_bits, self.weight.q_period = start_bits, target_bits, quantization_period
        self.weight_quantization_enabled_in_forward = weight_quantization_enabled_in_forward
        if self.weight_quantization_enabled_in_forward:
            if self.weight.target_bits >= 3:
                self.weight_quantizer = SymQuantizer.apply if quantization_type == "symmetric" else AsymQuantizer.apply
            elif self.weight.target_bits == 2:
                assert quantization_type == "symmetric", "Only symmetric quantization is supported for ternary weight quantization"
                self.weight_quantizer = TernaryQuantizer.apply
            elif self.weight.target_bits == 1:
                assert quantization_type == "symmetric", "Only symmetric quantization is supported for binary weight quantization"
                self.weight_quantizer = BinaryQuantizer.apply
            self.weight_quantize_num_groups = self.weight.size(0)        # one group per token row

    def fix_weight_quantization(self):
        self.weight.data = self.weight_quantizer(self.weight, self.weight.target_bits, None, None,
                                                 self.weight_quantize_num_groups).data
        self.weight_quantization_enabled_in_forward = False
        return None

    def forward(self, input):
        w = self.weight
        if self.weight_quantization_enabled_in_forward and self.weight_quantization_enabled:
            w = self.weight_quantizer(self.weight, self.weight.target_bits, None, None, self.weight_quantize_num_groups)
        return F.embedding(input, w, self.padding_idx, self.max_norm, self.norm_type, self.scale_grad_by_freq, self.sparse)


class LinearLayer_Compress(nn.Linear):
    def __init__(self, *kargs, bias=True):
        super().__init__(*kargs, bias=bias)
        self.sparse_pruning_method = self.row_pruning_method = self.head_pruning_method = None
        self.activation_quantization_method = None
        self.weight.start_bits = self.weight.target_bits = None
        self.weight.q_period = 0
        self.weight_quantization_enabled_in_forward = False
        self.weight_quantization_enabled = False
        self.sparse_pruning_enabled = self.row_pruning_enabled = self.head_pruning_enabled = False
        self.activation_quantization_enabled = False

    def extra_repr(self):
        return "in_features={}, out_features={}, bias={}, sparse pruning={}, row pruning={}, head pruning={}, activation quantization={}, weight_quantization={}".format(
            self.in_features, self.out_features, self.bias is not None, self.sparse_pruning_method is not None,
            self.row_pruning_method is not None, self.head_pruning_method is not None,
            self.activation_quantization_method is not None, self.weight.target_bits)

    # --------------------------------------------------------------------------------- enable_*
    def enable_sparse_pruning(self, ratio, method):
        self.sparse_pruning_ratio, self.sparse_pruning_method = ratio, method
        if method == "l1":
            mask = self._l1_mask(self.weight.abs(), ratio).to(self.weight.device)
        elif method == "topk":
            mask = None
            self.sparse_mask_scores = nn.Parameter(torch.Tensor(self.weight.size()))
            self.sparse_mask_scores.data = self.sparse_mask_scores.data.to(self.weight.device)
            init.kaiming_uniform_(self.sparse_mask_scores, a=math.sqrt(5))
        else:
            raise NotImplementedError
        self.register_buffer("sparse_pruning_mask", mask)

    def enable_row_pruning(self, ratio, method):
        self.row_pruning_ratio, self.row_pruning_method = ratio, method
        if method == "l1":
            mask = self._l1_mask(torch.linalg.norm(self.weight.data, ord=1, dim=1), ratio).view(-1, 1).to(self.weight.device)
        elif method == "topk":
            mask = None
            self.row_mask_scores = nn.Parameter(torch.Tensor(self.weight.size(0), 1))
            self.row_mask_scores.data = self.row_mask_scores.data.to(self.weight.device)
            init.kaiming_uniform_(self.row_mask_scores, a=math.sqrt(5))
        else:
            raise NotImplementedError
        self.register_buffer("row_pruning_mask", mask)

    def enable_head_pruning(self, ratio, method, num_heads):
        self.num_heads, self.head_pruning_ratio, self.head_pruning_method = num_heads, ratio, method
        if method not in ("topk",):
            raise NotImplementedError
        self.head_pruning_scores = nn.Parameter(torch.Tensor(1, self.num_heads))
        self.head_pruning_scores.data = self.head_pruning_scores.data.to(self.weight.device)
        init.kaiming_uniform_(self.head_pruning_scores, a=math.sqrt(5))

    @staticmethod
    def _l1_mask(scores, dense_ratio):
        k = max(1, int(scores.numel() * dense_ratio))
        thresh = torch.topk(scores.flatten(), k, largest=True).values[-1]
        return (scores >= thresh).to(torch.bool)

    # ------------------------------------------------------------------------------------ fix_*
    def fix_sparse_pruning_helper(self):
        mask = self.get_mask(pruning_type="sparse")
        self.weight.data = self.weight.data * mask
        del self.sparse_pruning_mask
        if self.sparse_pruning_method == "topk":
            del self.sparse_mask_scores
        self.sparse_pruning_method = None
        self.sparse_pruning_enabled = False
        return None

    def fix_row_col_pruning_helper(self, mask=None, dim_reduction=False):
        """``mask`` None: prune my own rows (and return the mask for the consumer layer); else prune my COLUMNS with the
        producer's row mask."""
        if mask is None:
            mask = self.get_mask(pruning_type="row").bool()
            if dim_reduction:
                keep = mask.flatten().nonzero().flatten()
                self.weight = nn.Parameter(self.weight.data[keep, :])
                if self.bias is not None:
                    self.bias = nn.Parameter(self.bias.data[keep])
                self.out_features = self.weight.size(0)
            else:
                self.weight.data = self.weight.data * mask.view(-1, 1)
                if self.bias is not None:
                    self.bias.data = self.bias.data * mask.view(-1)
            del self.row_pruning_mask
            if self.row_pruning_method == "topk":
                del self.row_mask_scores
            self.row_pruning_method = None
        else:
            keep = mask.flatten().nonzero().flatten()
            self.weight = nn.Parameter(self.weight.data[:, keep])
            self.in_features = self.weight.size(1)
            mask = None
        self.row_pruning_enabled = False
        return mask

    def fix_head_pruning_helper(self, mask=None, num_heads=None, dim_reduction=False):
        """Applies to the attention OUTPUT projection (columns grouped by head); with ``mask`` given, prunes the rows of a
        q/k/v projection accordingly."""
        num_heads = num_heads if num_heads else self.num_heads
        if mask is None:
            if self.head_pruning_method == "topk":
                mask = self.get_mask(pruning_type="head").bool()
                if dim_reduction:
                    shape = self.weight.size(0)
                    keep = mask.flatten().nonzero().flatten()
                    self.weight = nn.Parameter(self.weight.data.t().reshape(num_heads, -1)[keep, :].reshape(-1, shape).t())
                    self.in_features = self.weight.size(1)
                else:
                    shape = self.weight.size()
                    self.weight.data = (self.weight.data.t().reshape(self.num_heads, -1) * mask.view(-1, 1)).reshape(shape[1], shape[0]).t()
                del self.head_pruning_scores
                self.head_pruning_method = None
            else:
                raise NotImplementedError
        else:
            shape = self.weight.size(1)
            keep = mask.flatten().nonzero().flatten()
            self.weight = nn.Parameter(self.weight.data.reshape(num_heads, -1)[keep, :].reshape(-1, shape))
            if self.bias is not None:
                self.bias = nn.Parameter(self.bias.data.reshape(num_heads, -1)[keep, :].flatten())
            self.out_features = self.weight.size(0)
        self.head_pruning_enabled = False
        return mask

    def get_mask(self, pruning_type="row"):
        if pruning_type == "sparse":
            if self.sparse_pruning_method == "l1":
                return self.sparse_pruning_mask.to(self.weight.device)
            if self.sparse_pruning_method == "topk":
                return TopKBinarizer.apply(self.sparse_mask_scores, self.sparse_pruning_ratio, False)
            raise NotImplementedError
        if pruning_type == "row":
            if self.row_pruning_method == "l1":
                return self.row_pruning_mask.to(self.weight.device)
            if self.row_pruning_method == "topk":
                return TopKBinarizer.apply(self.row_mask_scores, self.row_pruning_ratio, False)
            raise NotImplementedError
        if pruning_type == "head":
            if self.head_pruning_method == "topk":
                return TopKBinarizer.apply(self.head_pruning_scores, self.head_pruning_ratio, False)
            raise NotImplementedError
        raise NotImplementedError

    # ----------------------------------------------------------------------------- quantisation
    def enable_weight_quantization(self, start_bits, target_bits, quantization_period, weight_quantization_enabled_in_forward,
                                   quantization_type, num_groups):
        self.weight.start_bits, self.weight.target_bits, self.weight.q_period = start_bits, target_bits, quantization_period
        self.weight_quantization_enabled_in_forward = weight_quantization_enabled_in_forward
        if self.weight_quantization_enabled_in_forward:
            if self.weight.target_bits >= 3:
                self.weight_quantizer = SymQuantizer.apply if quantization_type == "symmetric" else AsymQuantizer.apply
            elif self.weight.target_bits == 2:
                assert quantization_type == "symmetric"
                self.weight_quantizer = TernaryQuantizer.apply
            elif self.weight.target_bits == 1:
                assert quantization_type == "symmetric"
                self.weight_quantizer = BinaryQuantizer.apply
            self.weight_quantize_num_groups = num_groups

    def fix_weight_quantization(self):
        self.weight.data = self.weight_quantizer(self.weight, self.weight.target_bits, None, None,
                                                 self.weight_quantize_num_groups).data
        self.weight_quantization_enabled_in_forward = False
        return None

    def enable_activation_quantization(self, bits, quantization_type, range_calibration):
        assert bits in (4, 8), "Only 4/8 bits activation quantization are supported for now"
        self.activation_quantization_bits = bits
        self.activation_quantization_method = f"{quantization_type}_{range_calibration}"
        if range_calibration == "static":
            self.activation_quantizer = QuantAct(quant_mode=quantization_type)
        else:
            self.activation_quantizer = SymQuantizer.apply if quantization_type == "symmetric" else AsymQuantizer.apply

    def head_pruning_reshape(self, w, mask):
        shape = w.shape
        return (w.t().reshape(self.num_heads, -1) * mask.view(-1, 1)).reshape(shape[1], shape[0]).t()

    def forward(self, input, skip_bias_add=False):
        w, b = self.weight, self.bias
        if self.weight_quantization_enabled_in_forward and self.weight_quantization_enabled:
            w = self.weight_quantizer(w, self.weight.target_bits, None, None, self.weight_quantize_num_groups)
        if self.sparse_pruning_enabled and self.sparse_pruning_method:
            w = w * self.get_mask(pruning_type="sparse")
        if self.row_pruning_enabled and self.row_pruning_method:
            m = self.get_mask(pruning_type="row")
            w = w * m.view(-1, 1)
            if b is not None:
                b = b * m.view(-1)
        if self.head_pruning_enabled and self.head_pruning_method:
            w = self.head_pruning_reshape(w, self.get_mask(pruning_type="head"))
        if self.activation_quantization_enabled:
            if "dynamic" in self.activation_quantization_method:
                groups = input.numel() // input.size(-1)        # one range per token
                input = self.activation_quantizer(input, self.activation_quantization_bits, None, None, groups)
            else:
                input = self.activation_quantizer(input, self.activation_quantization_bits)
        if skip_bias_add:
            return F.linear(input, w, None), b
        return F.linear(input, w, b)


class Conv2dLayer_Compress(nn.Conv2d):
    def __init__(self, *kargs):
        super().__init__(*kargs)
        self.sparse_pruning_method = self.channel_pruning_method = self.activation_quantization_method = None
        self.weight.start_bits = self.weight.target_bits = None
        self.weight.q_period = 0
        self.weight_quantization_enabled_in_forward = self.weight_quantization_enabled = False
        self.sparse_pruning_enabled = self.channel_pruning_enabled = self.activation_quantization_enabled = False

    def enable_sparse_pruning(self, ratio, method):
        self.sparse_pruning_ratio, self.sparse_pruning_method = ratio, method
        if method == "l1":
            mask = LinearLayer_Compress._l1_mask(self.weight.abs(), ratio).to(self.weight.device)
        elif method == "topk":
            mask = None
            self.sparse_mask_scores = nn.Parameter(torch.Tensor(self.weight.size()))
            init.kaiming_uniform_(self.sparse_mask_scores, a=math.sqrt(5))
        else:
            raise NotImplementedError
        self.register_buffer("sparse_pruning_mask", mask)

    def enable_channel_pruning(self, ratio, method):
        self.channel_pruning_ratio, self.channel_pruning_method = ratio, method
        if method == "l1":
            norms = torch.linalg.norm(self.weight.data.reshape(self.weight.size(0), -1), ord=1, dim=1)
            mask = LinearLayer_Compress._l1_mask(norms, ratio).view(-1, 1, 1, 1).to(self.weight.device)
        elif method == "topk":
            mask = None
            self.channel_mask_scores = nn.Parameter(torch.Tensor(self.weight.size(0), 1, 1, 1))
            init.kaiming_uniform_(self.channel_mask_scores, a=math.sqrt(5))
        else:
            raise NotImplementedError
        self.register_buffer("channel_pruning_mask", mask)

    def fix_sparse_pruning_helper(self):
        self.weight.data = self.weight.data * self.get_mask(pruning_type="sparse")
        del self.sparse_pruning_mask
        if self.sparse_pruning_method == "topk":
            del self.sparse_mask_scores
        self.sparse_pruning_method, self.sparse_pruning_enabled = None, False
        return None

    def fix_channel_pruning_helper(self, mask=None, dim_reduction=False):
        if mask is None:
            mask = self.get_mask(pruning_type="channel").bool()
            if dim_reduction:
                keep = mask.flatten().nonzero().flatten()
                self.weight = nn.Parameter(self.weight.data[keep, ...])
                if self.bias is not None:
                    self.bias = nn.Parameter(self.bias.data[keep])
                self.out_channels = self.weight.size(0)
            else:
                self.weight.data = self.weight.data * mask.view(-1, 1, 1, 1)
                if self.bias is not None:
                    self.bias.data = self.bias.data * mask.view(-1)
            del self.channel_pruning_mask
            if self.channel_pruning_method == "topk":
                del self.channel_mask_scores
            self.channel_pruning_method = None
        else:
            keep = mask.flatten().nonzero().flatten()
            self.weight = nn.Parameter(self.weight.data[:, keep, ...])
            self.in_channels = self.weight.size(1)
            mask = None
        self.channel_pruning_enabled = False
        return mask

    def get_mask(self, pruning_type="sparse"):
        if pruning_type == "sparse":
            if self.sparse_pruning_method == "l1":
                return self.sparse_pruning_mask.to(self.weight.device)
            return TopKBinarizer.apply(self.sparse_mask_scores, self.sparse_pruning_ratio, False)
        if pruning_type == "channel":
            if self.channel_pruning_method == "l1":
                return self.channel_pruning_mask.to(self.weight.device)
            return TopKBinarizer.apply(self.channel_mask_scores, self.channel_pruning_ratio, False)
        raise NotImplementedError

    enable_weight_quantization = LinearLayer_Compress.enable_weight_quantization
    fix_weight_quantization = LinearLayer_Compress.fix_weight_quantization

    def enable_activation_quantization(self, bits, quantization_type, range_calibration):
        assert bits in (4, 8) and range_calibration == "static", "Conv2d activation quantisation supports static ranges only"
        self.activation_quantization_bits = bits
        self.activation_quantization_method = f"{quantization_type}_{range_calibration}"
        self.activation_quantizer = QuantAct(quant_mode=quantization_type)

    def forward(self, input):
        w, b = self.weight, self.bias
        if self.weight_quantization_enabled_in_forward and self.weight_quantization_enabled:
            w = self.weight_quantizer(w, self.weight.target_bits, None, None, self.weight_quantize_num_groups)
        if self.sparse_pruning_enabled and self.sparse_pruning_method:
            w = w * self.get_mask(pruning_type="sparse")
        if self.channel_pruning_enabled:
            m = self.get_mask(pruning_type="channel")
            w = w * m.view(-1, 1, 1, 1)
            if b is not None:
                b = b * m.view(-1)
        if self.activation_quantization_enabled:
            input = self.activation_quantizer(input, self.activation_quantization_bits)
        return F.conv2d(input, w, b, self.stride, self.padding, self.dilation, self.groups)


class BNLayer_Compress(nn.BatchNorm2d):
    def fix_channel_pruning_helper(self, mask, dim_reduction=True):
        keep = mask.flatten().nonzero().flatten()
        self.weight = nn.Parameter(self.weight.data[keep])
        self.bias = nn.Parameter(self.bias.data[keep])
        self.running_mean = self.running_mean[keep]
        self.running_var = self.running_var[keep]
        self.num_features = self.weight.size(0)


# ================================================================================= Megatron-style tensor-parallel variants
# Reference ``compression/basic_layer.py:600-840``: Megatron's ColumnParallelLinear / RowParallelLinear re-based on
# LinearLayer_Compress so QAT / pruning also applies to tensor-parallel models.  ``mpu`` is the Megatron model-parallel
# unit (``get_model_parallel_world_size / _rank / _group``).
def _mp_group(mpu):
    return mpu.get_model_parallel_group() if hasattr(mpu, "get_model_parallel_group") else mpu.get_tensor_model_parallel_group()


def _mp_world(mpu) -> int:
    return mpu.get_model_parallel_world_size() if hasattr(mpu, "get_model_parallel_world_size") \
        else mpu.get_tensor_model_parallel_world_size()


def _mp_rank(mpu) -> int:
    return mpu.get_model_parallel_rank() if hasattr(mpu, "get_model_parallel_rank") else mpu.get_tensor_model_parallel_rank()


def split_tensor_along_last_dim(tensor, num_partitions, contiguous_split_chunks=False):
    assert tensor.size(-1) % num_partitions == 0, f"{tensor.size(-1)} is not divisible by {num_partitions}"
    parts = torch.split(tensor, tensor.size(-1) // num_partitions, dim=-1)
    return tuple(p.contiguous() for p in parts) if contiguous_split_chunks else parts


def _reduce(input_, mpu):
    if _mp_world(mpu) > 1:
        from .. import comm as dist
        dist.all_reduce(input_, group=_mp_group(mpu))
    return input_


def _split(input_, mpu):
    world = _mp_world(mpu)
    return input_ if world == 1 else split_tensor_along_last_dim(input_, world)[_mp_rank(mpu)].contiguous()


def _gather(input_, mpu):
    world = _mp_world(mpu)
    if world == 1:
        return input_
    from .. import comm as dist
    parts = [torch.empty_like(input_) for _ in range(world)]
    parts[_mp_rank(mpu)] = input_
    dist.all_gather(parts, input_.contiguous(), group=_mp_group(mpu))
    return torch.cat(parts, dim=-1).contiguous()


class _CopyToModelParallelRegion(torch.autograd.Function):
    """Identity forward, all-reduce backward (input of a column-parallel layer)."""
    @staticmethod
    def forward(ctx, input_, mpu):
        ctx.mpu = mpu
        return input_

    @staticmethod
    def backward(ctx, grad_output):
        return _reduce(grad_output.clone(), ctx.mpu), None


class _ReduceFromModelParallelRegion(torch.autograd.Function):
    """All-reduce forward, identity backward (output of a row-parallel layer)."""
    @staticmethod
    def forward(ctx, input_, mpu):
        return _reduce(input_.clone(), mpu)

    @staticmethod
    def backward(ctx, grad_output):
        return grad_output, None


class _ScatterToModelParallelRegion(torch.autograd.Function):
    """Keep this rank's chunk of the last dimension; gather in backward."""
    @staticmethod
    def forward(ctx, input_, mpu):
        ctx.mpu = mpu
        return _split(input_, mpu)

    @staticmethod
    def backward(ctx, grad_output):
        return _gather(grad_output, ctx.mpu), None


class _GatherFromModelParallelRegion(torch.autograd.Function):
    """Concatenate the ranks' chunks along the last dimension; split in backward."""
    @staticmethod
    def forward(ctx, input_, mpu):
        ctx.mpu = mpu
        return _gather(input_, mpu)

    @staticmethod
    def backward(ctx, grad_output):
        return _split(grad_output, ctx.mpu), None


def copy_to_model_parallel_region(input_, mpu): return _CopyToModelParallelRegion.apply(input_, mpu)
def reduce_from_model_parallel_region(input_, mpu): return _ReduceFromModelParallelRegion.apply(input_, mpu)
def scatter_to_model_parallel_region(input_, mpu): return _ScatterToModelParallelRegion.apply(input_, mpu)
def gather_from_model_parallel_region(input_, mpu): return _GatherFromModelParallelRegion.apply(input_, mpu)


class ColumnParallelLinear_Compress(LinearLayer_Compress):
    """``Y = X A + b`` with ``A`` split by output columns over the model-parallel group."""

    def __init__(self, mpu, input_size, output_size, bias=True, gather_output=True, skip_bias_add=False):
        self.mpu = mpu
        self.input_size, self.output_size = input_size, output_size
        self.gather_output, self.skip_bias_add = gather_output, skip_bias_add
        world = _mp_world(mpu)
        assert output_size % world == 0, f"output size {output_size} is not divisible by the model-parallel size {world}"
        self.output_size_per_partition = output_size // world
        super().__init__(input_size, self.output_size_per_partition, bias=bias)

    def forward(self, input_):
        x = copy_to_model_parallel_region(input_, self.mpu)
        bias = self.bias if not self.skip_bias_add else None
        out, b = super().forward(x, True)
        if bias is not None and b is not None:
            out = out + b
        if self.gather_output:
            out = gather_from_model_parallel_region(out, self.mpu)
        return out, (self.bias if self.skip_bias_add else None)


class RowParallelLinear_Compress(LinearLayer_Compress):
    """``Y = X A + b`` with ``A`` split by input rows: partial products are all-reduced, the bias added once."""

    def __init__(self, mpu, input_size, output_size, bias=True, input_is_parallel=False, skip_bias_add=False):
        self.mpu = mpu
        self.input_size, self.output_size = input_size, output_size
        self.input_is_parallel, self.skip_bias_add = input_is_parallel, skip_bias_add
        world = _mp_world(mpu)
        assert input_size % world == 0, f"input size {input_size} is not divisible by the model-parallel size {world}"
        self.input_size_per_partition = input_size // world
        super().__init__(self.input_size_per_partition, output_size, bias=bias)

    def forward(self, input_):
        x = input_ if self.input_is_parallel else scatter_to_model_parallel_region(input_, self.mpu)
        partial, b = super().forward(x, True)
        out = reduce_from_model_parallel_region(partial, self.mpu)
        if not self.skip_bias_add and b is not None:
            out = out + b
            return out, None
        return out, b
