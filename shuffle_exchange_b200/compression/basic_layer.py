"""Compression-aware layers: drop-in ``nn.Linear`` / ``nn.Conv2d`` / ``nn.Embedding`` / ``nn.BatchNorm2d`` that can, per
schedule, fake-quantise weights and activations (QAT) and apply sparse / row / head / channel pruning masks; masks can
later be made permanent (``fix_*_helper`` with ``dim_reduction``) to shrink the layer.

Public surface as in the reference (``deepspeed/compression/basic_layer.py:17-840``: class names, the ``enable_*`` /
``fix_*`` / ``get_mask`` methods, the ``*_enabled`` flags the scheduler flips, and the state-dict names of masks and
scores).  Built differently: every layer type mixes in :class:`_Compressible`, which keeps ONE table of pruning kinds
(how a mask of that kind is scored with L1 and how it broadcasts over the weight) and one weight-quantiser chooser, so
Linear / Conv2d / Embedding share the mechanics and differ only in their tables and in how a fixed mask shrinks them.
"""
from __future__ import annotations

import math
from typing import Callable, Dict, NamedTuple, Optional, Tuple

import torch
import torch.nn.functional as F
from torch import nn

from .utils import AsymQuantizer, BinaryQuantizer, SymQuantizer, TernaryQuantizer, TopKBinarizer


def _uniform_quantizer(kind: str):
    return SymQuantizer.apply if kind == "symmetric" else AsymQuantizer.apply


def _weight_quantizer(bits: int, kind: str):
    """>= 3 bits: uniform; 2: ternary; 1: binary (the last two exist only in symmetric form)."""
    if bits >= 3:
        return _uniform_quantizer(kind)
    if kind != "symmetric":
        raise AssertionError(f"Only symmetric quantization is supported for {'ternary' if bits == 2 else 'binary'} "
                             "weight quantization")
    return TernaryQuantizer.apply if bits == 2 else BinaryQuantizer.apply


def _largest(scores: torch.Tensor, dense_ratio: float) -> torch.Tensor:
    """Boolean mask keeping the ``dense_ratio`` fraction of entries with the largest score."""
    k = max(1, int(scores.numel() * dense_ratio))
    cut = torch.topk(scores.flatten(), k, largest=True).values[-1]
    return scores >= cut


class _Kind(NamedTuple):
    """One pruning kind of a layer type."""
    scores_name: str                                   # name of the learnable-score parameter (method ``topk``)
    mask_name: Optional[str]                           # name of the fixed-mask buffer (method ``l1``); None: no l1 form
    shape: Callable[[nn.Module], Tuple[int, ...]]      # shape of mask / scores
    importance: Optional[Callable[[torch.Tensor], torch.Tensor]]   # L1 importance of the weight, in ``shape``


class QuantAct(nn.Module):
    """Activation fake-quantiser with a calibrated (static) range: exponential moving average of min/max collected in
    training, frozen in eval."""

    def __init__(self, act_range_momentum=0.95, quant_mode="symmetric"):
        super().__init__()
        self.act_range_momentum = act_range_momentum
        self.quant_mode = quant_mode
        self.act_function = _uniform_quantizer(quant_mode)
        self.register_buffer("x_min_max", torch.zeros(2))

    def forward(self, x, num_bits, *args):
        if self.training:
            seen = torch.stack(torch.aminmax(x.detach().float()))
            if self.x_min_max[0] == self.x_min_max[1]:          # first batch: start the average at the observed range
                self.x_min_max.copy_(seen)
            self.x_min_max.lerp_(seen.to(self.x_min_max.dtype), 1 - self.act_range_momentum)
        return self.act_function(x, num_bits, self.x_min_max[0], self.x_min_max[1])


class _Compressible:
    """Mixin over an ``nn`` layer with a ``weight``: pruning-kind table + weight / activation quantisation state."""

    _KINDS: Dict[str, _Kind] = {}

    def _init_compression(self):
        for kind in self._KINDS:
            setattr(self, f"{kind}_pruning_method", None)
            setattr(self, f"{kind}_pruning_enabled", False)
        self.activation_quantization_method = None
        self.activation_quantization_enabled = False
        self.weight.start_bits = self.weight.target_bits = None
        self.weight.q_period = 0
        self.weight_quantization_enabled_in_forward = False
        self.weight_quantization_enabled = False

    # ------------------------------------------------------------------------------------ pruning
    def _enable_pruning(self, kind: str, ratio: float, method: str):
        spec = self._KINDS[kind]
        setattr(self, f"{kind}_pruning_ratio", ratio)
        setattr(self, f"{kind}_pruning_method", method)
        mask = None
        if method == "l1" and spec.mask_name is not None:
            mask = _largest(spec.importance(self.weight.data), ratio).view(spec.shape(self)).to(self.weight.device)
        elif method == "topk":
            scores = nn.Parameter(torch.empty(spec.shape(self), device=self.weight.device))
            nn.init.kaiming_uniform_(scores, a=math.sqrt(5))
            setattr(self, spec.scores_name, scores)
        else:
            raise NotImplementedError(f"{kind} pruning has no method {method!r}")
        if spec.mask_name is not None:
            self.register_buffer(spec.mask_name, mask)

    def get_mask(self, pruning_type):
        spec = self._KINDS.get(pruning_type)
        method = getattr(self, f"{pruning_type}_pruning_method", None) if spec else None
        if method == "l1" and spec.mask_name is not None:
            return getattr(self, spec.mask_name).to(self.weight.device)
        if method == "topk":
            return TopKBinarizer.apply(getattr(self, spec.scores_name), getattr(self, f"{pruning_type}_pruning_ratio"), False)
        raise NotImplementedError(f"no {pruning_type} pruning mask (method {method!r})")

    def _active(self, kind: str) -> bool:
        return bool(getattr(self, f"{kind}_pruning_enabled") and getattr(self, f"{kind}_pruning_method"))

    def _retire_pruning(self, kind: str):
        """Forget a pruning kind after its mask was burnt into the weight."""
        spec = self._KINDS[kind]
        if spec.mask_name is not None and hasattr(self, spec.mask_name):
            delattr(self, spec.mask_name)
        if getattr(self, f"{kind}_pruning_method") == "topk":
            delattr(self, spec.scores_name)
        setattr(self, f"{kind}_pruning_method", None)
        setattr(self, f"{kind}_pruning_enabled", False)

    def enable_sparse_pruning(self, ratio, method):
        self._enable_pruning("sparse", ratio, method)

    def fix_sparse_pruning_helper(self):
        self.weight.data = self.weight.data * self.get_mask("sparse")
        self._retire_pruning("sparse")
        return None

    # ------------------------------------------------------------------------------- quantisation
    def _default_groups(self, num_groups):
        return num_groups

    def enable_weight_quantization(self, start_bits, target_bits, quantization_period, weight_quantization_enabled_in_forward,
                                   quantization_type, num_groups):
        w = self.weight
        w.start_bits, w.target_bits, w.q_period = start_bits, target_bits, quantization_period
        self.weight_quantization_enabled_in_forward = weight_quantization_enabled_in_forward
        if weight_quantization_enabled_in_forward:
            self.weight_quantizer = _weight_quantizer(target_bits, quantization_type)
            self.weight_quantize_num_groups = self._default_groups(num_groups)

    def _quantized_weight(self):
        return self.weight_quantizer(self.weight, self.weight.target_bits, None, None, self.weight_quantize_num_groups)

    def fix_weight_quantization(self):
        self.weight.data = self._quantized_weight().data
        self.weight_quantization_enabled_in_forward = False
        return None

    def _forward_weight(self):
        if self.weight_quantization_enabled_in_forward and self.weight_quantization_enabled:
            return self._quantized_weight()
        return self.weight

    def _enable_activation_quantization(self, bits, quantization_type, range_calibration, allow_dynamic: bool):
        if bits not in (4, 8):
            raise AssertionError("Only 4/8 bits activation quantization are supported for now")
        if range_calibration != "static" and not allow_dynamic:
            raise AssertionError(f"{type(self).__name__} supports static activation ranges only")
        self.activation_quantization_bits = bits
        self.activation_quantization_method = f"{quantization_type}_{range_calibration}"
        self.activation_quantizer = QuantAct(quant_mode=quantization_type) if range_calibration == "static" \
            else _uniform_quantizer(quantization_type)

    def _forward_input(self, x):
        if not self.activation_quantization_enabled:
            return x
        if "dynamic" in self.activation_quantization_method:        # one range per token
            return self.activation_quantizer(x, self.activation_quantization_bits, None, None, x.numel() // x.size(-1))
        return self.activation_quantizer(x, self.activation_quantization_bits)

    def _flags_repr(self):
        bits = [f"{kind} pruning={getattr(self, f'{kind}_pruning_method') is not None}" for kind in self._KINDS]
        bits.append(f"activation quantization={self.activation_quantization_method is not None}")
        bits.append(f"weight_quantization={self.weight.target_bits}")
        return ", ".join(bits)


def _kept(mask: torch.Tensor) -> torch.Tensor:
    return mask.flatten().nonzero().flatten()


class Embedding_Compress(_Compressible, nn.Embedding):
    """Embedding with per-row (one group per token id) weight fake-quantisation."""

    def __init__(self, *kargs):
        super().__init__(*kargs)
        self._init_compression()

    def _default_groups(self, num_groups):
        return self.weight.size(0)

    def extra_repr(self):
        return f"num_embeddings={self.num_embeddings}, embedding_dim={self.embedding_dim}, " \
               f"weight_quantization={self.weight.target_bits}"

    def forward(self, input):
        return F.embedding(input, self._forward_weight(), self.padding_idx, self.max_norm, self.norm_type,
                           self.scale_grad_by_freq, self.sparse)


class LinearLayer_Compress(_Compressible, nn.Linear):
    _KINDS = {
        "sparse": _Kind("sparse_mask_scores", "sparse_pruning_mask", lambda m: tuple(m.weight.shape), torch.abs),
        "row": _Kind("row_mask_scores", "row_pruning_mask", lambda m: (m.weight.size(0), 1),
                     lambda w: w.abs().sum(dim=1, keepdim=True)),
        "head": _Kind("head_pruning_scores", None, lambda m: (1, m.num_heads), None),
    }

    def __init__(self, *kargs, bias=True):
        super().__init__(*kargs, bias=bias)
        self._init_compression()

    def extra_repr(self):
        return f"in_features={self.in_features}, out_features={self.out_features}, bias={self.bias is not None}, " \
               + self._flags_repr()

    # --------------------------------------------------------------------------------- enable_*
    def enable_row_pruning(self, ratio, method):
        self._enable_pruning("row", ratio, method)

    def enable_head_pruning(self, ratio, method, num_heads):
        self.num_heads = num_heads
        if method != "topk":
            raise NotImplementedError("head pruning is learnt (method 'topk') only")
        self._enable_pruning("head", ratio, method)

    def enable_activation_quantization(self, bits, quantization_type, range_calibration):
        self._enable_activation_quantization(bits, quantization_type, range_calibration, allow_dynamic=True)

    # ------------------------------------------------------------------------------------ fix_*
    def fix_row_col_pruning_helper(self, mask=None, dim_reduction=False):
        """Without ``mask``: burn MY row mask in (dropping the rows when ``dim_reduction``) and hand it back for the
        consumer layer.  With ``mask`` (the producer's row mask): drop my matching input columns."""
        if mask is not None:
            self.weight = nn.Parameter(self.weight.data[:, _kept(mask)])
            self.in_features = self.weight.size(1)
            self.row_pruning_enabled = False
            return None
        mask = self.get_mask("row").bool()
        has_bias = self.bias is not None
        if dim_reduction:
            rows = _kept(mask)
            self.weight = nn.Parameter(self.weight.data[rows])
            if has_bias:
                self.bias = nn.Parameter(self.bias.data[rows])
            self.out_features = self.weight.size(0)
        else:
            self.weight.data.mul_(mask.view(-1, 1))
            if has_bias:
                self.bias.data.mul_(mask.view(-1))
        self._retire_pruning("row")
        return mask

    def _by_head(self, w, heads):
        """Input columns of an attention output projection grouped per head: ``[heads, head_dim * out]``."""
        return w.t().reshape(heads, -1)

    def head_pruning_reshape(self, w, mask):
        out_f, in_f = w.shape
        return (self._by_head(w, self.num_heads) * mask.view(-1, 1)).reshape(in_f, out_f).t()

    def fix_head_pruning_helper(self, mask=None, num_heads=None, dim_reduction=False):
        """Without ``mask``: this is the attention OUTPUT projection, whose input columns are grouped by head.  With
        ``mask``: this is a q/k/v projection, whose output rows are dropped for the pruned heads."""
        heads = num_heads or self.num_heads
        if mask is not None:
            alive = _kept(mask)
            in_f = self.weight.size(1)
            self.weight = nn.Parameter(self.weight.data.reshape(heads, -1)[alive].reshape(-1, in_f))
            if self.bias is not None:
                self.bias = nn.Parameter(self.bias.data.reshape(heads, -1)[alive].flatten())
            self.out_features = self.weight.size(0)
            self.head_pruning_enabled = False
            return mask
        if self.head_pruning_method != "topk":
            raise NotImplementedError
        mask = self.get_mask("head").bool()
        if dim_reduction:
            out_f = self.weight.size(0)
            self.weight = nn.Parameter(self._by_head(self.weight.data, heads)[_kept(mask)].reshape(-1, out_f).t())
            self.in_features = self.weight.size(1)
        else:
            self.weight.data = self.head_pruning_reshape(self.weight.data, mask)
        self._retire_pruning("head")
        return mask

    def forward(self, input, skip_bias_add=False):
        w, b = self._forward_weight(), self.bias
        if self._active("sparse"):
            w = w * self.get_mask("sparse")
        if self._active("row"):
            rows = self.get_mask("row")
            w = w * rows.view(-1, 1)
            b = None if b is None else b * rows.view(-1)
        if self._active("head"):
            w = self.head_pruning_reshape(w, self.get_mask("head"))
        y = F.linear(self._forward_input(input), w, None if skip_bias_add else b)
        return (y, b) if skip_bias_add else y


class Conv2dLayer_Compress(_Compressible, nn.Conv2d):
    _KINDS = {
        "sparse": _Kind("sparse_mask_scores", "sparse_pruning_mask", lambda m: tuple(m.weight.shape), torch.abs),
        "channel": _Kind("channel_mask_scores", "channel_pruning_mask", lambda m: (m.weight.size(0), 1, 1, 1),
                         lambda w: w.abs().sum(dim=(1, 2, 3), keepdim=True)),
    }

    def __init__(self, *kargs):
        super().__init__(*kargs)
        self._init_compression()

    def extra_repr(self):
        return super().extra_repr() + ", " + self._flags_repr()

    def enable_channel_pruning(self, ratio, method):
        self._enable_pruning("channel", ratio, method)

    def enable_activation_quantization(self, bits, quantization_type, range_calibration):
        self._enable_activation_quantization(bits, quantization_type, range_calibration, allow_dynamic=False)

    def fix_channel_pruning_helper(self, mask=None, dim_reduction=False):
        """Without ``mask``: burn in / drop MY output channels and return the mask; with the producer's ``mask``: drop
        the matching input channels."""
        if mask is not None:
            self.weight = nn.Parameter(self.weight.data[:, _kept(mask)])
            self.in_channels = self.weight.size(1)
            self.channel_pruning_enabled = False
            return None
        mask = self.get_mask("channel").bool()
        has_bias = self.bias is not None
        if dim_reduction:
            chans = _kept(mask)
            self.weight = nn.Parameter(self.weight.data[chans])
            if has_bias:
                self.bias = nn.Parameter(self.bias.data[chans])
            self.out_channels = self.weight.size(0)
        else:
            self.weight.data.mul_(mask.view(-1, 1, 1, 1))
            if has_bias:
                self.bias.data.mul_(mask.view(-1))
        self._retire_pruning("channel")
        return mask

    def forward(self, input):
        w, b = self._forward_weight(), self.bias
        if self._active("sparse"):
            w = w * self.get_mask("sparse")
        if self.channel_pruning_enabled:
            chans = self.get_mask("channel")
            w = w * chans.view(-1, 1, 1, 1)
            b = None if b is None else b * chans.view(-1)
        return F.conv2d(self._forward_input(input), w, b, self.stride, self.padding, self.dilation, self.groups)


class BNLayer_Compress(nn.BatchNorm2d):
    def fix_channel_pruning_helper(self, mask, dim_reduction=True):
        """Follow the channel mask of the convolution that feeds this norm."""
        chans = _kept(mask)
        for name in ("weight", "bias"):
            setattr(self, name, nn.Parameter(getattr(self, name).data[chans]))
        for name in ("running_mean", "running_var"):
            setattr(self, name, getattr(self, name)[chans])
        self.num_features = chans.numel()


# ================================================================================= tensor-parallel variants
# Reference ``compression/basic_layer.py:600-840`` re-bases Megatron's ColumnParallelLinear / RowParallelLinear on
# LinearLayer_Compress so QAT / pruning also applies to tensor-parallel models.  ``mpu`` is the Megatron model-parallel
# unit (``get_model_parallel_world_size / _rank / _group``, or the ``tensor_model_parallel`` spellings).
class _MP:
    """The three facts the regions need from an ``mpu``, under either naming."""

    def __init__(self, mpu):
        pick = lambda what: getattr(mpu, f"get_model_parallel_{what}", None) or getattr(mpu, f"get_tensor_model_parallel_{what}")
        self.world, self.rank, self.group = pick("world_size")(), pick("rank")(), pick("group")()


def split_tensor_along_last_dim(tensor, num_partitions, contiguous_split_chunks=False):
    last = tensor.size(-1)
    if last % num_partitions:
        raise AssertionError(f"{last} is not divisible by {num_partitions}")
    chunks = tensor.split(last // num_partitions, dim=-1)
    return tuple(c.contiguous() for c in chunks) if contiguous_split_chunks else chunks


def _identity(x, mp):
    return x


def _sum_over_group(x, mp):
    if mp.world == 1:
        return x
    from .. import comm as dist
    x = x.clone()
    dist.all_reduce(x, group=mp.group)
    return x


def _my_chunk(x, mp):
    return x if mp.world == 1 else split_tensor_along_last_dim(x, mp.world)[mp.rank].contiguous()


def _all_chunks(x, mp):
    if mp.world == 1:
        return x
    from .. import comm as dist
    x = x.contiguous()
    chunks = [torch.empty_like(x) for _ in range(mp.world)]
    dist.all_gather(chunks, x, group=mp.group)
    chunks[mp.rank] = x
    return torch.cat(chunks, dim=-1)


class _Region(torch.autograd.Function):
    """A boundary of a tensor-parallel region: ``fwd`` on the way in, its transpose ``bwd`` on the gradient."""

    @staticmethod
    def forward(ctx, x, mpu, fwd, bwd):
        ctx.mp, ctx.bwd = _MP(mpu), bwd
        return fwd(x, ctx.mp)

    @staticmethod
    def backward(ctx, g):
        return ctx.bwd(g, ctx.mp), None, None, None


def copy_to_model_parallel_region(input_, mpu):
    """Identity forward, all-reduce backward (input of a column-parallel layer)."""
    return _Region.apply(input_, mpu, _identity, _sum_over_group)


def reduce_from_model_parallel_region(input_, mpu):
    """All-reduce forward, identity backward (output of a row-parallel layer)."""
    return _Region.apply(input_, mpu, _sum_over_group, _identity)


def scatter_to_model_parallel_region(input_, mpu):
    """Keep this rank's chunk of the last dimension; gather in backward."""
    return _Region.apply(input_, mpu, _my_chunk, _all_chunks)


def gather_from_model_parallel_region(input_, mpu):
    """Concatenate the ranks' chunks along the last dimension; split in backward."""
    return _Region.apply(input_, mpu, _all_chunks, _my_chunk)


def _share(total: int, mpu, what: str) -> int:
    world = _MP(mpu).world
    if total % world:
        raise AssertionError(f"{what} size {total} is not divisible by the model-parallel size {world}")
    return total // world


class ColumnParallelLinear_Compress(LinearLayer_Compress):
    """``Y = X A + b`` with ``A`` split by output columns over the model-parallel group.  Returns ``(Y, bias or None)``."""

    def __init__(self, mpu, input_size, output_size, bias=True, gather_output=True, skip_bias_add=False):
        self.mpu = mpu
        self.input_size, self.output_size = input_size, output_size
        self.gather_output, self.skip_bias_add = gather_output, skip_bias_add
        self.output_size_per_partition = _share(output_size, mpu, "output")
        super().__init__(input_size, self.output_size_per_partition, bias=bias)

    def forward(self, input_):
        y, b = super().forward(copy_to_model_parallel_region(input_, self.mpu), True)
        if not self.skip_bias_add and b is not None:
            y = y + b
        if self.gather_output:
            y = gather_from_model_parallel_region(y, self.mpu)
        return y, (self.bias if self.skip_bias_add else None)


class RowParallelLinear_Compress(LinearLayer_Compress):
    """``Y = X A + b`` with ``A`` split by input rows: partial products are all-reduced, the bias added once."""

    def __init__(self, mpu, input_size, output_size, bias=True, input_is_parallel=False, skip_bias_add=False):
        self.mpu = mpu
        self.input_size, self.output_size = input_size, output_size
        self.input_is_parallel, self.skip_bias_add = input_is_parallel, skip_bias_add
        self.input_size_per_partition = _share(input_size, mpu, "input")
        super().__init__(self.input_size_per_partition, output_size, bias=bias)

    def forward(self, input_):
        x = input_ if self.input_is_parallel else scatter_to_model_parallel_region(input_, self.mpu)
        partial, b = super().forward(x, True)
        y = reduce_from_model_parallel_region(partial, self.mpu)
        if self.skip_bias_add or b is None:
            return y, b
        return y + b, None
