"""Tensor-parallel building blocks used by AutoTP (training and inference).

Reference: ``deepspeed/module_inject/layers.py`` — autograd glue ``RowParallel / ColumnParallel /
AsyncColumnParallel`` (:64-152), ``TensorParallel_Layer`` base with gather/partition of parameters (:155-316),
``LinearAllreduce`` (row-parallel: partial products summed over the TP group, :388-462), ``LinearLayer``
(column-parallel, :465-547), fused-QKV / gate-up packed variants (:559-652), ``LmHeadLinearAllreduce`` (:655-680),
``GatherReplacedLayerParams`` (:326-385) and small replacement modules (``Normalize``, ``EmbeddingLayer``,
``RMSNormalize``).

B200 path: the row-parallel GEMM writes its partial product straight into a symmetric-memory buffer and the sum over
the TP group is ONE NVLink kernel (``sxb_symm_subset_mean`` with NVLS in-switch reduction when the fabric offers
multicast) instead of an NCCL all-reduce; everywhere else (CPU, fp32, groups without symmetric memory) the
``comm`` facade is used.
"""
from __future__ import annotations

from abc import ABC, abstractmethod
from copy import deepcopy
from typing import Any, Iterable, List, Optional, Tuple

import torch
import torch.nn as nn
from torch.nn import functional as TF

from .. import comm as dist
from ..ops import functional as F

_TP_TRAINING = False


def get_auto_tp_mode() -> bool:
    return _TP_TRAINING


def is_autotp_training_mode() -> bool:
    return _TP_TRAINING


def set_autotp_mode(training: bool = False) -> None:
    """Training mode keeps autograd-aware collectives; inference mode may use in-place / fused reductions."""
    global _TP_TRAINING
    _TP_TRAINING = bool(training)


def add_bias(input, bias):
    if bias is None:
        return input
    if is_autotp_training_mode():
        return input + bias
    input += bias
    return input


def _project(x: torch.Tensor, weight: torch.Tensor) -> torch.Tensor:
    """``x @ weight^T`` of a TP linear: the framework's tcgen05 linear (forward, dgrad and wgrad) for bf16 CUDA operands
    with 16-byte-aligned rows; PyTorch for CPU tensors and other dtypes."""
    if x.is_cuda and x.dtype == torch.bfloat16 and weight.dtype == torch.bfloat16 \
            and x.shape[-1] % 8 == 0 and weight.shape[0] % 8 == 0:
        return F.linear(x, weight)
    return TF.linear(x, weight)


def _tp_world(group) -> int:
    return dist.get_world_size(group=group) if (group is not None and dist.is_initialized()) else 1


def _all_reduce_sum(x: torch.Tensor, group, inplace: bool) -> torch.Tensor:
    if _tp_world(group) == 1:
        return x
    fused = _symm_reducer(x, group)
    if fused is not None:
        return fused(x)
    out = x if inplace else x.clone()
    dist.all_reduce(out.contiguous(), group=group)
    return out


_SYMM_CACHE = {}


def _symm_reducer(x: torch.Tensor, group):
    """One-shot NVLink sum for bf16 CUDA activations when symmetric memory is available for ``group``."""
    if not (x.is_cuda and x.dtype == torch.bfloat16 and x.is_contiguous()):
        return None
    try:
        from ..comm import symm
        if not symm.available(group):
            return None
        sg = symm.SymmGroup.get(group)
    except Exception:
        return None
    n = x.numel()
    if n % 8:
        return None

    def run(t):
        buf = sg.buffer("tp_allreduce", n, t.dtype)
        sg.wait_readers(buf, symm.CH_AVG_READ)
        buf.tensor[:n].copy_(t.view(-1))
        out = torch.empty_like(t)
        # large activations (prefill): two-phase form, 2 (w-1)/w instead of (w-1) tensor sizes over NVLink; small ones
        # (decode) stay on the single-pass kernel, which has one handshake less
        sg.subset_mean(buf, list(range(sg.world)), out.view(-1), n, shared=n * t.element_size() >= (8 << 20))
        return out.mul_(sg.world)               # kernel returns the mean
    return run


class RowParallel(torch.autograd.Function):
    """Forward: sum over the TP group.  Backward: identity (every rank already holds the full upstream gradient)."""

    @staticmethod
    def symbolic(graph, input):
        return input

    @staticmethod
    def forward(ctx: Any, group, input: torch.Tensor, is_inference_mode: bool) -> torch.Tensor:
        ctx.group = group
        if group is None:
            return input
        return _all_reduce_sum(input, group, inplace=is_inference_mode)

    @staticmethod
    def backward(ctx: Any, grad_output: torch.Tensor) -> Tuple[None, torch.Tensor, None]:
        return None, grad_output, None


class ColumnParallel(torch.autograd.Function):
    """Forward: identity.  Backward: sum the input gradient over the TP group."""

    @staticmethod
    def symbolic(graph, input):
        return input

    @staticmethod
    def forward(ctx: Any, group, input: torch.Tensor) -> torch.Tensor:
        ctx.group = group
        return input

    @staticmethod
    def backward(ctx: Any, grad_output: torch.Tensor) -> Tuple[None, torch.Tensor]:
        if ctx.group is None:
            return None, grad_output
        return None, _all_reduce_sum(grad_output.contiguous(), ctx.group, inplace=True)


class AsyncColumnParallel(torch.autograd.Function):
    """Column-parallel linear whose input-gradient all-reduce overlaps the weight-gradient GEMM."""

    @staticmethod
    def forward(ctx: Any, group, input: torch.Tensor, weight, bias) -> torch.Tensor:
        ctx.use_bias = bias is not None
        ctx.group = group
        out = F.matmul_nt(input, weight)
        if bias is not None:
            out = out + bias
        ctx.save_for_backward(input, weight)
        return out

    @staticmethod
    def backward(ctx: Any, grad_output: torch.Tensor):
        input, weight = ctx.saved_tensors
        grad_input = F.matmul_nn(grad_output, weight).contiguous()       # reduced in place while the wgrad GEMM runs
        handle = dist.all_reduce(grad_input, group=ctx.group, async_op=True) if _tp_world(ctx.group) > 1 else None
        g2 = grad_output.reshape(-1, grad_output.shape[-1])
        grad_weight = F.matmul_tn(g2, input).to(weight.dtype)
        grad_bias = g2.sum(0) if ctx.use_bias else None
        if handle is not None:
            handle.wait()
        return None, grad_input, grad_weight, grad_bias


class TensorParallel_Layer(nn.Module, ABC):
    """Base of the TP linears: knows its group, how to cut (``_tp_partition``) and re-assemble (``gather_params``) its
    parameters, and tags them (``tensor_model_parallel``, ``ds_is_replaced_module``) for checkpointing / optimizers."""
    keep_module_on_host: bool = False
    tp_overlap_comm: bool = False      # runtime switch (config ``tensor_parallel.tp_overlap_comm``), set for all layers

    def __init__(self, mp_group, **kwargs: Any):
        super().__init__()
        self.support_training: bool = False
        self.mp_group = mp_group
        self.tp_world_size: int = _tp_world(mp_group)
        self.tp_index: int = dist.get_rank(group=mp_group) if self.tp_world_size > 1 else 0
        self.name = kwargs.get("name")
        self.tp_world_size = kwargs.get("tp_world_size", self.tp_world_size)
        self.tp_index = kwargs.get("tp_index", self.tp_index)

    @classmethod
    def set_keep_module_on_host(cls, value: bool):
        cls.keep_module_on_host = value

    @abstractmethod
    def forward(self, input):
        ...

    @abstractmethod
    def gather_params(self, params_list):
        ...

    @abstractmethod
    def _tp_partition(self, params_list: List[torch.Tensor]):
        ...

    def config_requires_grad(self, weight):
        if weight is not None and self.is_training_mode():
            if weight.requires_grad is None:
                weight.requires_grad = True
        elif weight is not None:
            weight.requires_grad = False

    def config_tp_params(self, weight):
        if weight is None:
            return
        self.config_requires_grad(weight)
        weight.gather_params = self.gather_params
        weight._tp_partition = self._tp_partition
        setattr(weight, "tensor_model_parallel", True)
        weight.ds_is_replaced_module = True
        weight.ds_is_preleace_module = True

    def is_training_mode(self):
        return is_autotp_training_mode() and self.support_training

    def extra_repr(self):
        if getattr(self, "weight", None) is None:
            return ""
        o, i = self.weight.shape[-2:] if self.weight.dim() >= 2 else (None, None)
        return "in_features={}, out_features={}, bias={}, dtype={}".format(i, o, getattr(self, "bias", None) is not None,
                                                                             self.weight.dtype)

    def move(self, tensor):
        """Detach a slice from the full tensor's storage and place it on the accelerator (or keep it on the host)."""
        if tensor.is_meta or self.keep_module_on_host:
            return tensor
        dev = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else tensor.device
        cloned = tensor.detach().clone() if tensor.device == dev else tensor.detach()
        return cloned.to(dev, copy=True) if tensor.device != dev else cloned

    def _slice(self, t: torch.Tensor, dim: int) -> torch.Tensor:
        from .tp_shard import get_shard_size_list
        sizes = get_shard_size_list(t.shape[dim], self.tp_world_size, self.name)
        return torch.split(t, sizes, dim=dim)[self.tp_index]


def configure_tensor_parallel_runtime(config):
    """Apply the runtime part of the ``tensor_parallel`` config block to every TP layer: the grain size of uneven
    splits and ``tp_overlap_comm`` (column-parallel layers overlap the input-gradient all-reduce with the
    weight-gradient GEMM; reference :319-323)."""
    set_autotp_mode(training=True)
    from .tp_shard import set_tp_grain_size
    set_tp_grain_size(getattr(getattr(config, "tp", config), "tp_grain_size", 1))
    TensorParallel_Layer.tp_overlap_comm = bool(getattr(config, "tp_overlap_comm", False))


class GatherReplacedLayerParams:
    """Context that temporarily re-assembles the full tensors of TP-sharded parameters (for saving / inspection) and
    re-partitions them on exit (reference :326-385)."""

    def __init__(self, params: Iterable[torch.Tensor], module: torch.nn.Module, enabled: bool = True) -> None:
        self.enabled = enabled
        self.module = module
        if not enabled:
            return
        if isinstance(params, torch.Tensor):
            params = [params]
        self.params = [p for p in params if getattr(p, "ds_is_replaced_module", False)]
        self.params = self.params if self.params else None

    def __enter__(self) -> None:
        if self.enabled and self.params:
            self.params[0].gather_params(self.params)

    def __exit__(self, exc_type, exc_value, traceback) -> None:
        if self.enabled and self.params:
            self.params[0]._tp_partition(self.params)


class LinearAllreduce(TensorParallel_Layer):
    """Row-parallel linear: the INPUT features are sharded; partial outputs are summed over the TP group; the bias is
    added once, after the reduction."""

    def __init__(self, module, mp_group, **kwargs):
        super().__init__(mp_group, **kwargs)
        self.weight, self.bias = module.weight, module.bias
        self._tp_partition([self.weight, self.bias])
        self.support_training = True
        self.config_tp_params(self.weight)
        if self.bias is not None:
            self.config_requires_grad(self.bias)

    def forward(self, input):
        out = _project(input, self.weight)
        out = RowParallel.apply(self.mp_group, out, not self.is_training_mode())
        return add_bias(out, self.bias)

    @torch.no_grad()
    def gather_params(self, params_list):
        for p in params_list:
            if p is None or p.dim() < 2:
                continue
            parts = [torch.empty_like(p.data) for _ in range(self.tp_world_size)]
            dist.all_gather(parts, p.data.contiguous(), group=self.mp_group)
            p.data = torch.cat(parts, dim=1)

    @torch.no_grad()
    def _tp_partition(self, params_list):
        w = params_list[0]
        if w is not None and not w.is_meta:
            w.data = self.move(self._slice(w.data, 1)).detach()
        b = params_list[1] if len(params_list) > 1 else None
        if b is not None and not b.is_meta:
            b.data = self.move(b.data).detach()


class LinearLayer(TensorParallel_Layer):
    """Column-parallel linear: the OUTPUT features are sharded; no forward communication."""

    def __init__(self, module, mp_group=None, skip_partition=False, **kwargs):
        super().__init__(mp_group, **kwargs)
        self.weight, self.bias = module.weight, module.bias
        if not skip_partition:
            self._tp_partition([self.weight, self.bias])
        self.support_training = True
        self.config_tp_params(self.weight)
        if self.bias is not None:
            self.config_tp_params(self.bias)

    def forward(self, input):
        if type(self).tp_overlap_comm and self.is_training_mode():
            return AsyncColumnParallel.apply(self.mp_group, input, self.weight, self.bias)
        input = ColumnParallel.apply(self.mp_group, input)
        out = _project(input, self.weight)
        return add_bias(out, self.bias)

    @torch.no_grad()
    def gather_params(self, params_list):
        for p in params_list:
            if p is None:
                continue
            parts = [torch.empty_like(p.data) for _ in range(self.tp_world_size)]
            dist.all_gather(parts, p.data.contiguous(), group=self.mp_group)
            p.data = torch.cat(parts, dim=0)

    def _split_rows(self, t: torch.Tensor) -> torch.Tensor:
        return self._slice(t, 0)

    @torch.no_grad()
    def _tp_partition(self, params_list):
        for p in params_list:
            if p is not None and not p.is_meta:
                p.data = self.move(self._split_rows(p.data)).detach()

    @classmethod
    def from_weights(cls, weight_shape=None, dtype=torch.half, weight=None, bias=None):
        if weight is None:
            weight = torch.empty(weight_shape, dtype=dtype)
        holder = nn.Module()
        holder.weight = nn.Parameter(weight, requires_grad=False) if not isinstance(weight, nn.Parameter) else weight
        holder.bias = None if bias is None else (nn.Parameter(bias, requires_grad=False) if not isinstance(bias, nn.Parameter) else bias)
        return cls(holder, skip_partition=True)


class fused_LinearLayer(LinearLayer):
    """Column-parallel over a FUSED projection (q|k|v or gate|up stacked along the output dimension): every component
    is sliced separately so that each rank keeps matching q/k/v rows (reference :559-577 + fusedqkv_utils.py)."""

    def __init__(self, module, mp_group, skip_partition=False, **kwargs):
        self.fused_parts = kwargs.get("fused_parts", 3)
        super().__init__(module, mp_group, skip_partition, **kwargs)

    def _split_rows(self, t: torch.Tensor) -> torch.Tensor:
        parts = torch.chunk(t, self.fused_parts, dim=0)
        return torch.cat([self._slice(p, 0) for p in parts], dim=0)


class GateUpPack_LinearLayer(fused_LinearLayer):
    def __init__(self, module, mp_group, skip_partition=False, **kwargs):
        kwargs["fused_parts"] = 2
        super().__init__(module, mp_group, skip_partition, **kwargs)


class conv_LinearLayer(LinearLayer):
    """GPT-2 ``Conv1D`` ([in, out] weight): transpose, then column-parallel."""

    @torch.no_grad()
    def _tp_partition(self, params_list):
        w, b = params_list[0], params_list[1] if len(params_list) > 1 else None
        if w is not None and not w.is_meta:
            w.data = self.move(self._slice(w.data.transpose(-1, -2).contiguous(), 0)).detach()
        if b is not None and not b.is_meta:
            b.data = self.move(self._slice(b.data, 0)).detach()


class Conv_LinearALlreduce(LinearAllreduce):
    @torch.no_grad()
    def _tp_partition(self, params_list):
        w, b = params_list[0], params_list[1] if len(params_list) > 1 else None
        if w is not None and not w.is_meta:
            w.data = self.move(self._slice(w.data.transpose(-1, -2).contiguous(), 1)).detach()
        if b is not None and not b.is_meta:
            b.data = self.move(b.data).detach()


class Yuan_LinearAllreduce(LinearAllreduce):
    pass


class Yuan_LinearLayer(LinearLayer):
    pass


class LmHeadLinearAllreduce(LinearAllreduce):
    """Vocabulary projection with the HIDDEN dimension sharded: every rank multiplies its slice of the hidden state
    with its slice of the weight columns and the logits are summed."""

    def __init__(self, module, mp_group, **kwargs):
        self.rank = dist.get_rank(mp_group) if _tp_world(mp_group) > 1 else 0
        self.world_size = _tp_world(mp_group)
        super().__init__(module, mp_group, **kwargs)

    def forward(self, input):
        from .tp_shard import get_shard_size, get_shard_size_list
        width = get_shard_size(input.shape[-1], self.world_size, "lm_head")
        start = sum(get_shard_size_list(input.shape[-1], self.world_size, "lm_head")[: self.rank])
        out = _project(input[..., start: start + width], self.weight)
        out = RowParallel.apply(self.mp_group, out, not self.is_training_mode())
        return add_bias(out, self.bias)


class FusedModuleWrapper:
    def __init__(self, fused_module: nn.Module):
        self.fused_module = fused_module

    def __getattr__(self, module):
        return self.fused_module


class TensorParallelConv2d(nn.Module):
    def __init__(self, conv, rank, world_size, shard_by_oc):
        super().__init__()
        self.rank, self.world_size, self.shard_by_oc = rank, world_size, shard_by_oc
        self.shard_weights(conv)

    def shard_weights(self, conv):
        if self.shard_by_oc:
            per = conv.out_channels // self.world_size
            w = conv.weight.data[self.rank * per: (self.rank + 1) * per]
            b = conv.bias.data[self.rank * per: (self.rank + 1) * per] if conv.bias is not None else None
            self.conv = nn.Conv2d(conv.in_channels, per, conv.kernel_size, conv.stride, conv.padding, conv.dilation,
                                  conv.groups, b is not None, conv.padding_mode)
        else:
            per = conv.in_channels // self.world_size
            w = conv.weight.data[:, self.rank * per: (self.rank + 1) * per]
            b = conv.bias.data if conv.bias is not None else None
            self.conv = nn.Conv2d(per, conv.out_channels, conv.kernel_size, conv.stride, conv.padding, conv.dilation,
                                  conv.groups, b is not None and self.rank == 0, conv.padding_mode)
            if self.rank != 0:
                b = None
        self.conv.weight.data.copy_(w)
        if b is not None and self.conv.bias is not None:
            self.conv.bias.data.copy_(b)
        del conv

    def forward(self, input: torch.Tensor) -> torch.Tensor:
        return self.conv(input)


class TensorParallelOcShardConv2d(TensorParallelConv2d):
    def __init__(self, conv, rank, world_size):
        super().__init__(conv, rank, world_size, True)


class TensorParallelIcShardConv2d(TensorParallelConv2d):
    def __init__(self, conv, rank, world_size):
        super().__init__(conv, rank, world_size, False)

    def forward(self, input: torch.Tensor) -> torch.Tensor:
        out = self.conv(input)
        if self.world_size > 1:
            dist.all_reduce(out)
        return out


class Normalize(nn.Module):
    def __init__(self, dim=None, dtype=torch.float, eps=1e-5, weight=None, bias=None):
        super().__init__()
        if weight is not None:
            self.weight, self.bias = weight, bias
        else:
            self.norm = nn.LayerNorm(dim, eps=eps).to(dtype)
            self.weight, self.bias = self.norm.weight, self.norm.bias
        self.eps = eps

    def forward(self, input):
        return TF.layer_norm(input, input.shape[-1:], self.weight, self.bias, eps=self.eps)


class EmbeddingLayer(nn.Module):
    def __init__(self, weight_shape=None, dtype=torch.half, weight=None, bias=None):
        super().__init__()
        self.weight = nn.Parameter(torch.empty(weight_shape[0], weight_shape[1], dtype=dtype)) if weight is None else weight
        self.weight.ds_tensor_parallel = False if not hasattr(self.weight, "ds_tensor_parallel") else self.weight.ds_tensor_parallel

    def forward(self, input):
        return TF.embedding(input, self.weight)


class OPTEmbedding(EmbeddingLayer):
    """Positional embedding of OPT: positions are offset by 2 and derived from the attention mask."""

    def __init__(self, weight_shape=None, weight=None, bias=None):
        self.offset = 2
        super().__init__(weight_shape, weight=weight)

    def forward(self, attention_mask: torch.LongTensor, past_key_values_length: int = 0, position_ids: int = 0):
        attention_mask = attention_mask.long()
        positions = (torch.cumsum(attention_mask, dim=1).type_as(attention_mask) * attention_mask).long() - 1
        positions = positions[:, past_key_values_length:]
        return super().forward(positions + self.offset)


class RMSNormalize(nn.Module):
    def __init__(self, dim=None, dtype=torch.float, eps=1e-5, weight=None):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(dim, dtype=dtype)) if weight is None else weight
        self.eps = eps

    def forward(self, hidden_states):
        if hidden_states.is_cuda and hidden_states.dtype == torch.bfloat16 and self.weight.dtype == torch.bfloat16:
            return F.rms_norm(hidden_states, self.weight, self.eps)[0]
        var = hidden_states.to(torch.float32).pow(2).mean(-1, keepdim=True)
        out = hidden_states * torch.rsqrt(var + self.eps)
        if self.weight.dtype in (torch.float16, torch.bfloat16):
            out = out.to(self.weight.dtype)
        return out * self.weight
