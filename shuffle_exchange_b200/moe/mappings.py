"""Token drop / gather across the tensor-parallel group for MoE with non-expert TP
(reference: deepspeed/moe/mappings.py:17-118): before the expert all-to-all every TP rank keeps a
1/tp slice of the capacity dimension (the activations are replicated across TP), afterwards the
slices are all-gathered back."""
import torch

from ..comm import comm as dist
from ..utils import groups


def _tp():
    g = groups.get_tensor_model_parallel_group()
    return g, (dist.get_world_size(g) if g is not None else 1), (dist.get_rank(g) if g is not None else 0)


def _gather_tokens(x, dim=0):
    g, world, _ = _tp()
    if world == 1:
        return x
    xt = x.transpose(0, dim).contiguous()
    out = torch.empty((world * xt.shape[0],) + tuple(xt.shape[1:]), dtype=x.dtype, device=x.device)
    dist.all_gather_into_tensor(out, xt, group=g)
    return out.transpose(0, dim).contiguous()


def _drop_tokens(x, dim=0):
    g, world, rank = _tp()
    if world == 1:
        return x
    total = x.shape[dim]
    assert total % world == 0, f"input dimension {dim} ({total}) is not divisible by tensor parallel world size ({world})"
    chunk = total // world
    return torch.narrow(x, dim, rank * chunk, chunk).contiguous()


class _GatherTokens(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, dim):
        ctx.dim = dim
        return _gather_tokens(x, dim)

    @staticmethod
    def backward(ctx, grad):
        return _drop_tokens(grad, ctx.dim), None


class _DropTokens(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, dim):
        ctx.dim = dim
        return _drop_tokens(x, dim)

    @staticmethod
    def backward(ctx, grad):
        return _gather_tokens(grad, ctx.dim), None


def gather_tokens(x, dim=0):
    if groups.mpu is None or groups.get_tensor_model_parallel_world_size() == 1:
        return x
    return _GatherTokens.apply(x, dim)


def drop_tokens(x, dim=0):
    if groups.mpu is None or groups.get_tensor_model_parallel_world_size() == 1:
        return x
    return _DropTokens.apply(x, dim)
