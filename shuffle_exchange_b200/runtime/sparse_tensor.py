"""Row-sparse gradient container and its data-parallel reduction.

Reference: deepspeed/runtime/sparse_tensor.py:13-68 (``SparseTensor``: non-zero row ``indices`` + ``values``)
and engine.py:2763-2824 (``sparse_allreduce``: every rank all-gathers its indices and values, padded to the
group-wide max row count, and concatenates them; duplicates are summed when densified).
"""
from __future__ import annotations

import torch

from .. import comm as dist


class SparseTensor:
    """Rows of a 2-D tensor that are not identically zero."""

    def __init__(self, dense_tensor: torch.Tensor = None):
        self.orig_dense_tensor = dense_tensor
        self.indices = self.values = self.dense_size = None
        self.dtype = None if dense_tensor is None else dense_tensor.dtype
        self.is_sparse = False if dense_tensor is None else dense_tensor.is_sparse
        if dense_tensor is None:
            return
        if dense_tensor.is_sparse:
            coo = dense_tensor.coalesce()
            self.indices, self.values = coo.indices().flatten(), coo.values()
        else:
            self.indices = dense_tensor.abs().sum(dim=1).nonzero().flatten()
            self.values = dense_tensor.index_select(0, self.indices)
        self.dense_size = list(dense_tensor.size())

    @staticmethod
    def type() -> str:
        return "deepspeed.SparseTensor"

    def to_coo_tensor(self):
        return torch.sparse_coo_tensor(self.indices.unsqueeze(0), self.values, self.dense_size)

    def to_dense(self):
        out = self.values.new_zeros(self.dense_size)
        return out.index_add_(0, self.indices, self.values)

    def sparse_size(self):
        return self.indices.numel() + self.values.numel(), self.dense_size[0] * self.dense_size[1]

    def add(self, other: "SparseTensor"):
        assert self.dense_size == other.dense_size
        self.indices = torch.cat([self.indices, other.indices])
        self.values = torch.cat([self.values, other.values])

    def __repr__(self):
        sp, dn = self.sparse_size()
        return (f"DeepSpeed.SparseTensor(indices_size={tuple(self.indices.size())}, values_size={tuple(self.values.size())}, "
                f"dense_size={self.dense_size}, device={self.indices.device}, reduction_factor={dn / max(sp, 1):.2f})")

    __str__ = __repr__


def _all_gather_padded(t: torch.Tensor, group) -> list:
    """All-gather tensors whose dim-0 differs across ranks (pad to the max, trim on arrival)."""
    world = dist.get_world_size(group=group)
    n = torch.tensor([t.shape[0]], dtype=torch.long, device=t.device)
    sizes = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(sizes, n, group=group)
    sizes = [int(s.item()) for s in sizes]
    mx = max(sizes)
    if t.shape[0] < mx:
        pad = t.new_zeros((mx - t.shape[0],) + tuple(t.shape[1:]))
        t = torch.cat([t, pad])
    outs = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(outs, t.contiguous(), group=group)
    return [o[:s] for o, s in zip(outs, sizes)]


def sparse_allreduce(grad: torch.Tensor, group, postscale: bool = True, predivide_factor: float = 1.0) -> torch.Tensor:
    """Average a (torch-sparse or row-sparse dense) gradient over ``group`` and write it back densely
    into ``grad`` (or return the dense result when ``grad`` is a torch sparse tensor)."""
    world = dist.get_world_size(group=group)
    sp = SparseTensor(grad)
    vals = sp.values.float()
    if postscale:
        vals = vals / (predivide_factor if predivide_factor != 1.0 else 1.0)
    idx = torch.cat(_all_gather_padded(sp.indices, group))
    val = torch.cat(_all_gather_padded(vals, group))
    if postscale:
        val = val * (predivide_factor / world if predivide_factor != 1.0 else 1.0 / world)
    else:
        val = val / world
    sp.indices, sp.values = idx, val.to(sp.dtype)
    dense = sp.to_dense()
    if not grad.is_sparse:
        grad.copy_(dense)
        return grad
    return dense
