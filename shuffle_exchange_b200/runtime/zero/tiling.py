"""``TiledLinear``: split one huge linear layer into ``in_splits x out_splits`` independent ``nn.Linear`` tiles so
that ZeRO-3 gathers (and frees) one tile at a time (reference: ``runtime/zero/tiling.py:12-296``)."""
from __future__ import annotations

import torch

from ...runtime.utils import partition_uniform as partition


def split_tensor_along_last_dim(tensor, partitions, contiguous_split_chunks=False):
    """``partitions``: boundary list of length n+1.  Returns n views (or contiguous copies)."""
    sizes = [partitions[i + 1] - partitions[i] for i in range(len(partitions) - 1)]
    chunks = torch.split(tensor, sizes, dim=tensor.dim() - 1)
    return tuple(c.contiguous() for c in chunks) if contiguous_split_chunks else chunks


class TiledLinear(torch.nn.Module):
    def __init__(self, in_features, out_features, bias=True, in_splits=1, out_splits=1, input_is_already_split=False,
                 combine_out_splits=True, linear_cls=torch.nn.Linear, init_linear=None, **kwargs):
        super().__init__()
        if not (1 <= in_splits <= in_features):
            raise RuntimeError("in splits must be in range [1, in_features].")
        if not (1 <= out_splits <= out_features):
            raise RuntimeError("out splits must be in range [1, out_features].")
        self.in_features, self.out_features, self.use_bias = in_features, out_features, bias
        self.in_splits, self.out_splits = in_splits, out_splits
        self.input_is_already_split, self.combine_out_splits = input_is_already_split, combine_out_splits
        self.in_parts = partition(num_items=in_features, num_parts=in_splits)
        self.out_parts = partition(num_items=out_features, num_parts=out_splits)
        assert len(self.out_parts) == out_splits + 1 and len(self.in_parts) == in_splits + 1
        assert self.out_parts[0] == 0 and self.out_parts[out_splits] == out_features
        assert self.in_parts[in_splits] == in_features
        self.linears = torch.nn.ModuleList()
        for o in range(out_splits):
            row = torch.nn.ModuleList()
            for i in range(in_splits):
                # only the last tile of a row adds the bias, so it is applied once per output element
                row.append(linear_cls(self.in_parts[i + 1] - self.in_parts[i], self.out_parts[o + 1] - self.out_parts[o],
                                      bias=bias if i == in_splits - 1 else False, **kwargs))
            self.linears.append(row)
        if init_linear is not None:
            self.copy_params_from(init_linear)

    def forward(self, input_):
        if self.in_splits > 1 and not self.input_is_already_split:
            inputs = self._split_global_input(input_)
        elif self.in_splits > 1:
            inputs = input_
            assert len(inputs) == self.in_splits, f"Col splits {self.in_splits} does not match input splits {len(inputs)}"
        else:
            inputs = [input_]
        outputs = [None] * self.out_splits
        for o in range(self.out_splits):
            for i in range(self.in_splits):
                local = self.linears[o][i](inputs[i])
                outputs[o] = self._reduce_local_output(i, o, outputs[o], local)
        return self._combine_output_splits(outputs) if self.combine_out_splits else outputs

    def _split_global_input(self, input, split_sizes=None):
        return split_tensor_along_last_dim(input, self.in_parts)

    def _reduce_local_output(self, in_id, out_id, current_out, new_out):
        return new_out if current_out is None else current_out + new_out

    def _combine_output_splits(self, outputs):
        assert len(outputs) == self.out_splits
        return torch.cat(outputs, dim=-1)

    @torch.no_grad()
    def copy_params_from(self, other):
        """Slice a full ``nn.Linear`` into the tiles."""
        assert hasattr(other, "weight") and other.weight.size() == (self.out_features, self.in_features)
        if self.use_bias:
            assert other.bias is not None and other.bias.size() == (self.out_features,)
        else:
            assert other.bias is None
        for o in range(self.out_splits):
            r0, r1 = self.out_parts[o], self.out_parts[o + 1]
            for i in range(self.in_splits):
                c0, c1 = self.in_parts[i], self.in_parts[i + 1]
                tile = self.linears[o][i]
                tile.weight.copy_(other.weight[r0:r1, c0:c1])
                if tile.bias is not None:
                    tile.bias.copy_(other.bias[r0:r1])


class TiledLinearReturnBias(TiledLinear):
    """Tiles whose forward returns ``(output, bias)`` (Megatron-style layers)."""

    def _reduce_local_output(self, in_id, out_id, current_out, new_out):
        if current_out is None:
            return new_out
        old_t, old_b = current_out
        new_t, new_b = new_out
        assert old_t is not None and new_t is not None
        t = old_t + new_t
        if old_b is None:
            b = new_b
        elif new_b is None:
            b = old_b
        else:
            b = old_b + new_b
        return t, b

    def _combine_output_splits(self, outputs):
        tensors = [o[0] for o in outputs]
        tensor = super()._combine_output_splits(tensors)
        biases = [o[1] for o in outputs if o[1] is not None]
        bias = super(TiledLinearReturnBias, self)._combine_output_splits(biases) if biases else None
        return tensor, bias
