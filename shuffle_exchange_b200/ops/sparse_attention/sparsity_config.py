"""Block-sparsity layouts for sparse self-attention.

Same classes, constructor arguments and resulting ``[num_heads, S/block, S/block]`` 0/1 layouts as the reference
(``deepspeed/ops/sparse_attention/sparsity_config.py``: Dense :63, Fixed :95, Variable :239, BigBird :411,
BSLongformer :546, LocalSlidingWindow :674).  Layouts are built with whole-tensor index arithmetic (window ids,
band masks) instead of per-element Python loops; random blocks still use :mod:`random` so a seeded run reproduces.
"""
from __future__ import annotations

import random
from typing import List, Optional, Sequence

import torch

_ATTN = ("unidirectional", "bidirectional")


def _check_attention(attention: str, horizontal: bool = False) -> None:
    if attention not in _ATTN:
        raise NotImplementedError('only "uni/bi-directional" attentions are supported for now!')
    if attention != "bidirectional" and horizontal:
        raise ValueError('only "bi-directional" attentions can support horizontal global attention!')


def _window_mask(window_id: torch.Tensor, causal: bool) -> torch.Tensor:
    """Blocks of the same local window see each other (lower triangle only when causal)."""
    same = window_id[:, None] == window_id[None, :]
    if causal:
        n = window_id.numel()
        same &= torch.ones(n, n, dtype=torch.bool).tril()
    return same


def _band_mask(nb: int, half_width: int, causal_to_diag: bool = False) -> torch.Tensor:
    r = torch.arange(nb)[:, None]
    c = torch.arange(nb)[None, :]
    m = (c >= r - half_width) & (c <= r + half_width)
    if causal_to_diag:
        m &= c <= r
    return m


def _check_span(starts: Sequence[int], ends: Optional[Sequence[int]]) -> None:
    if ends is None:
        return
    if len(starts) != len(ends):
        raise ValueError(f"Global block start indices length, {len(starts)}, must be same as global block end "
                         f"indices length, {len(ends)}!")
    for s, e in zip(starts, ends):
        if s >= e:
            raise ValueError(f"Global block start index, {s}, must be smaller than global block end index, {e}!")


class SparsityConfig:
    """Shared properties: head count, block size, and whether heads may differ."""

    def __init__(self, num_heads, block=16, different_layout_per_head=False):
        self.num_heads = num_heads
        self.block = block
        self.different_layout_per_head = different_layout_per_head
        self.num_layout_heads = num_heads if different_layout_per_head else 1

    def setup_layout(self, seq_len):
        if seq_len % self.block != 0:
            raise ValueError(f"Sequence Length, {seq_len}, needs to be dividable by Block size {self.block}!")
        nb = seq_len // self.block
        return torch.zeros((self.num_heads, nb, nb), dtype=torch.int64)

    def check_and_propagate_first_head_layout(self, layout):
        if not self.different_layout_per_head:
            layout[1:self.num_heads] = layout[0]
        return layout

    # helpers shared by the concrete configs -------------------------------------------------------------
    def _random_blocks(self, layout, h, count, causal=False):
        nb = layout.shape[1]
        if nb < count:
            raise ValueError(f"Number of random blocks, {count}, must be smaller than overall number of blocks in a "
                             f"row, {nb}!")
        for row in range(nb):
            pool = range(0, row + 1) if causal else range(0, nb)
            cols = random.sample(pool, count)
            layout[h, row, cols] = 1
        return layout

    def _sliding_window(self, layout, h, width, causal_to_diag=False):
        nb = layout.shape[1]
        if nb < width:
            raise ValueError(f"Number of sliding window blocks, {width}, must be smaller than overall number of "
                             f"blocks in a row, {nb}!")
        layout[h] |= _band_mask(nb, width // 2, causal_to_diag).to(layout.dtype)
        return layout


class DenseSparsityConfig(SparsityConfig):
    """Every block present (testing / comparison baseline)."""

    def make_layout(self, seq_len):
        layout = self.setup_layout(seq_len)
        layout[:] = 1
        return layout


class FixedSparsityConfig(SparsityConfig):
    """`Generative Modeling with Sparse Transformers` fixed pattern: local windows + per-window global blocks."""

    def __init__(self, num_heads, block=16, different_layout_per_head=False, num_local_blocks=4, num_global_blocks=1,
                 attention="bidirectional", horizontal_global_attention=False, num_different_global_patterns=1):
        super().__init__(num_heads, block, different_layout_per_head)
        if num_local_blocks % num_global_blocks != 0:
            raise ValueError(f"Number of blocks in a local window, {num_local_blocks}, must be dividable by number "
                             f"of global blocks, {num_global_blocks}!")
        _check_attention(attention, horizontal_global_attention)
        if num_different_global_patterns > 1 and not different_layout_per_head:
            raise ValueError("Number of different layouts cannot be more than one when you have set a single layout "
                             "for all heads! Set different_layout_per_head to True.")
        if num_different_global_patterns > num_local_blocks // num_global_blocks:
            raise ValueError(f"Number of layout versions (num_different_global_patterns), "
                             f"{num_different_global_patterns}, cannot be larger than number of local window blocks "
                             f"divided by number of global blocks, {num_local_blocks} / {num_global_blocks} = "
                             f"{num_local_blocks // num_global_blocks}!")
        self.num_local_blocks = num_local_blocks
        self.num_global_blocks = num_global_blocks
        self.attention = attention
        self.horizontal_global_attention = horizontal_global_attention
        self.num_different_global_patterns = num_different_global_patterns

    def set_local_layout(self, h, layout):
        nb = layout.shape[1]
        wid = torch.arange(nb) // self.num_local_blocks
        layout[h] |= _window_mask(wid, self.attention == "unidirectional").to(layout.dtype)
        return layout

    def set_global_layout(self, h, layout):
        nb = layout.shape[1]
        g = self.num_global_blocks
        first = self.num_local_blocks - (1 + h % self.num_different_global_patterns) * g
        full_end = nb - (nb % self.num_local_blocks)
        starts = list(range(first, full_end, self.num_local_blocks))
        if full_end < nb:   # short trailing window: its representative is clamped inside the sequence
            starts.append(min(full_end + first, nb - g))
        for s in starts:
            top = 0 if self.attention == "bidirectional" else s
            layout[h, top:, s:s + g] = 1
            if self.horizontal_global_attention:
                layout[h, s:s + g, :] = 1
        return layout

    def make_layout(self, seq_len):
        layout = self.setup_layout(seq_len)
        for h in range(self.num_layout_heads):
            layout = self.set_local_layout(h, layout)
            layout = self.set_global_layout(h, layout)
        return self.check_and_propagate_first_head_layout(layout)


class VariableSparsityConfig(SparsityConfig):
    """Random + variable-size local windows + explicit global block (ranges)."""

    def __init__(self, num_heads, block=16, different_layout_per_head=False, num_random_blocks=0,
                 local_window_blocks: List[int] = [4], global_block_indices: List[int] = [0],
                 global_block_end_indices: Optional[List[int]] = None, attention="bidirectional",
                 horizontal_global_attention=False):
        super().__init__(num_heads, block, different_layout_per_head)
        _check_span(global_block_indices, global_block_end_indices)
        _check_attention(attention, horizontal_global_attention)
        self.num_random_blocks = num_random_blocks
        self.local_window_blocks = local_window_blocks
        self.global_block_indices = global_block_indices
        self.global_block_end_indices = global_block_end_indices
        self.attention = attention
        self.horizontal_global_attention = horizontal_global_attention

    def set_random_layout(self, h, layout):
        return self._random_blocks(layout, h, self.num_random_blocks)

    def set_local_layout(self, h, layout):
        nb = layout.shape[1]
        # window id per block: the listed sizes first, then the last size repeated to the end
        wid = torch.empty(nb, dtype=torch.int64)
        pos, w = 0, 0
        sizes = list(self.local_window_blocks)
        while pos < nb:
            size = sizes[w] if w < len(sizes) else sizes[-1]
            wid[pos:pos + size] = w
            pos += size
            w += 1
        layout[h] |= _window_mask(wid, self.attention == "unidirectional").to(layout.dtype)
        return layout

    def set_global_layout(self, h, layout):
        nb = layout.shape[1]
        ends = self.global_block_end_indices
        spans = [(s, s + 1) for s in self.global_block_indices] if ends is None else \
            list(zip(self.global_block_indices, ends))
        for s, e in spans:
            if s >= nb:
                continue
            e = min(e, nb)
            if self.horizontal_global_attention:
                layout[h, s:e, :] = 1
            top = 0 if self.attention == "bidirectional" else s
            layout[h, top:, s:e] = 1
        return layout

    def make_layout(self, seq_len):
        layout = self.setup_layout(seq_len)
        for h in range(self.num_layout_heads):
            layout = self.set_random_layout(h, layout)
            layout = self.set_local_layout(h, layout)
            layout = self.set_global_layout(h, layout)
        return self.check_and_propagate_first_head_layout(layout)


class BigBirdSparsityConfig(SparsityConfig):
    """BigBird ITC: random + sliding window + leading global blocks."""

    def __init__(self, num_heads, block=16, different_layout_per_head=False, num_random_blocks=1,
                 num_sliding_window_blocks=3, num_global_blocks=1, attention="bidirectional"):
        super().__init__(num_heads, block, different_layout_per_head)
        _check_attention(attention)
        self.num_random_blocks = num_random_blocks
        self.num_sliding_window_blocks = num_sliding_window_blocks
        self.num_global_blocks = num_global_blocks
        self.attention = attention

    def set_random_layout(self, h, layout):
        return self._random_blocks(layout, h, self.num_random_blocks, causal=self.attention == "unidirectional")

    def set_sliding_window_layout(self, h, layout):
        return self._sliding_window(layout, h, self.num_sliding_window_blocks)

    def set_global_layout_itc(self, h, layout):
        nb = layout.shape[1]
        if nb < self.num_global_blocks:
            raise ValueError(f"Number of global blocks, {self.num_global_blocks}, must be smaller than overall "
                             f"number of blocks in a row, {nb}!")
        layout[h, :self.num_global_blocks, :] = 1
        layout[h, :, :self.num_global_blocks] = 1
        if self.attention == "unidirectional":
            layout = torch.tril(layout)
        return layout

    def make_layout(self, seq_len):
        layout = self.setup_layout(seq_len)
        for h in range(self.num_layout_heads):
            layout = self.set_random_layout(h, layout)
            layout = self.set_sliding_window_layout(h, layout)
            layout = self.set_global_layout_itc(h, layout)
        return self.check_and_propagate_first_head_layout(layout)


class BSLongformerSparsityConfig(SparsityConfig):
    """Block-sparse Longformer: sliding window + symmetric global rows/columns."""

    def __init__(self, num_heads, block=16, different_layout_per_head=False, num_sliding_window_blocks=3,
                 global_block_indices: List[int] = [0], global_block_end_indices: Optional[List[int]] = None,
                 attention="bidirectional"):
        super().__init__(num_heads, block, different_layout_per_head)
        _check_span(global_block_indices, global_block_end_indices)
        self.num_sliding_window_blocks = num_sliding_window_blocks
        self.global_block_indices = global_block_indices
        self.global_block_end_indices = global_block_end_indices
        self.attention = attention

    def set_sliding_window_layout(self, h, layout):
        return self._sliding_window(layout, h, self.num_sliding_window_blocks)

    def set_global_layout(self, h, layout):
        nb = layout.shape[1]
        ends = self.global_block_end_indices
        spans = [(s, s + 1) for s in self.global_block_indices] if ends is None else \
            list(zip(self.global_block_indices, ends))
        for s, e in spans:
            if s < nb:
                e = min(e, nb)
                layout[h, s:e, :] = 1
                layout[h, :, s:e] = 1
        if self.attention == "unidirectional":
            layout = torch.tril(layout)
        return layout

    def make_layout(self, seq_len):
        layout = self.setup_layout(seq_len)
        for h in range(self.num_layout_heads):
            layout = self.set_sliding_window_layout(h, layout)
            layout = self.set_global_layout(h, layout)
        return self.check_and_propagate_first_head_layout(layout)


class LocalSlidingWindowSparsityConfig(SparsityConfig):
    """Purely local sliding window (causal by default)."""

    def __init__(self, num_heads, block=16, num_sliding_window_blocks=3, attention="unidirectional"):
        super().__init__(num_heads, block)
        self.num_sliding_window_blocks = num_sliding_window_blocks
        self.attention = attention

    def set_sliding_window_layout(self, h, layout):
        return self._sliding_window(layout, h, self.num_sliding_window_blocks,
                                    causal_to_diag=self.attention != "bidirectional")

    def make_layout(self, seq_len):
        layout = self.setup_layout(seq_len)
        for h in range(self.num_layout_heads):
            layout = self.set_sliding_window_layout(h, layout)
        return self.check_and_propagate_first_head_layout(layout)
