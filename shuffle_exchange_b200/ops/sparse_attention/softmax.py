"""Block-sparse softmax (reference ``deepspeed/ops/sparse_attention/softmax.py:123-296``).

One CUDA pass per direction (``csrc/softmax_dropout.cu: bsr_softmax_fwd/bwd``): a warp owns one query row, walks the
non-zero blocks of its block-row through a CSR table built once per layout, and applies scale, relative position
embedding, key-padding mask and attention mask on the fly."""
from __future__ import annotations

import torch

from ..fused_softmax import BsrLayout, bsr_softmax


class Softmax:

    def __init__(self, layout: torch.Tensor, block: int, bench: bool = False):
        self.layout = layout
        self.block = int(block)
        self.lay = BsrLayout(layout, block)
        self.num_blocks = self.lay.nnz
        self.spdims = layout.shape
        self.bench = bench

    def __call__(self, x, scale=1., rpe=None, key_padding_mask=None, attn_mask=None, key_padding_mask_mode='add',
                 attn_mask_mode='add'):
        if rpe is not None and rpe.dtype != x.dtype:
            raise ValueError('relative position embedding must be %s' % x.dtype)
        if attn_mask is not None and attn_mask.dtype != x.dtype:
            raise ValueError('Attention mask must be %s' % x.dtype)
        if key_padding_mask is not None and key_padding_mask.dtype != x.dtype:
            raise ValueError('Key padding mask must be %s' % x.dtype)
        for mode in (key_padding_mask_mode, attn_mask_mode):
            if mode not in ('add', 'mul'):
                raise ValueError("mask modes are 'add' or 'mul'")
        return bsr_softmax(x, self.lay, scale, rpe, key_padding_mask, attn_mask, key_padding_mask_mode,
                           attn_mask_mode)
