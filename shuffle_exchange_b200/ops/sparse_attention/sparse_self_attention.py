"""``SparseSelfAttention``: QK^T on the non-zero blocks, block-sparse softmax, probabilities x V.

Reference: ``deepspeed/ops/sparse_attention/sparse_self_attention.py:12-149`` (same constructor, master layout buffer
broadcast from rank 0 on first use, per-sequence-length op cache, mask squeezing rules).  fp16, bf16 and fp32 inputs
are accepted (the reference asserts fp16)."""
from __future__ import annotations

import torch
import torch.nn as nn
from torch import distributed as dist

from .sparsity_config import SparsityConfig
from .matmul import MatMul
from .softmax import Softmax


class SparseSelfAttention(nn.Module):
    ops = dict()

    def __init__(self, sparsity_config=SparsityConfig(num_heads=4), key_padding_mask_mode='add',
                 attn_mask_mode='mul', max_seq_length=2048):
        super().__init__()
        self.sparsity_config = sparsity_config
        self.register_buffer("master_layout", self.sparsity_config.make_layout(max_seq_length))
        self._need_layout_synchronization = True
        self.key_padding_mask_mode = key_padding_mask_mode
        self.attn_mask_mode = attn_mask_mode

    def get_layout(self, L):
        if self._need_layout_synchronization and dist.is_available() and dist.is_initialized():
            # random layouts must agree across ranks
            buf = self.master_layout
            if dist.get_backend() == "nccl" and not buf.is_cuda:
                tmp = buf.cuda()
                dist.broadcast(tmp, src=0)
                buf.copy_(tmp.cpu())
            else:
                dist.broadcast(buf, src=0)
            self._need_layout_synchronization = False
        block = self.sparsity_config.block
        if L % block != 0:
            raise ValueError(f'Sequence Length, {L}, needs to be dividable by Block size {block}!')
        nb = L // block
        return self.master_layout[..., :nb, :nb].cpu()

    def get_ops(self, H, L):
        key = (id(self), L)
        if key not in SparseSelfAttention.ops:
            layout = self.get_layout(L)
            block = self.sparsity_config.block
            SparseSelfAttention.ops[key] = (MatMul(layout, block, 'sdd', trans_a=False, trans_b=True),
                                            MatMul(layout, block, 'dsd', trans_a=False, trans_b=False),
                                            Softmax(layout, block))
        return SparseSelfAttention.ops[key]

    def transpose_key_for_scores(self, x, L):
        if x.size(2) != L:
            return x.permute(0, 1, 3, 2)
        return x

    def transpose_mask_for_sparse(self, qtype, x, is_key_padding_mask=False):
        x = x.type(qtype)
        if is_key_padding_mask:
            for d in range(x.dim() - 1, 0, -1):
                x = x.squeeze(dim=d)
            return x
        return x.squeeze()

    def forward(self, query, key, value, rpe=None, key_padding_mask=None, attn_mask=None):
        """query/key/value: ``[batch, heads, seq, head_dim]`` -> context of the same shape."""
        bsz, num_heads, tgt_len, head_dim = query.size()
        key = self.transpose_key_for_scores(key, tgt_len)
        if query.shape != key.shape or key.shape != value.shape:
            raise NotImplementedError('only self-attention is supported for now')
        if key_padding_mask is not None:
            key_padding_mask = self.transpose_mask_for_sparse(query.dtype, key_padding_mask, True)
        if attn_mask is not None:
            attn_mask = self.transpose_mask_for_sparse(query.dtype, attn_mask)
        sdd_nt, dsd_nn, softmax = self.get_ops(num_heads, tgt_len)
        scores = sdd_nt(query, key)
        probs = softmax(scores, scale=float(head_dim) ** -0.5, rpe=rpe, key_padding_mask=key_padding_mask,
                        attn_mask=attn_mask, key_padding_mask_mode=self.key_padding_mask_mode,
                        attn_mask_mode=self.attn_mask_mode)
        return dsd_nn(probs, value)
