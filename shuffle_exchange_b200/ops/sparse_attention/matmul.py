"""Block-sparse matrix multiplication: ``sdd`` (dense x dense -> sparse), ``dsd`` (sparse x dense -> dense) and
``dds`` (dense x sparse -> dense).

Reference: ``deepspeed/ops/sparse_attention/matmul.py:196-819`` (legacy-Triton kernels + look-up tables).  Here the
non-zero blocks are gathered into one ``[B, nnz, block, *]`` batch and multiplied by a single batched tensor-core GEMM;
results of ``dsd``/``dds`` are scattered back with an ``index_add`` over (head, block-row) or (head, block-col).  All
steps are differentiable, so the three backward products of the reference (:632-700) follow from autograd.

Sparse operands use the reference's storage: ``[B, nnz, block, block]`` with blocks in ``layout.nonzero()`` order.
"""
from __future__ import annotations

import torch

from ..fused_softmax import BsrLayout


class MatMul:

    def __init__(self, layout: torch.Tensor, block: int, mode: str, trans_a: bool = False, trans_b: bool = False,
                 bench: bool = False):
        if mode not in ("sdd", "dsd", "dds"):
            raise NotImplementedError("Supported modes are: sdd, dsd, dds")
        self.mode, self.block = mode, int(block)
        self.trans_a, self.trans_b = trans_a, trans_b
        self.layout = layout
        self.lay = BsrLayout(layout, block)
        self.spdims = layout.shape
        self.bench = bench

    # -- helpers ------------------------------------------------------------------------------------------
    def _blocks_rows(self, x: torch.Tensor, n: int) -> torch.Tensor:
        """[B, H, n*blk, K] -> [B, H, n, blk, K]"""
        B, H, S, K = x.shape
        return x.reshape(B, H, n, self.block, K)

    def _pad_batch(self, x: torch.Tensor) -> torch.Tensor:
        while x.dim() < 4:
            x = x.unsqueeze(0)
        return x

    def __call__(self, a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
        lay = self.lay
        _, _, hi, ri, ci = lay.on(a.device)
        blk = self.block
        if self.mode == "sdd":
            a = self._pad_batch(a)
            b = self._pad_batch(b)
            am = a.transpose(-1, -2) if self.trans_a else a            # [B,H,M,K]
            bm = b if self.trans_b else b.transpose(-1, -2)            # [B,H,N,K]
            ag = self._blocks_rows(am, lay.nrb)[:, hi, ri]             # [B,nnz,blk,K]
            bg = self._blocks_rows(bm, lay.ncb)[:, hi, ci]             # [B,nnz,blk,K]
            return torch.matmul(ag, bg.transpose(-1, -2))              # [B,nnz,blk,blk]
        if self.mode == "dsd":
            b = self._pad_batch(b)
            bm = b.transpose(-1, -2) if self.trans_b else b            # [B,H,Kdim,N]
            B_, H, _, N = bm.shape
            if self.trans_a:      # A^T: block (r,c) of the layout acts as block (c,r), transposed
                blocks, out_idx, in_idx, n_out, n_in = a.transpose(-1, -2), ci, ri, lay.ncb, lay.nrb
            else:
                blocks, out_idx, in_idx, n_out, n_in = a, ri, ci, lay.nrb, lay.ncb
            bg = self._blocks_rows(bm, n_in)[:, hi, in_idx]            # [B,nnz,blk,N]
            prod = torch.matmul(blocks, bg)                            # [B,nnz,blk,N]
            out = torch.zeros(B_, H * n_out, blk, N, dtype=prod.dtype, device=prod.device)
            out = out.index_add(1, hi * n_out + out_idx, prod)
            return out.reshape(B_, H, n_out * blk, N)
        # dds
        a = self._pad_batch(a)
        am = a.transpose(-1, -2) if self.trans_a else a                # [B,H,M,Kdim]
        B_, H, M, _ = am.shape
        if self.trans_b:
            blocks, out_idx, in_idx, n_out, n_in = b.transpose(-1, -2), ri, ci, lay.nrb, lay.ncb
        else:
            blocks, out_idx, in_idx, n_out, n_in = b, ci, ri, lay.ncb, lay.nrb
        ag = am.reshape(B_, H, M, n_in, blk).permute(0, 1, 3, 2, 4)[:, hi, in_idx]   # [B,nnz,M,blk]
        prod = torch.matmul(ag, blocks)                                # [B,nnz,M,blk]
        out = torch.zeros(B_, H * n_out, M, blk, dtype=prod.dtype, device=prod.device)
        out = out.index_add(1, hi * n_out + out_idx, prod)
        return out.reshape(B_, H, n_out, M, blk).permute(0, 1, 3, 2, 4).reshape(B_, H, M, n_out * blk)
