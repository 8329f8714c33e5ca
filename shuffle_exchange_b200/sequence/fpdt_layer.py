"""FPDT — fully pipelined distributed transformer blocks for multi-million-token sequences.

Reference: ``deepspeed/sequence/fpdt_layer.py`` — load-balanced chunk assignment (``FPDT_InputConstruct`` :79-131),
chunked Ulysses attention with online-softmax merging and optional host offload of the chunk tensors
(``_FPDTGPUAttentionImpl_`` :134-459, ``_FPDTGPUOffloadingAttentionImpl_`` :510-968, ``FPDT_Attention`` :971-1041),
chunked FFN (``FPDT_FFN`` :1056-1134) and chunked logits + loss (``FPDT_LogitsLoss`` :1137-1225).

B200 design
-----------
* Pair kernels: every (q-chunk, kv-chunk) product is one launch of the TMA/tcgen05 flash-attention kernels through
  the raw ``attention_fwd_lse`` / ``attention_bwd_raw`` entry points; partial results are merged with the
  log-sum-exp rule.  The backward uses the FINAL merged ``out``/``lse`` of a q chunk for every pair it took part in,
  so no gradient ever flows through the merge itself and each pair is again one kernel launch.
* The chunk loop is software-pipelined over two CUDA streams: the all-to-all of chunk ``i+1`` (NCCL over NVLink) and
  the host<->device staging of offloaded chunks run on a side stream while the tensor cores work on chunk ``i``.
* With 180 GB of HBM per GPU, offloading is only needed past ~1M tokens/GPU; ``enable_offloading=False`` keeps every
  chunk resident and skips the PCIe traffic altogether.
* QKV projection, rotary embedding and the head/sequence all-to-all stay ordinary differentiable ops (tcgen05 GEMM,
  fused RoPE kernel, ``_SeqAllToAll``), evaluated chunk by chunk so no full-sequence QKV tensor is ever alive.
"""
from __future__ import annotations

import math
from typing import Any, List, Optional, Tuple

import torch
import torch.distributed as dist
from torch import Tensor

from ..ops import functional as F
from .layer import _SeqAllToAll, apply_rotary_pos_emb


# ------------------------------------------------------------------------------------------ small helpers
def _rotate_half_backward(x):
    x = x.reshape(*x.shape[:-1], 2, x.shape[-1] // 2)
    x1, x2 = x.unbind(dim=-2)
    return torch.cat((x2, -x1), dim=-1)


def apply_rotary_pos_emb_backward(grad_output, freqs_cos, freqs_sin):
    """Gradient of ``apply_rotary_pos_emb`` w.r.t. its input (the rotation is orthogonal: apply the inverse)."""
    rot_dim = freqs_cos.shape[-1]
    grad, grad_pass = grad_output[..., :rot_dim], grad_output[..., rot_dim:]
    grad_t = (grad * freqs_cos) + (_rotate_half_backward(grad * freqs_sin))
    return grad_t if grad_pass.shape[-1] == 0 else torch.cat((grad_t, grad_pass), dim=-1)


def update_out_and_lse(out: Optional[Tensor], lse: Optional[Tensor], block_out: Tensor,
                       block_lse: Tensor) -> Tuple[Tensor, Tensor]:
    """Merge one partial attention result into the running ``(out fp32 [B,S,H,D], lse fp32 [B,H,S])``."""
    block_out = block_out.float()
    if out is None:
        return block_out, block_lse.float()
    new_lse = torch.logaddexp(lse, block_lse)
    w_old = torch.exp(lse - new_lse).transpose(1, 2).unsqueeze(-1)          # [B,S,H,1]
    w_new = torch.exp(block_lse - new_lse).transpose(1, 2).unsqueeze(-1)
    return out * w_old + block_out * w_new, new_lse


class FPDT_InputConstruct(torch.nn.Module):
    """Re-orders the global sequence so that every rank owns chunks spread evenly over the sequence: chunk ``c`` of the
    ``num_chunk_per_gpu * sp_size`` chunks goes to rank ``c % sp_size`` (causal work is then balanced)."""

    def __init__(self, tokens, labels, loss_mask, attention_mask, position_ids, args, sp_size, sp_rank) -> None:
        super().__init__()
        self.tokens, self.labels, self.loss_mask = tokens, labels, loss_mask
        self.attention_mask, self.position_ids = attention_mask, position_ids
        global_seq_len = tokens.shape[1]
        chunk_cfg = args.ds_sequence_parallel_fpdt_chunk_size
        assert global_seq_len % sp_size == 0
        assert global_seq_len % chunk_cfg == 0
        num_chunk_per_gpu = global_seq_len // chunk_cfg
        local_seq_len = global_seq_len // sp_size
        assert local_seq_len % num_chunk_per_gpu == 0
        self.num_chunk_per_gpu = num_chunk_per_gpu
        self.chunk_size = local_seq_len // num_chunk_per_gpu
        self.sp_size, self.sp_rank = sp_size, sp_rank
        self.global_seq_len, self.local_seq_len = global_seq_len, local_seq_len
        self.batch_size = tokens.shape[0]
        self.device = tokens.device

    def _order(self) -> Tensor:
        """Token indices in load-balanced order for ALL ranks ([sp_size * num_chunk_per_gpu, chunk_size])."""
        total = self.global_seq_len // self.chunk_size
        chunk_ids = torch.arange(total, device=self.device).reshape(self.num_chunk_per_gpu, self.sp_size).t()
        offs = torch.arange(self.chunk_size, device=self.device)
        return (chunk_ids.reshape(-1, 1) * self.chunk_size + offs).long()

    def generate(self):
        order = self._order()
        flat_all = order.flatten()
        mine = order[self.num_chunk_per_gpu * self.sp_rank:self.num_chunk_per_gpu * (self.sp_rank + 1)].flatten()
        pick = lambda t, idx: t[:, idx] if t is not None else None
        return (pick(self.tokens, mine), pick(self.labels, mine), pick(self.loss_mask, flat_all),
                self.attention_mask, pick(self.position_ids, mine))


# ------------------------------------------------------------------------------------------ host staging
class SequenceChunk:
    """One chunk tensor that can live on the device or in pinned host memory (double-buffered by the caller)."""

    def __init__(self, chunk: Tensor, device=None, is_in_use: bool = False):
        self.device = chunk.device if device is None else device
        self.shape, self.dtype = chunk.shape, chunk.dtype
        self.gpu_chunk: Optional[Tensor] = chunk if is_in_use or not chunk.is_cuda else None
        self.cpu_chunk: Optional[Tensor] = None
        if self.gpu_chunk is None:
            self.cpu_chunk = torch.empty(chunk.shape, dtype=chunk.dtype, device="cpu", pin_memory=True)
            self.cpu_chunk.copy_(chunk, non_blocking=True)

    def load_to_gpu(self):
        if self.gpu_chunk is None:
            self.gpu_chunk = torch.empty(self.shape, dtype=self.dtype, device=self.device)
            self.gpu_chunk.copy_(self.cpu_chunk, non_blocking=True)

    def get_gpu_chunk(self) -> Tensor:
        assert self.gpu_chunk is not None and self.gpu_chunk.device == self.device
        return self.gpu_chunk

    def check_gpu_chunk(self) -> bool:
        assert (self.gpu_chunk is not None) and (self.gpu_chunk.device == self.device), \
            f"gpu chunk {self.gpu_chunk is not None} on {self.device}"
        return True

    def offload(self):
        if self.cpu_chunk is not None:
            self.gpu_chunk = None

    def overwrite_to_cpu(self):
        assert self.gpu_chunk is not None and self.cpu_chunk is not None
        self.cpu_chunk.copy_(self.gpu_chunk, non_blocking=True)


class _ChunkStore:
    """List of chunk tensors, resident or staged through pinned host memory on a side stream."""

    def __init__(self, offload: bool):
        self.offload = offload
        self.items: List[Any] = []
        self.stream = torch.cuda.Stream() if offload and torch.cuda.is_available() else None

    def append(self, t: Tensor):
        if self.offload and t.is_cuda:
            self.stream.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(self.stream):
                t.record_stream(self.stream)
                self.items.append(SequenceChunk(t))
        else:
            self.items.append(t)

    def prefetch(self, i: int):
        it = self.items[i]
        if isinstance(it, SequenceChunk) and it.gpu_chunk is None:
            with torch.cuda.stream(self.stream):
                it.load_to_gpu()

    def get(self, i: int) -> Tensor:
        it = self.items[i]
        if isinstance(it, SequenceChunk):
            self.prefetch(i)
            torch.cuda.current_stream().wait_stream(self.stream)
            return it.get_gpu_chunk()
        return it

    def release(self, i: int):
        it = self.items[i]
        if isinstance(it, SequenceChunk):
            it.offload()

    def __len__(self):
        return len(self.items)


# ------------------------------------------------------------------------------------------ chunked attention core
class _FPDTChunkedAttention(torch.autograd.Function):
    """Causal attention over ``n`` equally sized global chunks: chunk ``i`` attends chunks ``0..i`` (the diagonal pair
    causally).  Inputs/outputs are flat tuples ``q_0..q_{n-1}, k_0.., v_0..`` of ``[B, Lc, H, D]`` tensors."""

    @staticmethod
    def forward(ctx, scale: float, offload: bool, n: int, *qkv: Tensor):
        qs, ks, vs = qkv[:n], qkv[n:2 * n], qkv[2 * n:]
        kstore, vstore, qstore = _ChunkStore(offload), _ChunkStore(offload), _ChunkStore(offload)
        outs, lses = [], []
        for i in range(n):
            kstore.append(ks[i].detach())
            vstore.append(vs[i].detach())
            out, lse = None, None
            for j in range(i + 1):
                if j + 1 <= i:
                    kstore.prefetch(j + 1)
                    vstore.prefetch(j + 1)
                kj = ks[j] if j == i else kstore.get(j)
                vj = vs[j] if j == i else vstore.get(j)
                o, l = F.attention_fwd_lse(qs[i], kj, vj, causal=(j == i), scale=scale)
                out, lse = update_out_and_lse(out, lse, o, l)
                if j != i:
                    kstore.release(j)
                    vstore.release(j)
            outs.append(out.to(qs[i].dtype))
            lses.append(lse)
            qstore.append(qs[i].detach())
        ctx.scale, ctx.n = scale, n
        ctx.q, ctx.k, ctx.v = qstore, kstore, vstore
        ctx.save_for_backward(*outs, *lses)
        return tuple(outs)

    @staticmethod
    def backward(ctx, *douts: Tensor):
        n, scale = ctx.n, ctx.scale
        saved = ctx.saved_tensors
        outs, lses = saved[:n], saved[n:]
        dq: List[Optional[Tensor]] = [None] * n
        dk: List[Optional[Tensor]] = [None] * n
        dv: List[Optional[Tensor]] = [None] * n
        # kv-major order: dK_j / dV_j are finished after one sweep over q chunks i >= j (one resident accumulator each)
        for j in range(n):
            kj, vj = ctx.k.get(j), ctx.v.get(j)
            acc_k = torch.zeros(kj.shape, dtype=torch.float32, device=kj.device)
            acc_v = torch.zeros(vj.shape, dtype=torch.float32, device=vj.device)
            for i in range(j, n):
                if i + 1 < n:
                    ctx.q.prefetch(i + 1)
                qi = ctx.q.get(i)
                gq, gk, gv = F.attention_bwd_raw(qi, kj, vj, outs[i], douts[i].contiguous(), lses[i],
                                                 causal=(i == j), scale=scale)
                dq[i] = gq.float() if dq[i] is None else dq[i] + gq.float()
                acc_k += gk.float()
                acc_v += gv.float()
                if i != j:
                    ctx.q.release(i)
            dk[j], dv[j] = acc_k.to(kj.dtype), acc_v.to(vj.dtype)
            ctx.k.release(j)
            ctx.v.release(j)
        dq = [g.to(outs[i].dtype) for i, g in enumerate(dq)]
        return (None, None, None, *dq, *dk, *dv)


def fpdt_chunked_attention(qs: List[Tensor], ks: List[Tensor], vs: List[Tensor], scale: float,
                           offload: bool = False) -> List[Tensor]:
    n = len(qs)
    return list(_FPDTChunkedAttention.apply(scale, offload, n, *qs, *ks, *vs))


class FPDT_Attention(torch.nn.Module):
    """Chunked Ulysses attention block: QKV projection, rotary, head-scatter all-to-all, chunk-pair flash attention,
    sequence-scatter all-to-all and output projection — every stage evaluated per chunk.

    ``layernorm_output`` is ``[s_local, b, h]`` (Megatron layout, as in the reference); returns ``(output, bias)``."""

    def __init__(self, config, first_weight, first_bias, second_weight, second_bias, sequence_process_group,
                 gather_idx: int = 0, scatter_idx: int = 2, return_bias=True, chunk_size=65536,
                 enable_offloading=True) -> None:
        super().__init__()
        self.spg = sequence_process_group
        self.scatter_idx, self.gather_idx = scatter_idx, gather_idx
        self.config = config
        self.projection_size = config.kv_channels * config.num_attention_heads
        self.hidden_size_per_attention_head = self.projection_size // config.num_attention_heads
        self.kv_projection_size = config.kv_channels * config.num_key_value_heads
        self.hidden_size = config.hidden_size
        self.qkv_linear_weight, self.qkv_linear_bias = first_weight, first_bias
        self.qkv_dense_weight, self.qkv_dense_bias = second_weight, second_bias
        self.reture_bias = return_bias
        self.dropout = getattr(config, "attention_dropout", 0.0)
        self.chunk_size = chunk_size
        self.double_buffer = enable_offloading

    def _world(self) -> int:
        return dist.get_world_size(self.spg) if (dist.is_available() and dist.is_initialized()) else 1

    def forward(self, layernorm_output, attention_mask, inference_params, rotary_pos_emb,
                cpu_offloading=True) -> Tuple[Tensor, Optional[Tensor]]:
        assert attention_mask is None, "FPDT attention is causal; an explicit mask is not supported"
        sp = self._world()
        s_local, b, _ = layernorm_output.shape
        num_chunks = max(1, s_local * sp // self.chunk_size)
        assert s_local % num_chunks == 0
        lc = s_local // num_chunks
        hd = self.hidden_size_per_attention_head
        ps, kps = self.projection_size, self.kv_projection_size
        cos = sin = None
        if rotary_pos_emb is not None:
            cos, sin = rotary_pos_emb[0].permute(1, 0, 2, 3), rotary_pos_emb[1].permute(1, 0, 2, 3)

        def a2a(t):   # [b, l_local, nh, hd] -> [b, l_local * sp, nh / sp, hd]
            return _SeqAllToAll.apply(self.spg, t, self.scatter_idx, 1, 0) if sp > 1 else t

        qs, ks, vs = [], [], []
        for i in range(num_chunks):
            x = layernorm_output[i * lc:(i + 1) * lc]
            qkv = F.linear(x, self.qkv_linear_weight, self.qkv_linear_bias)             # [lc, b, ps + 2 kps]
            split = lambda t: t.reshape(lc, b, -1, hd).permute(1, 0, 2, 3).contiguous()  # b, l, nh, hd
            q = a2a(split(qkv[..., :ps]))
            k = a2a(split(qkv[..., ps:ps + kps]))
            v = a2a(split(qkv[..., ps + kps:]))
            if cos is not None:
                g = q.shape[1]
                q = apply_rotary_pos_emb(q, cos[:, g * i:g * (i + 1)], sin[:, g * i:g * (i + 1)])
                k = apply_rotary_pos_emb(k, cos[:, g * i:g * (i + 1)], sin[:, g * i:g * (i + 1)])
            qs.append(q)
            ks.append(k)
            vs.append(v)
        offload = bool(cpu_offloading and self.double_buffer and num_chunks > 1 and layernorm_output.is_cuda)
        outs = fpdt_chunked_attention(qs, ks, vs, hd ** -0.5, offload)
        back = []
        for o in outs:   # [b, l_global_chunk, nh/sp, hd] -> [b, lc, nh, hd]
            back.append(_SeqAllToAll.apply(self.spg, o, 1, self.scatter_idx, 0) if sp > 1 else o)
        output = torch.cat(back, dim=1)                                                   # b, s_local, nh, hd
        output = output.flatten(2).permute(1, 0, 2).contiguous()                          # s_local, b, ps
        output = F.linear(output, self.qkv_dense_weight, None if self.reture_bias else self.qkv_dense_bias)
        return output, (self.qkv_dense_bias if self.reture_bias else None)


# ------------------------------------------------------------------------------------------ chunked FFN
def bias_gelu(x):
    return x * 0.5 * (1.0 + torch.tanh(0.79788456 * x * (1 + 0.044715 * x * x)))


def bias_gelu_back(g, x):
    t = torch.tanh(0.79788456 * x * (1 + 0.044715 * x * x))
    ff = 0.5 * x * ((1 - t * t) * (0.79788456 + 0.1070322243 * x * x)) + 0.5 * (1 + t)
    return ff * g


class FPDT_FFN(torch.autograd.Function):
    """``gelu(x W1^T + b1) W2^T (+ b2)`` evaluated ``chunk_size`` sequence rows at a time; the ``4h`` intermediate is
    never alive for more than one chunk and is recomputed in backward.  Returns ``(y, b2 or None)``."""

    @staticmethod
    def forward(ctx: Any, x, w1, b1, w2, b2, add_bias, chunk_size):
        num_chunk = x.shape[0] // chunk_size
        assert chunk_size * num_chunk == x.shape[0]
        ctx.add_bias, ctx.num_chunk = add_bias, num_chunk
        result = torch.empty_like(x)
        with torch.no_grad():
            for i in range(num_chunk):
                sl = slice(i * chunk_size, (i + 1) * chunk_size)
                inter = F.bias_act(F.linear(x[sl], w1), b1, "gelu") if x.is_cuda else bias_gelu(
                    F.matmul_nt(x[sl], w1) + b1)
                result[sl] = F.linear(inter, w2, b2 if add_bias else None)
                del inter
        ctx.save_for_backward(x, w1, b1, w2, b2)
        return result, (b2 if not add_bias else None)

    @staticmethod
    def backward(ctx, grad_output, grad_bias):
        x, w1, b1, w2, b2 = ctx.saved_tensors
        n = ctx.num_chunk
        cs = x.shape[0] // n
        gw1 = torch.zeros(w1.shape, device=x.device, dtype=torch.float32)
        gb1 = torch.zeros(b1.shape, device=x.device, dtype=torch.float32)
        gw2 = torch.zeros(w2.shape, device=x.device, dtype=torch.float32)
        gb2 = torch.zeros(b2.shape, device=x.device, dtype=torch.float32)
        gx = torch.empty_like(x)
        for i in range(n):
            sl = slice(i * cs, (i + 1) * cs)
            xc, go = x[sl], grad_output[sl]
            pre = F.matmul_nt(xc, w1) + b1
            act = bias_gelu(pre)
            F.accumulate_tn(gw2, go, act)
            ginter = bias_gelu_back(F.matmul_nn(go, w2), pre)
            del pre, act
            F.accumulate_tn(gw1, ginter, xc)
            gb1.add_(ginter.reshape(-1, ginter.shape[-1]).sum(0).float())
            gx[sl] = F.matmul_nn(ginter, w1)
            if ctx.add_bias:
                gb2.add_(go.reshape(-1, go.shape[-1]).sum(0).float())
        dt = x.dtype
        return gx, gw1.to(dt), gb1.to(dt), gw2.to(dt), gb2.to(dt), None, None


# ------------------------------------------------------------------------------------------ chunked logits + loss
class FPDT_LogitsLoss(torch.autograd.Function):
    """Per-token NLL of ``lm_output W^T`` computed ``num_chunk`` sequence slices at a time (the ``[s, b, vocab]`` logits
    never exist), then all-gathered over the sequence-parallel group: returns ``[s_local * sp, b]`` fp32 losses."""

    @staticmethod
    def forward(ctx: Any, lm_output, labels, logit_weights, rank, spg_size, spg, num_chunk):
        labels = labels.t()                                           # [s_local, b]
        s_local, b = lm_output.shape[0], lm_output.shape[1]
        cs = s_local // num_chunk
        assert cs * num_chunk == s_local
        loss = torch.empty((s_local, b), dtype=torch.float32, device=lm_output.device)
        with torch.no_grad():
            for i in range(num_chunk):
                sl = slice(i * cs, (i + 1) * cs)
                logits = F.linear(lm_output[sl], logit_weights).float()
                lse = torch.logsumexp(logits, dim=-1)
                picked = logits.gather(-1, labels[sl].unsqueeze(-1)).squeeze(-1)
                loss[sl] = lse - picked
                del logits
        ctx.num_chunk, ctx.cs, ctx.rank, ctx.s_local = num_chunk, cs, rank, s_local
        ctx.save_for_backward(lm_output, labels, logit_weights)
        if spg_size > 1:
            loss_all = torch.empty(s_local * spg_size, b, dtype=loss.dtype, device=loss.device)
            dist.all_gather_into_tensor(loss_all, loss.contiguous(), group=spg)
            return loss_all
        return loss

    @staticmethod
    def backward(ctx, grad_output):
        lm_output, labels, w = ctx.saved_tensors
        cs, n = ctx.cs, ctx.num_chunk
        grad_output = grad_output[ctx.rank * ctx.s_local:(ctx.rank + 1) * ctx.s_local]
        gx = torch.empty_like(lm_output)
        gw = torch.zeros(w.shape, device=w.device, dtype=torch.float32)
        for i in range(n):
            sl = slice(i * cs, (i + 1) * cs)
            xc = lm_output[sl]
            probs = torch.softmax(F.matmul_nt(xc, w).float(), dim=-1)
            probs.scatter_add_(-1, labels[sl].unsqueeze(-1), -torch.ones_like(probs[..., :1]))
            probs.mul_(grad_output[sl].unsqueeze(-1))
            g = probs.to(lm_output.dtype)
            gx[sl] = F.matmul_nn(g, w)
            F.accumulate_tn(gw, g, xc)
        return gx, None, gw.to(w.dtype), None, None, None, None
