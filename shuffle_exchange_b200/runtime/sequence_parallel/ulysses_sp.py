"""ALST / Ulysses sequence parallelism for Hugging-Face-style models.

Reference: ``deepspeed/runtime/sequence_parallel/ulysses_sp.py`` — ``UlyssesSPAttentionHF`` (:47-423, head<->sequence
all-to-all around any attention function, KV-head replication when ``kv_heads < sp``), ``UlyssesSPDataLoaderAdapter``
(:426-551, every SP group processes one rank's sample at a time, sharded on the sequence dim), generic tiled compute
(``sequence_tiled_compute`` :554, ``SequenceTiledCompute`` :608, ``TiledMLP`` :757), SFT-aware tiled loss
(``TiledLoss`` :915) and the forward/loss/backward driver ``UlyssesSPFwdLossBwdWithLogits`` (:1064).

B200 notes: the default core attention is the framework's TMA/tcgen05 flash-attention kernel on ``[B, S, H, D]`` (no
layout round trip through ``[S, B, H, D]``); Q, K and V travel in ONE all-to-all (concatenated on the head dim) instead
of three, and the context tensor in a second one — NVSwitch gives every pair full bandwidth, so the cost is launch
latency, not link count.  Tiled compute recomputes each tile in backward, so peak activation memory is one tile.
"""
from __future__ import annotations

import math
from collections import defaultdict
from typing import Any, Callable, Dict, List, Optional, Tuple

import torch
import torch.distributed as dist
from torch import Tensor

from ...ops import functional as F
from ...sequence.layer import _SeqAllToAll


# =========================================================================================== attention wrapper
class UlyssesSPAttentionHF(torch.nn.Module):
    """Wraps a core attention callable with the Ulysses exchange.

    ``forward(module, query, key, value, attention_mask, *args, **kwargs)`` follows the HF attention-interface: inputs
    are ``[batch, heads, seq_local, head_dim]``; the wrapped ``attn`` is called with ``[batch, heads_local, seq_global,
    head_dim]`` tensors and must return ``(context [batch, seq_global, heads_local, head_dim], weights_or_None)``.
    The result is ``(context [batch, seq_local, heads, head_dim], weights)``."""

    def __init__(self, attn, local_seq_length: int, global_seq_length: int, batch_size: int, attn_head_count: int,
                 attn_head_size: int, kv_head_count: int, num_hidden_layers: int, process_group,
                 seq_length_is_variable: bool = False) -> None:
        super().__init__()
        self.attn = attn
        self.process_group = process_group
        self.world_size = dist.get_world_size(process_group)
        self.sp_rank = dist.get_rank(process_group)
        self.local_seq_length, self.global_seq_length = local_seq_length, global_seq_length
        self.batch_size = batch_size
        self.seq_length_is_variable = seq_length_is_variable
        self.attn_head_size, self.attn_head_count = attn_head_size, attn_head_count
        self.global_kv_head_count = kv_head_count
        self.num_hidden_layers = num_hidden_layers
        self.skip_all_but_last_attention_debug_mode = False
        self.rotating_layer_counter = 0
        if attn_head_count % self.world_size != 0:
            raise ValueError(f"Attention head count {attn_head_count} is not divisible by SP size {self.world_size}")
        if not (kv_head_count % self.world_size == 0 or self.world_size % kv_head_count == 0):
            raise ValueError(f"KV attention head count {kv_head_count} is not divisible by SP size "
                             f"{self.world_size} or vice versa")
        self.local_q_head_count = attn_head_count // self.world_size
        # 4 kv heads on sp=8: every kv head is replicated 2x so each rank still owns one
        self.kv_replication_factor = max(1, self.world_size // kv_head_count)
        self.local_kv_head_count = 1 if self.kv_replication_factor > 1 else kv_head_count // self.world_size
        self.required_query_shape = torch.Size([local_seq_length, batch_size, attn_head_count, attn_head_size])
        self.required_key_value_shape = torch.Size([local_seq_length, batch_size, kv_head_count, attn_head_size])
        self.required_context_shape = torch.Size(
            [global_seq_length, batch_size, attn_head_size * attn_head_count // self.world_size])

    # -- exchanges ----------------------------------------------------------------------------------------
    def _combine_local_sequences(self, query, key, value) -> Tuple[Tensor, Tensor, Tensor]:
        """[b, s_local, h, d] x3 -> [b, s_global, h/sp, d] x3 with a single all-to-all."""
        P = self.world_size
        if self.kv_replication_factor > 1:
            key = key.repeat_interleave(self.kv_replication_factor, dim=2)
            value = value.repeat_interleave(self.kv_replication_factor, dim=2)
        hq, hk = query.shape[2] // P, key.shape[2] // P
        b, s, _, d = query.shape
        # interleave so that rank r's slice of the head dim holds its q, k and v heads contiguously
        packed = torch.cat([query.reshape(b, s, P, hq, d), key.reshape(b, s, P, hk, d),
                            value.reshape(b, s, P, hk, d)], dim=3).reshape(b, s, P * (hq + 2 * hk), d)
        out = _SeqAllToAll.apply(self.process_group, packed, 2, 1, 0)            # [b, s*P, hq+2hk, d]
        return out[:, :, :hq], out[:, :, hq:hq + hk], out[:, :, hq + hk:]

    def _partition_global_sequence(self, context) -> Tensor:
        """[b, s_global, h/sp, d] -> [b, s_local, h, d]"""
        return _SeqAllToAll.apply(self.process_group, context, 1, 2, 0)

    def forward(self, module: torch.nn.Module, query: Tensor, key: Tensor, value: Tensor, attention_mask: Tensor,
                *args: Any, **kwargs: Any):
        # HF gives [b, h, s_local, d]
        if not self.seq_length_is_variable:
            b, h, s, d = query.shape
            assert (s, b, h, d) == tuple(self.required_query_shape), \
                f"[{self.sp_rank}]: query input tensor does not match the required shape " \
                f"({self.required_query_shape}): {(s, b, h, d)}"
        q, k, v = (t.transpose(1, 2) for t in (query, key, value))               # [b, s_local, h, d]
        q, k, v = self._combine_local_sequences(q.contiguous(), k.contiguous(), v.contiguous())
        self.rotating_layer_counter = (self.rotating_layer_counter + 1) % max(self.num_hidden_layers, 1)
        if self.skip_all_but_last_attention_debug_mode and self.rotating_layer_counter != 0:
            context, weights = q.contiguous(), None                              # debug: bypass the core attention
        else:
            # the core sees HF layout [b, h_local, s_global, d]; masks are rebuilt by the core from position ids /
            # causality (a local-sequence mask would be wrong for the gathered sequence)
            context, weights = self.attn(module, q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2), None,
                                         *args, **kwargs)
        context = self._partition_global_sequence(context.contiguous())          # [b, s_local, h, d]
        return context, weights

    # -- registration -------------------------------------------------------------------------------------
    @classmethod
    def register_with_transformers(cls, model_name_or_path, core_attn_implementation, sequence_parallel_size,
                                   max_length, micro_batch_size, seq_length_is_variable=True):
        """Install the Ulysses wrapper as the attention implementation ``"ulysses"`` of ``transformers`` and return the
        parallel-state module (``mpu``) whose groups it uses.  ``model_name_or_path`` may be a path/hub id or an
        already-built ``PretrainedConfig``.  ``core_attn_implementation``: a name registered in
        ``transformers.AttentionInterface`` (``"sdpa"``, ``"flash_attention_2"``…) or ``"sxb_flash"`` for the
        framework's tcgen05 kernel."""
        from . import parallel_state_sp as mpu
        mpu.initialize_sequence_parallel(sequence_parallel_size=sequence_parallel_size)
        import transformers
        from transformers import AutoConfig
        from transformers.modeling_utils import ALL_ATTENTION_FUNCTIONS
        cfg = model_name_or_path if hasattr(model_name_or_path, "num_attention_heads") else \
            AutoConfig.from_pretrained(model_name_or_path)
        if core_attn_implementation == "sxb_flash":
            core = sxb_flash_core_attention
        else:
            if core_attn_implementation not in ALL_ATTENTION_FUNCTIONS:
                raise ValueError(f"{core_attn_implementation} is not a valid attn_implementation; choices: "
                                 f"{list(ALL_ATTENTION_FUNCTIONS.keys()) + ['sxb_flash']}")
            core = ALL_ATTENTION_FUNCTIONS[core_attn_implementation]
        heads = cfg.num_attention_heads
        uattn = cls(attn=core, local_seq_length=max_length // sequence_parallel_size, global_seq_length=max_length,
                    batch_size=micro_batch_size, attn_head_count=heads,
                    attn_head_size=getattr(cfg, "head_dim", None) or cfg.hidden_size // heads,
                    kv_head_count=getattr(cfg, "num_key_value_heads", heads),
                    num_hidden_layers=cfg.num_hidden_layers, process_group=mpu.get_sequence_parallel_group(),
                    seq_length_is_variable=seq_length_is_variable)

        def uattn_wrapper(module, query, key, value, attention_mask, *args, **kwargs):
            return uattn(module, query, key, value, attention_mask, *args, **kwargs)

        transformers.AttentionInterface.register("ulysses", uattn_wrapper)
        return mpu


def sxb_flash_core_attention(module, query, key, value, attention_mask, *args, scaling: Optional[float] = None,
                             is_causal: Optional[bool] = None, **kwargs):
    """HF attention-interface adapter of the framework's flash-attention kernel: ``[b, h, s, d]`` in,
    ``([b, s, h, d], None)`` out."""
    causal = True if is_causal is None else bool(is_causal)
    q, k, v = (t.transpose(1, 2) for t in (query, key, value))
    return F.flash_attention(q, k, v, causal=causal, scale=scaling), None


# =========================================================================================== dataloader adapter
class UlyssesSPDataLoaderAdapter:
    """Turns a per-rank DataLoader into an SP-over-DP loader: the SP group consumes rank 0's sample (each rank holding
    ``1/sp`` of its sequence), then rank 1's, … — ``sp`` micro-steps per wrapped-loader step.  ``labels`` are
    pre-shifted into ``shift_labels`` before sharding so the loss of a shard never needs a neighbour's token."""

    def __init__(self, dl, sp_rank: int, sp_group, sp_world_size, device):
        self.dl, self.sp_rank, self.sp_group = dl, sp_rank, sp_group
        self.sp_world_size, self.device = sp_world_size, device
        self.iter = iter(dl)
        self.micro_batches: List[Any] = []

    def __len__(self):
        return len(self.dl) * self.sp_world_size

    def __iter__(self):
        return self

    def __next__(self):
        if not self.micro_batches:
            self.refill()
        return self.micro_batches.pop(0)

    def refill(self):
        batch = next(self.iter)            # StopIteration propagates
        P = self.sp_world_size
        per_rank: Dict[int, Dict[str, Any]] = defaultdict(dict)
        seqlens = _gather_seqlens(batch["input_ids"].shape[1], P, self.sp_group, self.device)
        for k, val in batch.items():
            if torch.is_tensor(val):
                val = val.to(self.device)
                parts = [torch.zeros((val.shape[0], seqlens[i]) + tuple(val.shape[2:]), dtype=val.dtype,
                                     device=val.device) for i in range(P)]
                dist.all_gather(parts, val.contiguous(), group=self.sp_group)
            else:
                parts = [None] * P
                dist.all_gather_object(parts, val, group=self.sp_group)
            for r, t in enumerate(parts):
                per_rank[r][k] = t
        for r in range(P):
            self.micro_batches.append(_shard_batch(per_rank[r], P, self.sp_rank, keep_on_cpu=True))


def _gather_seqlens(seqlen: int, P: int, group, device) -> List[int]:
    mine = torch.tensor([seqlen], dtype=torch.int64, device=device)
    out = [torch.zeros(1, dtype=torch.int64, device=device) for _ in range(P)]
    dist.all_gather(out, mine, group=group)
    return [int(x.item()) for x in out]


def _shard_batch(batch: Dict[str, Any], P: int, sp_rank: int, keep_on_cpu: bool) -> Dict[str, Any]:
    seq_length = batch["input_ids"].shape[1]
    if seq_length % P != 0:
        raise ValueError(f"batch's seqlen={seq_length} isn't divisible by sp-size={P}")
    chunk = seq_length // P
    batch = dict(batch)
    labels = torch.nn.functional.pad(batch.pop("labels"), (0, 1), value=-100)
    batch["shift_labels"] = labels[..., 1:].contiguous()
    for k, val in batch.items():
        if torch.is_tensor(val):
            piece = val[:, chunk * sp_rank:chunk * (sp_rank + 1)]
            batch[k] = piece.cpu() if keep_on_cpu else piece
    return batch


# =========================================================================================== tiled compute
def _gather_zero3(params):
    """ZeRO-3 parameters must be whole while a tile is recomputed outside the module hooks."""
    import contextlib
    params = [p for p in (params or []) if hasattr(p, "ds_id") or hasattr(p, "sxb_zero3")]
    if not params:
        return contextlib.nullcontext()
    from ..zero.partition_parameters import GatheredParameters
    return GatheredParameters(params, modifier_rank=None)


def sequence_tiled_compute(fn, seqlen, shards, kwargs_to_shard, kwargs_to_pass, grad_requiring_tensor_key,
                           compute_params=None, output_unshard_dimension=1, output_reduction="mean"):
    """Run ``fn(**sharded, **passed)`` over ``shards`` slices of the sequence dim (dim 1) with one slice of activations
    alive at a time; see :class:`SequenceTiledCompute`.  ``output_reduction``: ``None`` | ``"mean"`` | ``"sum"``."""
    keys_to_shard, keys_to_pass = list(kwargs_to_shard.keys()), list(kwargs_to_pass.keys())
    return SequenceTiledCompute.apply(fn, seqlen, shards, keys_to_shard, keys_to_pass, grad_requiring_tensor_key,
                                      compute_params, output_unshard_dimension, output_reduction,
                                      *kwargs_to_shard.values(), *kwargs_to_pass.values())


class SequenceTiledCompute(torch.autograd.Function):
    """Generic tiled forward + tile-by-tile recompute backward for any function of ``[bs, seqlen, ...]`` kwargs."""

    @staticmethod
    def forward(ctx, fn, seqlen, shards, keys_to_shard, keys_to_pass, grad_requiring_tensor_key, compute_params,
                output_unshard_dimension, output_reduction, *args) -> Tensor:
        if output_reduction not in (None, "mean", "sum"):
            raise ValueError(f"unknown value {output_reduction}: valid values are: none/mean/sum")
        args = list(args)
        n_shard = len(keys_to_shard)
        ctx.fn, ctx.seqlen, ctx.shards = fn, seqlen, shards
        ctx.key = grad_requiring_tensor_key
        ctx.key_index = (keys_to_shard + keys_to_pass).index(grad_requiring_tensor_key)
        ctx.total_args = len(args)
        ctx.compute_params = [p for p in (compute_params or []) if p.requires_grad]
        ctx.unshard_dim, ctx.reduction = output_unshard_dimension, output_reduction
        ctx.to_shard = dict(zip(keys_to_shard, args[:n_shard]))
        ctx.to_pass = dict(zip(keys_to_pass, args[n_shard:]))
        step = math.ceil(seqlen / shards)
        outs = []
        with torch.no_grad():
            for i in range(shards):
                outs.append(fn(**{k: v[:, i * step:(i + 1) * step] for k, v in ctx.to_shard.items()}, **ctx.to_pass))
        if output_unshard_dimension == 0:
            full = torch.stack(outs, dim=0) if outs[0].dim() == 0 else torch.cat([o.unsqueeze(0) for o in outs], 0)
        else:
            full = torch.cat(outs, dim=output_unshard_dimension)
        ctx.out_numel = full.numel()
        if output_reduction is None:
            return full
        return full.mean() if output_reduction == "mean" else full.sum()

    @staticmethod
    def backward(ctx, *grads):
        fn, shards, step = ctx.fn, ctx.shards, math.ceil(ctx.seqlen / ctx.shards)
        x = ctx.to_shard[ctx.key]
        x_grad = torch.zeros_like(x) if x.requires_grad else None
        incoming = grads[0]
        with _gather_zero3(ctx.compute_params):
            for i in range(shards):
                sl = slice(i * step, (i + 1) * step)
                kw = {k: v[:, sl] for k, v in ctx.to_shard.items()}
                xi = kw[ctx.key].detach().requires_grad_(x.requires_grad)
                kw[ctx.key] = xi
                with torch.enable_grad():
                    out = fn(**kw, **ctx.to_pass)
                if ctx.reduction is None:
                    if ctx.unshard_dim == 0:
                        g = incoming[i]
                    else:
                        g = incoming.narrow(ctx.unshard_dim, i * step, out.shape[ctx.unshard_dim])
                elif ctx.reduction == "mean":
                    g = incoming.expand_as(out) / ctx.out_numel if out.dim() else incoming / ctx.out_numel
                else:
                    g = incoming.expand_as(out) if out.dim() else incoming
                torch.autograd.backward(out, g)      # accumulates into parameters' .grad directly
                if x_grad is not None:
                    x_grad[:, sl] = xi.grad
        ret: List[Optional[Tensor]] = [None] * (9 + ctx.total_args)
        ret[9 + ctx.key_index] = x_grad
        return tuple(ret)


class TiledMLP(torch.autograd.Function):
    """``fn(self, x)`` over ``shards`` slices of the sequence dim of ``x [bs, seqlen, hidden]``; each slice is
    recomputed in backward, so only one slice's ``[.., intermediate]`` activations exist at any time."""

    @staticmethod
    def forward(ctx, fn, self, x, shards, compute_params) -> Tensor:
        ctx.fn, ctx.self, ctx.shards = fn, self, shards
        ctx.compute_params = [p for p in (compute_params or []) if p.requires_grad]
        ctx.save_for_backward(x)
        with torch.no_grad():
            outs = [fn(self, xs) for xs in torch.chunk(x, chunks=shards, dim=-2)]
        return torch.cat(outs, dim=-2)

    @staticmethod
    def backward(ctx, *grads):
        (x,) = ctx.saved_tensors
        fn, self, shards = ctx.fn, ctx.self, ctx.shards
        x_grad = torch.zeros_like(x)
        xs = list(torch.chunk(x, chunks=shards, dim=-2))
        gs = list(torch.chunk(grads[0], chunks=shards, dim=-2))
        gxs = list(torch.chunk(x_grad, chunks=shards, dim=-2))
        with _gather_zero3(ctx.compute_params):
            for xi, gi, gxi in zip(xs, gs, gxs):
                xi = xi.detach().requires_grad_(x.requires_grad)
                with torch.enable_grad():
                    out = fn(self, xi)
                torch.autograd.backward(out, gi)
                if xi.grad is not None:
                    gxi.copy_(xi.grad)
        return None, None, x_grad, None, None


class TiledFusedLogitsLoss(torch.autograd.Function):
    """SFT-aware tiled loss: ``loss_fn(logits=…, labels=None, vocab_size=…, shift_labels=…)`` evaluated per sequence
    tile, tiles whose labels are all ``-100`` skipped, tile means re-weighted by their number of real labels.  (The
    reference names this ``TiledLoss``; both names are exported.)"""

    @staticmethod
    def forward(ctx, loss_fn, logits, vocab_size, shift_labels, shards) -> Tensor:
        ctx.save_for_backward(logits, shift_labels)
        ctx.loss_fn, ctx.vocab_size, ctx.shards = loss_fn, vocab_size, shards
        step = math.ceil(shift_labels.shape[1] / shards)
        total, good = None, 0
        with torch.no_grad():
            for i in range(shards):
                lab = shift_labels[:, i * step:(i + 1) * step]
                n = int((lab != -100).sum())
                if n == 0:
                    continue
                l = loss_fn(logits=logits[:, i * step:(i + 1) * step, :], labels=None, vocab_size=vocab_size,
                            shift_labels=lab) * n
                total = l if total is None else total + l
                good += n
        ctx.good = good
        if total is None:
            return (logits.sum() * 0.0).float()
        return total / good

    @staticmethod
    def backward(ctx, *grads):
        logits, shift_labels = ctx.saved_tensors
        step = math.ceil(shift_labels.shape[1] / ctx.shards)
        g_logits = torch.zeros_like(logits)
        if ctx.good == 0:
            return None, g_logits, None, None, None
        for i in range(ctx.shards):
            sl = slice(i * step, (i + 1) * step)
            lab = shift_labels[:, sl]
            n = int((lab != -100).sum())
            if n == 0:
                continue
            xi = logits[:, sl, :].detach().requires_grad_(True)
            with torch.enable_grad():
                l = ctx.loss_fn(logits=xi, labels=None, vocab_size=ctx.vocab_size, shift_labels=lab)
            (gi,) = torch.autograd.grad(l, xi, grads[0] * (n / ctx.good))
            g_logits[:, sl, :] = gi
        return None, g_logits, None, None, None


TiledLoss = TiledFusedLogitsLoss


# =========================================================================================== fwd / loss / bwd driver
class UlyssesSPFwdLossBwdWithLogits:
    """Drives one optimizer step's worth of SP micro-steps for a model that returns full logits: gathers every rank's
    sample, and for each of them runs forward on this rank's sequence shard, a (tiled) loss, a label-count-weighted
    loss average over the SP group, and backward — gradient accumulation boundary only after the last sample."""

    def __init__(self, model, model_unwrapped, device, num_loss_logit_shards="auto", **kwargs):
        self.model, self.model_unwrapped, self.device = model, model_unwrapped, device
        self.num_loss_logit_shards = num_loss_logit_shards
        self.kwargs = kwargs
        from ...utils import groups
        self.sp_group = groups._get_sequence_parallel_group()
        self.sp_world_size = groups._get_sequence_parallel_world_size()
        self.sp_rank = groups._get_sequence_parallel_rank()

    def sp_fwd_loss_bwd(self, batch) -> float:
        if not (batch["input_ids"].shape == batch["position_ids"].shape == batch["labels"].shape):
            raise ValueError(f'Borked batch {batch["input_ids"].shape=} != {batch["position_ids"].shape=} != '
                             f'{batch["labels"].shape=}) in DataLoader->iter->next, cannot continue with Ulysses '
                             "Sequence parallelism")
        P = self.sp_world_size
        seqlens = _gather_seqlens(batch["input_ids"].shape[1], P, self.sp_group, self.device)
        per_rank: Dict[int, Dict[str, Tensor]] = defaultdict(dict)
        with torch.no_grad():
            for k in list(batch.keys()):
                val = batch[k].to(self.device)
                parts = [torch.zeros((val.shape[0], seqlens[i]), dtype=val.dtype, device=val.device)
                         for i in range(P)]
                dist.all_gather(parts, val.contiguous(), group=self.sp_group)
                for r, t in enumerate(parts):
                    per_rank[r][k] = t
        if hasattr(self.model, "set_gradient_accumulation_boundary"):
            self.model.set_gradient_accumulation_boundary(False)
        losses = []
        for sub in range(P):
            full = per_rank[sub]
            seq_length = full["input_ids"].shape[1]
            if seq_length % P != 0:
                raise ValueError(f"sub_step_id={sub}: batch's seqlen={seq_length} isn't divisible by sp-size={P}")
            chunk = seq_length // P
            counts = []
            for r in range(P):
                n = int((full["labels"][:, chunk * r:chunk * (r + 1)] != -100).sum())
                counts.append(n - 1 if n > 1 else n)
            shard = _shard_batch(full, P, self.sp_rank, keep_on_cpu=False)
            shift_labels = shard.pop("shift_labels").to(self.device)
            self.forward({k: v.to(self.device) for k, v in shard.items()})
            loss = self.compute_loss(labels=None, shift_labels=shift_labels)
            from torch.distributed.nn.functional import all_gather as ag
            per = ag(loss, group=self.sp_group)
            denom = max(sum(counts), 1)
            self.loss = sum(per[r] * counts[r] for r in range(P)) / denom
            self.backward()
            losses.append(float(self.loss.detach()))
        if hasattr(self.model, "set_gradient_accumulation_boundary"):
            self.model.set_gradient_accumulation_boundary(True)
        return sum(losses) / len(losses) if losses else float("nan")

    def forward(self, batch):
        outputs = self.model(**batch, use_cache=False)
        self.logits = outputs.logits
        return outputs

    def compute_loss(self, labels, shift_labels):
        mu = self.model_unwrapped
        if bool((shift_labels == -100).all()):
            loss = (self.logits.sum() * 0.0).float()
        else:
            if self.num_loss_logit_shards == "auto":
                self.num_loss_logit_shards = math.ceil(self.logits.numel() * 4 / 2 ** 30)     # 1 GiB fp32 slices
            if self.num_loss_logit_shards > 1:
                loss = TiledFusedLogitsLoss.apply(mu.loss_function, self.logits, mu.config.vocab_size, shift_labels,
                                                  self.num_loss_logit_shards)
            else:
                loss = mu.loss_function(logits=self.logits, labels=None, vocab_size=mu.config.vocab_size,
                                        shift_labels=shift_labels)
        self.loss = loss
        return loss

    def backward(self):
        if hasattr(self.model, "backward"):
            self.model.backward(self.loss)
        else:
            self.loss.backward()
