"""Ragged-batch inference model for the Llama family (Llama-2/3, Mistral: dense; Mixtral / Qwen2-MoE: sparse MLP).

Reference: ``inference/v2/model_implementations/{llama_v2,mistral,mixtral,qwen_v2,qwen_v2_moe}/model.py`` — per layer:
RMSNorm -> fused QKV GEMM -> rotary + KV-cache append -> blocked flash attention -> output GEMM -> RMSNorm ->
gated MLP (or top-k MoE) -> residual; final norm, last-token gather, unembedding.

Here: GEMMs are the tcgen05 kernel (``ops.functional.linear``), norms/activations the fused elementwise kernels,
rotary + cache append ``sxb_kv_rotary_blocked``, attention ``sxb_paged_attention``, MoE routing
``sxb_top_k_gating / moe_scatter / moe_gather`` with one GEMM per expert over its contiguous row range.
Weights come from a training :class:`~shuffle_exchange_b200.models.llama.LlamaForCausalLM` (same fused
``wqkv`` / ``w_gate_up`` layout) or from a HuggingFace state dict via that class's ``load_hf_state_dict``.
"""
from __future__ import annotations

from typing import List, Optional

import torch

from ....ops import functional as F
from ....ops import moe_ops
from ..kernels import ragged_ops as R


class RaggedLlamaModel:
    def __init__(self, causal_lm, state_manager, moe_layers: Optional[List[dict]] = None, top_k: int = 2):
        self.lm = causal_lm
        self.cfg = causal_lm.cfg
        self.state_manager = state_manager
        self.moe_layers = moe_layers       # per layer: {"gate": [E, H], "w_gate_up": [E, 2F, H], "w_down": [E, H, F]} or None
        self.top_k = top_k

    # memory the engine must plan for ------------------------------------------------------------
    def kv_cache_config(self, block_size: int):
        from ..ragged.kv_cache import KVCacheConfig
        c = self.cfg
        return KVCacheConfig(block_size=block_size, cache_shape=(c.num_hidden_layers, c.num_key_value_heads, c.head_dim),
                             cache_dtype=self.lm.model.embed_tokens.dtype)

    def get_kv_requirements(self, sequence, max_new_tokens: int, max_new_blocks: int):
        bs = self.state_manager.kv_block_size
        total = sequence.seen_tokens + max_new_tokens
        need = (total + bs - 1) // bs - sequence.cur_allocated_blocks
        if need > max_new_blocks:
            fit = (sequence.cur_allocated_blocks + max_new_blocks) * bs - sequence.seen_tokens
            return max(0, fit), max_new_blocks
        return max_new_tokens, max(0, need)

    def maybe_allocate_kv(self, sequence, n_new_tokens: int) -> None:
        _, need = self.get_kv_requirements(sequence, n_new_tokens, self.state_manager.free_blocks)
        if need > 0:
            sequence.extend_kv_cache(self.state_manager.allocate_blocks(need))

    # forward -----------------------------------------------------------------------------------------
    def _mlp(self, layer_idx: int, layer, x: torch.Tensor) -> torch.Tensor:
        moe = self.moe_layers[layer_idx] if self.moe_layers is not None else None
        if moe is None:
            return F.linear(F.swiglu(F.linear(x, layer.w_gate_up)), layer.w_down)
        logits = F.linear(x, moe["gate"])
        scores, assign, offsets, counts = moe_ops.top_k_gating(logits, self.top_k, renormalize=True)
        from ....moe import fused_ep
        if fused_ep.grouped_inference_ok(x, moe["w_gate_up"], moe["w_down"]):
            return fused_ep.grouped_experts_inference(x, scores, assign, offsets, counts, moe["w_gate_up"],
                                                      moe["w_down"], F.swiglu, self.top_k)
        moe_in, cumsum, slots = moe_ops.moe_scatter(x, assign, offsets, counts)
        ends = cumsum.tolist()
        out = torch.empty_like(moe_in)
        lo = 0
        for e, hi in enumerate(ends):
            if hi > lo:
                h = F.swiglu(F.linear(moe_in[lo:hi], moe["w_gate_up"][e]))
                out[lo:hi] = F.linear(h, moe["w_down"][e])
            lo = hi
        return moe_ops.moe_gather(out, scores, slots)

    @torch.no_grad()
    def forward(self, batch) -> torch.Tensor:
        """``batch``: finalized :class:`RaggedBatchWrapper`.  Returns logits [n_seqs, vocab] of every sequence's last
        in-flight token."""
        c, m = self.cfg, self.lm.model
        ids, t2s, pos = batch.input_ids(), batch.token_to_seq(), batch.token_positions()
        tables = batch.block_tables()
        h = R.ragged_embed(ids, m.embed_tokens)
        residual = None
        nq, nkv, d = c.num_attention_heads, c.num_key_value_heads, c.head_dim
        for li, layer in enumerate(m.layers):
            x, residual = F.rms_norm(h, layer.input_layernorm, c.rms_norm_eps, residual)
            qkv = F.linear(x, layer.wqkv)
            cache = self.state_manager.get_cache(li)
            R.kv_rotary_blocked(qkv, cache, t2s, pos, tables, nq, nkv, d, d, c.rope_theta)
            q = qkv.view(qkv.shape[0], nq + 2 * nkv, d)[:, :nq]
            o = R.ragged_attention(q, cache, batch)
            h = F.linear(o.reshape(o.shape[0], nq * d), layer.wo)
            x, residual = F.rms_norm(h, layer.post_attention_layernorm, c.rms_norm_eps, residual)
            h = self._mlp(li, layer, x)
        x, _ = F.rms_norm(h, m.norm, c.rms_norm_eps, residual)
        last = R.logits_gather(x, batch.last_token_index())
        return F.linear(last, self.lm.head_weight())
