"""Sparse attention layouts + block-sparse ops, Evoformer attention and the training transformer layer (CPU paths)."""
import math
import random

import pytest
import torch

from shuffle_exchange_b200.ops.sparse_attention import (BertSparseSelfAttention, BigBirdSparsityConfig,
                                                        BSLongformerSparsityConfig, DenseSparsityConfig,
                                                        FixedSparsityConfig, LocalSlidingWindowSparsityConfig,
                                                        MatMul, Softmax, SparseAttentionUtils, SparseSelfAttention,
                                                        VariableSparsityConfig)
from shuffle_exchange_b200.ops.deepspeed4science import DS4Sci_EvoformerAttention
from shuffle_exchange_b200.ops.transformer import DeepSpeedTransformerConfig, DeepSpeedTransformerLayer


def _dense_mask(layout, block):
    return layout.repeat_interleave(block, 1).repeat_interleave(block, 2).bool()


def test_fixed_layout_structure():
    cfg = FixedSparsityConfig(num_heads=2, block=4, num_local_blocks=4, num_global_blocks=1,
                              attention="unidirectional")
    lay = cfg.make_layout(64)
    assert lay.shape == (2, 16, 16)
    assert torch.equal(lay[0], lay[1])                       # propagated first head
    assert torch.equal(lay[0], torch.tril(lay[0]))           # causal
    assert lay[0].diagonal().all()
    # block 3 of every window (its last) is the global column for all later rows
    assert lay[0, 3:, 3].all() and lay[0, 7:, 7].all()
    assert lay[0, 8, 4] == 0                                 # non-global block of an earlier window


def test_fixed_layout_rejects_bad_args():
    with pytest.raises(ValueError):
        FixedSparsityConfig(num_heads=2, num_local_blocks=4, num_global_blocks=3)
    with pytest.raises(ValueError):
        FixedSparsityConfig(num_heads=2, attention="unidirectional", horizontal_global_attention=True)
    with pytest.raises(NotImplementedError):
        FixedSparsityConfig(num_heads=2, attention="sideways")
    with pytest.raises(ValueError):
        DenseSparsityConfig(num_heads=2, block=16).make_layout(40)


def test_other_layouts():
    random.seed(0)
    assert DenseSparsityConfig(num_heads=3, block=4).make_layout(32).sum() == 3 * 8 * 8
    lw = LocalSlidingWindowSparsityConfig(num_heads=1, block=4, num_sliding_window_blocks=3).make_layout(32)[0]
    assert torch.equal(lw, torch.tril(lw)) and lw.sum() == 8 + 7       # diagonal + one sub-diagonal
    bb = BigBirdSparsityConfig(num_heads=2, block=4, num_random_blocks=1, num_sliding_window_blocks=3,
                               num_global_blocks=1).make_layout(64)
    assert bb[0, 0].all() and bb[0, :, 0].all()
    for r in range(1, 15):
        assert bb[0, r, r - 1:r + 2].all()
    lf = BSLongformerSparsityConfig(num_heads=2, block=4, global_block_indices=[0, 5]).make_layout(64)
    assert lf[0, 5].all() and lf[0, :, 5].all()
    var = VariableSparsityConfig(num_heads=2, block=4, local_window_blocks=[1, 3], global_block_indices=[2],
                                 attention="unidirectional").make_layout(64)
    assert var[0, 2:, 2].all() and var[0, 1, 0] == 0 and var[0, 3, 1] == 1


def _ref_attention(q, k, v, mask, rpe=None, kpm=None, am=None):
    s = (q @ k.transpose(-1, -2)) * q.shape[-1] ** -0.5
    s = s.masked_fill(~mask[None], float("-inf"))
    if rpe is not None:
        s = s + rpe
    if kpm is not None:
        s = s + kpm[:, None, None, :]
    if am is not None:
        s = s.masked_fill(am[None, None] == 0, float("-inf"))
    return torch.softmax(s, -1) @ v


def test_sparse_self_attention_matches_dense_masked():
    torch.manual_seed(0)
    cfg = FixedSparsityConfig(num_heads=2, block=4, num_local_blocks=2, num_global_blocks=1)
    att = SparseSelfAttention(cfg, max_seq_length=32, attn_mask_mode="mul")
    q, k, v = (torch.randn(2, 2, 16, 8, requires_grad=True) for _ in range(3))
    rpe = torch.randn(2, 16, 16)
    kpm = torch.zeros(2, 16)
    kpm[:, 13:] = -1e4
    am = torch.ones(16, 16).tril()
    out = att(q, k, v, rpe=rpe, key_padding_mask=kpm, attn_mask=am)
    mask = _dense_mask(att.get_layout(16), 4)
    ref = _ref_attention(q, k, v, mask, rpe, kpm, am)
    assert torch.allclose(out, ref, atol=1e-5)
    g = torch.randn_like(out)
    for a, b in zip(torch.autograd.grad(out, (q, k, v), g, retain_graph=True), torch.autograd.grad(ref, (q, k, v), g)):
        assert torch.allclose(a, b, atol=1e-5)


def test_matmul_modes_against_dense():
    torch.manual_seed(1)
    H, nb, blk, D = 2, 4, 4, 8
    layout = (torch.rand(H, nb, nb) > 0.4).long()
    layout[:, torch.arange(nb), torch.arange(nb)] = 1
    mask = _dense_mask(layout, blk)
    a = torch.randn(3, H, nb * blk, D)
    b = torch.randn(3, H, nb * blk, D)
    sdd = MatMul(layout, blk, "sdd", trans_b=True)(a, b)
    dense = (a @ b.transpose(-1, -2)) * mask[None]
    nz = layout.nonzero()
    for i, (h, r, c) in enumerate(nz.tolist()):
        assert torch.allclose(sdd[:, i], dense[:, h, r * blk:(r + 1) * blk, c * blk:(c + 1) * blk], atol=1e-5)
    dsd = MatMul(layout, blk, "dsd")(sdd, b)
    assert torch.allclose(dsd, dense @ b, atol=1e-4)
    dsd_t = MatMul(layout, blk, "dsd", trans_a=True)(sdd, b)
    assert torch.allclose(dsd_t, dense.transpose(-1, -2) @ b, atol=1e-4)
    x = torch.randn(3, H, 5, nb * blk)
    dds = MatMul(layout, blk, "dds")(x, sdd)
    assert torch.allclose(dds, x @ dense, atol=1e-4)
    dds_t = MatMul(layout, blk, "dds", trans_b=True)(x, sdd)
    assert torch.allclose(dds_t, x @ dense.transpose(-1, -2), atol=1e-4)


def test_softmax_rejects_dtype_mismatch():
    layout = torch.ones(1, 2, 2, dtype=torch.long)
    sm = Softmax(layout, 4)
    x = torch.randn(1, 4, 4, 4)
    with pytest.raises(ValueError):
        sm(x, rpe=torch.zeros(1, 8, 8, dtype=torch.float64))


def test_bert_sparse_self_attention_and_utils():
    class C:
        hidden_size, num_attention_heads = 32, 4
    torch.manual_seed(0)
    layer = BertSparseSelfAttention(C(), FixedSparsityConfig(num_heads=4, block=4, num_local_blocks=2))
    layer.sparse_self_attention = SparseSelfAttention(FixedSparsityConfig(num_heads=4, block=4, num_local_blocks=2),
                                                      max_seq_length=64)
    x = torch.randn(2, 16, 32)
    out = layer(x, torch.zeros(2, 16))
    assert out.shape == (2, 16, 32)
    ids = torch.ones(2, 10, dtype=torch.long)
    pad_len, ids2, am2, tt2, pos2, emb2 = SparseAttentionUtils.pad_to_block_size(
        4, ids, torch.ones(2, 10, dtype=torch.bool), torch.zeros(2, 10, dtype=torch.long), None, None, 0, None)
    assert pad_len == 2 and ids2.shape == (2, 12) and not am2[:, 10:].any() and tt2.shape == (2, 12)
    assert SparseAttentionUtils.unpad_sequence_output(pad_len, torch.zeros(2, 12, 8)).shape == (2, 10, 8)


def test_evoformer_attention_with_bias_grads():
    torch.manual_seed(0)
    B, N, L, H, D = 2, 3, 20, 2, 8
    Q, K, V = (torch.randn(B, N, L, H, D, requires_grad=True) for _ in range(3))
    b1 = torch.randn(B, N, 1, 1, L, requires_grad=True)
    b2 = torch.randn(B, 1, H, L, L, requires_grad=True)
    out = DS4Sci_EvoformerAttention(Q, K, V, [b1, b2])
    sc = torch.einsum("bnlhd,bnmhd->bnhlm", Q, K) * D ** -0.5 + b1 + b2
    ref = torch.einsum("bnhlm,bnmhd->bnlhd", torch.softmax(sc, -1), V)
    assert torch.allclose(out, ref, atol=1e-5)
    g = torch.randn_like(out)
    got = torch.autograd.grad(out, (Q, K, V, b1, b2), g, retain_graph=True)
    exp = torch.autograd.grad(ref, (Q, K, V, b1, b2), g)
    for a, b in zip(got, exp):
        assert torch.allclose(a, b, atol=1e-4)
    with pytest.raises(AssertionError):
        DS4Sci_EvoformerAttention(Q, K, V, [torch.zeros(B, N, 1, L)])


def _ref_layer(layer, x, mask):
    c = layer.config
    H, Dh = c.heads, c.hidden_size // c.heads
    ln = lambda t, w, b: torch.nn.functional.layer_norm(t, (c.hidden_size,), w, b, c.layer_norm_eps)
    h = ln(x, layer.attn_nw, layer.attn_nb) if c.pre_layer_norm else x
    qkv = h @ layer.attn_qkvw.t() + layer.attn_qkvb
    B, S, _ = x.shape
    q, k, v = qkv.view(B, S, 3, H, Dh).permute(2, 0, 3, 1, 4)
    s = q @ k.transpose(-1, -2) / math.sqrt(Dh)
    if mask is not None:
        s = s + mask
    ctx = (torch.softmax(s, -1) @ v).permute(0, 2, 1, 3).reshape(B, S, H * Dh)
    a = ctx @ layer.attn_ow.t() + layer.attn_ob + x
    if not c.pre_layer_norm:
        a = ln(a, layer.attn_nw, layer.attn_nb)
    h2 = ln(a, layer.norm_w, layer.norm_b) if c.pre_layer_norm else a
    inter = torch.nn.functional.gelu(h2 @ layer.inter_w.t() + layer.inter_b, approximate="tanh")
    o = inter @ layer.output_w.t() + layer.output_b + a
    if not c.pre_layer_norm:
        o = ln(o, layer.norm_w, layer.norm_b)
    return o


@pytest.mark.parametrize("pre_ln,ckpt", [(True, False), (False, False), (True, True)])
def test_transformer_layer_matches_reference_math(pre_ln, ckpt):
    torch.manual_seed(0)
    cfg = DeepSpeedTransformerConfig(batch_size=2, hidden_size=32, heads=4, attn_dropout_ratio=0.0,
                                     hidden_dropout_ratio=0.0, num_hidden_layers=2, initializer_range=0.05,
                                     pre_layer_norm=pre_ln, gelu_checkpoint=ckpt, normalize_invertible=ckpt,
                                     attn_dropout_checkpoint=ckpt, layer_norm_eps=1e-5)
    layer = DeepSpeedTransformerLayer(cfg)
    with torch.no_grad():
        for p in layer.parameters():
            if p.dim() == 1:
                p.add_(torch.randn_like(p) * 0.1)
    x = torch.randn(2, 10, 32, requires_grad=True)
    mask = torch.zeros(2, 1, 1, 10)
    mask[..., 7:] = -1e4
    y = layer(x, mask)
    ref = _ref_layer(layer, x, mask)
    assert torch.allclose(y, ref, atol=2e-4), (y - ref).abs().max()
    g = torch.randn_like(y)
    params = [x] + list(layer.parameters())
    got = torch.autograd.grad(y, params, g, retain_graph=True)
    exp = torch.autograd.grad(ref, params, g)
    for a, b in zip(got, exp):
        assert torch.allclose(a, b, atol=5e-4, rtol=2e-3)


def test_transformer_layer_config_roundtrip_and_ids():
    cfg = DeepSpeedTransformerConfig.from_dict({"hidden_size": 16, "heads": 2, "intermediate_size": 64,
                                                "num_hidden_layers": 1, "initializer_range": 0.02,
                                                "attn_dropout_ratio": 0.1, "hidden_dropout_ratio": 0.1})
    first = DeepSpeedTransformerLayer(cfg)
    second = DeepSpeedTransformerLayer(DeepSpeedTransformerConfig(hidden_size=16, heads=2, num_hidden_layers=1,
                                                                  initializer_range=0.02, attn_dropout_ratio=0.0,
                                                                  hidden_dropout_ratio=0.0, return_tuple=True))
    assert second.config.layer_id == first.config.layer_id + 1
    assert second.config.intermediate_size == 64
    out = second(torch.randn(1, 4, 16))
    assert isinstance(out, tuple) and out[0].shape == (1, 4, 16)
    first.train()
    y = first(torch.randn(1, 4, 16))           # dropout active on the CPU path
    assert y.shape == (1, 4, 16)
