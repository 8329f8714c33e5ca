"""module_inject helpers: fused-QKV slicing per layout, TP-aware ALiBi, checkpoint streaming, training-layer injection,
in-place INT8 layer quantisation."""
import pytest
import torch
import torch.nn as nn

from shuffle_exchange_b200.module_inject import auto_tp_model_utils as atp
from shuffle_exchange_b200.module_inject import fusedqkv_utils as fq
from shuffle_exchange_b200.module_inject import tp_shard


@pytest.fixture(autouse=True)
def _heads():
    tp_shard.set_num_kv_heads(4); tp_shard.set_num_attention_heads(4); tp_shard.set_n_embd(16)
    yield
    tp_shard.set_num_kv_heads(None); tp_shard.set_num_attention_heads(None); tp_shard.set_n_embd(None)


def _attn_with(wq, wk, wv, x, heads):
    B, S, H = x.shape
    d = wq.shape[0] // heads
    q, k, v = ((x @ w.t()).view(B, S, heads, d).transpose(1, 2) for w in (wq, wk, wv))
    return (torch.softmax(q @ k.transpose(-1, -2) / d ** 0.5, -1) @ v).transpose(1, 2).reshape(B, S, -1)


def test_fused_qkv_slices_give_each_rank_its_own_heads():
    torch.manual_seed(0)
    H, heads, d = 16, 4, 4
    wq, wk, wv = torch.randn(H, H), torch.randn(H, H), torch.randn(H, H)
    x = torch.randn(2, 5, H)
    full = _attn_with(wq, wk, wv, x, heads)
    # glmtype: [q | k | v]
    glm = torch.cat([wq, wk, wv])
    for r in range(2):
        part = fq.prepare_tp_fused_qkvw("GLMBlock", glm, 2, r)
        q, k, v = part.chunk(3)
        assert torch.allclose(_attn_with(q, k, v, x, heads // 2), full[..., r * 8:(r + 1) * 8], atol=1e-5)
    # bloomtype: [heads, 3, d]
    bloom = torch.stack([wq.view(heads, d, H), wk.view(heads, d, H), wv.view(heads, d, H)], dim=1).reshape(3 * H, H)
    for r in range(2):
        part = fq.prepare_tp_fused_qkvw("BloomBlock", bloom, 2, r).view(heads // 2, 3, d, H)
        q, k, v = (part[:, i].reshape(-1, H) for i in range(3))
        assert torch.allclose(_attn_with(q, k, v, x, heads // 2), full[..., r * 8:(r + 1) * 8], atol=1e-5)
    # bigcode MQA: queries cut, the single k/v head replicated
    mqa = torch.cat([wq, wk[:d], wv[:d]])
    part = fq.prepare_tp_fused_qkvw("GPTBigCodeBlock", mqa, 2, 1)
    assert part.shape[0] == 8 + 2 * d and torch.equal(part[:8], wq[8:]) and torch.equal(part[8:], mqa[H:])
    # phi3 GQA with 2 kv heads
    tp_shard.set_num_kv_heads(2)
    gqa = torch.cat([wq, wk[:2 * d], wv[:2 * d]])
    part = fq.prepare_tp_fused_qkvw("Phi3DecoderLayer", gqa, 2, 0)
    assert part.shape[0] == 8 + d + d and torch.equal(part[8:8 + d], wk[:d])
    assert fq.require_tp_fused_qkvw("h.0.self_attention.query_key_value", 2)
    assert not fq.require_tp_fused_qkvw("h.0.mlp.dense", 2) and not fq.require_tp_fused_qkvw("a.c_attn", 1)
    gu = torch.randn(12, 4)
    sh = fq.shard_chunk_mlp(gu, None, 1, 2)
    assert torch.equal(sh, torch.cat([gu[3:6], gu[9:12]]))
    fused = fq.split_by_qkvlist_and_refuse([wq, wk, wv], 8)
    assert len(fused) == 2 and torch.equal(fused[1], torch.cat([wq[8:], wk[8:], wv[8:]]))


def test_alibi_builders_return_the_ranks_heads():
    from transformers.models.bloom.modeling_bloom import build_alibi_tensor
    mask = torch.ones(2, 7, dtype=torch.long)
    mask[1, :2] = 0
    ref = build_alibi_tensor(mask, 8, torch.float32).view(2, 8, 1, 7)
    for r in range(2):
        got = atp.build_bloom_alibi_tensor(mask, 8, torch.float32, rank=r, world=2).view(2, 4, 1, 7)
        assert torch.allclose(got, ref[:, r * 4:(r + 1) * 4], atol=1e-6)
    full = atp.build_mpt_alibi_tensor(None, 6, 5, rank=0, world=1)
    assert full.shape == (6, 1, 5) and torch.all(full[..., -1] == 0) and torch.all(full[..., 0] < 0)
    assert torch.equal(atp.build_mpt_alibi_tensor(None, 6, 5, rank=1, world=2), full[3:])

    class B:
        n_head = 4
    m = atp.get_alibi_mask(B(), torch.zeros(1), 6, rank=1, world=2)
    assert m.shape == (2, 6, 6) and torch.isinf(m[0, 0, 1]) and torch.isfinite(m[0, 3, 3])


def test_load_model_with_checkpoint_materialises_meta_and_concats_tp_shards():
    from shuffle_exchange_b200.module_inject import load_model_with_checkpoint
    torch.manual_seed(0)
    ref = nn.Sequential(nn.Embedding(10, 8), nn.Linear(8, 12), nn.Linear(12, 8))
    sd = {"model." + k: v.clone() for k, v in ref.state_dict().items()}
    with torch.device("meta"):
        m = nn.Sequential(nn.Embedding(10, 8), nn.Linear(8, 12), nn.Linear(12, 8))
    missing = load_model_with_checkpoint(m, sd)
    assert missing == [] and all(p.device.type != "meta" for p in m.parameters())
    ids = torch.tensor([[1, 2, 3]])
    assert torch.allclose(m(ids), ref(ids))
    # two tensor-parallel shards of one checkpoint (column-parallel 1.weight, row-parallel 2.weight)
    full = ref.state_dict()
    shards = [{"0.weight": full["0.weight"], "1.weight": full["1.weight"][r * 6:(r + 1) * 6],
               "1.bias": full["1.bias"][r * 6:(r + 1) * 6], "2.weight": full["2.weight"][:, r * 6:(r + 1) * 6],
               "2.bias": full["2.bias"]} for r in range(2)]
    m2 = nn.Sequential(nn.Embedding(10, 8), nn.Linear(8, 12), nn.Linear(12, 8))
    assert load_model_with_checkpoint(m2, shards, ckpt_type="tp", ckpt_mp_size=2) == []
    assert torch.allclose(m2(ids), ref(ids))
    assert load_model_with_checkpoint(nn.Linear(3, 3), {}) == ["weight", "bias"]


def test_training_layer_injection_and_layer_quantisation():
    transformers = pytest.importorskip("transformers")
    from transformers import AutoConfig, AutoModel
    from transformers.models.bert.modeling_bert import BertLayer
    from shuffle_exchange_b200.module_inject import module_inject, quantize_transformer_layer
    from shuffle_exchange_b200.ops.transformer import DeepSpeedTransformerLayer
    torch.manual_seed(0)
    cfg = AutoConfig.for_model(model_type="bert", vocab_size=64, hidden_size=32, num_hidden_layers=2,
                               num_attention_heads=4, intermediate_size=64, max_position_embeddings=32,
                               hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0)
    m = AutoModel.from_config(cfg).float().eval()
    g = torch.Generator().manual_seed(11)
    with torch.no_grad():              # zero-initialised biases / unit LayerNorm scales would hide swapped tensors
        for n, p in m.named_parameters():
            if n.endswith("bias"):
                p.copy_(torch.randn(p.shape, generator=g) * 0.2)
            elif "LayerNorm.weight" in n:
                p.copy_(1 + torch.randn(p.shape, generator=g) * 0.1)
    layer = m.encoder.layer[0]
    x = torch.randn(2, 6, 32)
    with torch.no_grad():
        ref = layer(x)
        ref = ref[0] if isinstance(ref, tuple) else ref
    module_inject(BertLayer, m, cfg, micro_batch_size=2, max_seq_length=32, seed=1, preln=False, fp16=False)
    fused = m.encoder.layer[0]
    assert isinstance(fused, DeepSpeedTransformerLayer)
    with torch.no_grad():
        out = fused(x)
        out = out[0] if isinstance(out, tuple) else out
    assert (out - ref).abs().max() < 5e-3
    # in-place INT8 quantisation of a fresh model's layers
    m2 = AutoModel.from_config(cfg).float().eval()
    w = m2.encoder.layer[1].output.dense.weight.data.clone()
    quantize_transformer_layer(BertLayer, m2, groups=4)
    lin = m2.encoder.layer[1].output.dense
    assert lin.weight.dtype == torch.int8 and lin.weight_scale.numel() == 4
    assert (lin.dequantized_weight() - w).abs().max() <= lin.weight_scale.max() * 0.5 + 1e-6
    assert m2.encoder.layer[1](x) is not None
