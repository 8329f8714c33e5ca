"""FastGen v2 tensor parallelism: sharding arithmetic, and a 2-rank (gloo) engine whose logits equal HuggingFace's."""
import pytest
import torch

from conftest import run_dist
from shuffle_exchange_b200.inference.v2.model_implementations import sharding as sh


def test_shard_endpoints_cover_the_dimension_exactly_once():
    for dim, n, g in [(96, 2, 32), (4096, 8, 32), (11008, 3, 32), (129, 4, 1)]:
        ends = [sh.get_shard_endpoints(dim, r, n, g) for r in range(n)]
        assert ends[0][0] == 0 and ends[-1][1] == dim
        assert all(a[1] == b[0] for a, b in zip(ends, ends[1:]))
        sizes = [b - a for a, b in ends]
        assert max(sizes) - min(sizes) <= g and sizes == sorted(sizes, reverse=True)
    with pytest.raises(ValueError):
        sh.get_shard_endpoints(100, 0, 2, 32)


def test_head_distribution_mha_gqa_and_replicated_kv():
    assert [sh.get_local_heads(r, 3, 8) for r in range(3)] == [(3, 3), (3, 3), (2, 2)]
    assert [sh.get_local_heads(r, 2, 8, 4) for r in range(2)] == [(4, 2), (4, 2)]
    assert [sh.get_local_heads(r, 4, 8, 2) for r in range(4)] == [(2, 1)] * 4        # each KV head on two ranks
    assert [sh.head_ranges(r, 4, 8, 2) for r in range(4)] == [((0, 2), (0, 1)), ((2, 2), (0, 1)), ((4, 2), (1, 1)),
                                                              ((6, 2), (1, 1))]
    with pytest.raises(ValueError):
        sh.get_local_heads(0, 3, 8, 4)


def test_sharded_projections_recompose_the_dense_result():
    torch.manual_seed(0)
    H, hq, hkv, d, Fd, n = 64, 8, 2, 8, 128, 4
    x = torch.randn(5, H)
    wqkv, wo = torch.randn((hq + 2 * hkv) * d, H), torch.randn(H, hq * d)
    q_full = (x @ wqkv.t())[:, : hq * d]
    parts = []
    for r in range(n):
        w = sh.shard_qkv_param(wqkv, r, n, d, hq, hkv)
        ql, kl = sh.get_local_heads(r, n, hq, hkv)
        assert w.shape[0] == (ql + 2 * kl) * d == sh.qkv_out_features(H, r, n, d, hq, hkv)
        parts.append((x @ w.t())[:, : ql * d])
    assert torch.allclose(torch.cat(parts, 1), q_full, atol=1e-5)
    ctx = torch.randn(5, hq * d)
    acc = 0
    for r in range(n):
        w = sh.shard_attn_out_param(wo, r, n, d, hq, hkv)
        (q0, qn), _ = sh.head_ranges(r, n, hq, hkv)
        acc = acc + ctx[:, q0 * d:(q0 + qn) * d] @ w.t()
    assert torch.allclose(acc, ctx @ wo.t(), atol=1e-4)
    # gated MLP: gate|up halves are each cut, so silu(gate) * up stays aligned
    w1, w2 = torch.randn(2 * Fd, H), torch.randn(H, Fd)
    def mlp(a, b):
        g, u = (x @ a.t()).chunk(2, dim=-1)
        return (torch.nn.functional.silu(g) * u) @ b.t()
    acc = sum(mlp(sh.shard_mlp_1_param(w1, r, n, gated=True), sh.shard_mlp_2_param(w2, r, n)) for r in range(n))
    assert torch.allclose(acc, mlp(w1, w2), atol=1e-3)
    # biases: column-parallel bias is cut, row-parallel bias lives on rank 0 only
    assert sh.shard_mlp_2_param(torch.randn(H), 1, n) is None and sh.shard_mlp_2_param(torch.randn(H), 0, n) is not None
    e1 = torch.randn(3, 2 * Fd, H)
    assert sh.shard_mlp_1_param(e1, 1, n, gated=True, is_moe=True).shape == (3, 2 * Fd // n, H)
    assert sh.shard_unembed_param(torch.randn(131, H), 0, 4).shape[0] == 33
    assert sh.shard_embedding_param(torch.randn(131, H), 1, 2).shape == (131, H // 2)


def _tp_engine(rank, world, name):
    transformers = pytest.importorskip("transformers")
    from transformers import AutoConfig, AutoModelForCausalLM
    from shuffle_exchange_b200.inference.v2.config_v2 import DeepSpeedTPConfig, RaggedInferenceEngineConfig
    from shuffle_exchange_b200.inference.v2.engine_factory import build_engine_from_hf_model
    common = dict(vocab_size=131, hidden_size=64, num_hidden_layers=2, num_attention_heads=4, num_key_value_heads=2,
                  max_position_embeddings=128, intermediate_size=128)
    cfgs = {"llama": dict(model_type="llama", **common),
            "mixtral": dict(model_type="mixtral", num_local_experts=4, num_experts_per_tok=2, **common),
            "qwen2": dict(model_type="qwen2", **common),
            "opt": dict(model_type="opt", vocab_size=131, hidden_size=64, num_hidden_layers=2, num_attention_heads=4,
                        ffn_dim=128, max_position_embeddings=128, word_embed_proj_dim=64)}
    torch.manual_seed(0)
    m = AutoModelForCausalLM.from_config(AutoConfig.for_model(**cfgs[name])).float().eval()
    g = torch.Generator().manual_seed(11)
    with torch.no_grad():              # HF zero-initialises biases, which would hide bias-placement mistakes
        for n, p in m.named_parameters():
            if n.endswith("bias"):
                p.copy_(torch.randn(p.shape, generator=g) * 0.2)
            elif p.dim() == 1 and n.endswith("weight"):          # norm scales are initialised to exactly 1
                p.copy_(1 + torch.randn(p.shape, generator=g) * 0.1)
    ids = torch.randint(0, 131, (1, 9), generator=torch.Generator().manual_seed(1))
    with torch.no_grad():
        ref = m(ids).logits[0]
    eng = build_engine_from_hf_model(m, RaggedInferenceEngineConfig(tensor_parallel=DeepSpeedTPConfig(tp_size=world)))
    model = eng.model()
    assert model.tp_size == world and model.tp_rank == rank and model.n_heads_kv_local == (2 if name == "opt" else 1)
    out = eng.put([7], [ids[0, :6]])
    assert out.shape[-1] == 131 and (out[0] - ref[5]).abs().max() < 1e-4
    out = eng.put([7], [ids[0, 6:9]])
    assert (out[0] - ref[8]).abs().max() < 1e-4


@pytest.mark.parametrize("name", ["llama", "mixtral", "qwen2", "opt"])
def test_two_rank_tensor_parallel_engine_matches_huggingface(name):
    run_dist(_tp_engine, world=2, args=(name,))


def test_flatten_restore_round_trip_shares_tied_tensors():
    from shuffle_exchange_b200.inference.v2.inference_parameter import InferenceParameter
    from shuffle_exchange_b200.inference.v2.model_implementations.flat_model_helpers import (
        ModelMetadata, flatten_inference_model, pad_to_aligned_offset, restore_inference_model)
    from shuffle_exchange_b200.inference.v2.model_implementations.inference_transformer_base import LayerWeights
    torch.manual_seed(0)
    emb = torch.randn(10, 8)
    q = InferenceParameter.initialize(torch.randint(0, 255, (6, 3), dtype=torch.uint8), scales=torch.randn(6))
    w = {"embed": emb, "unembed": emb, "final_norm_w": torch.randn(8).to(torch.bfloat16), "_transformed": True,
         "layers": [LayerWeights(wqkv=torch.randn(4, 8), wo=q), LayerWeights(wqkv=torch.randn(4, 8))]}
    ref = {"embed": emb.clone(), "l0": w["layers"][0]["wqkv"].clone(), "q": q.clone(), "qs": q.scales.clone()}
    buf, md = flatten_inference_model(w)
    assert buf.dtype == torch.uint8 and buf.numel() == md.total_bytes == pad_to_aligned_offset(md.total_bytes)
    assert md.params["embed"].core_param.offset == md.params["unembed"].core_param.offset        # tied: stored once
    assert all(pm.core_param.offset % 256 == 0 for pm in md.params.values())
    assert w["embed"].untyped_storage().data_ptr() == buf.untyped_storage().data_ptr()           # rebound to views
    md2 = ModelMetadata.parse_raw(md.json())
    r = restore_inference_model(buf.clone(), md2)
    assert torch.equal(r["embed"], ref["embed"]) and torch.equal(r["layers"][0]["wqkv"], ref["l0"])
    assert torch.equal(r["layers"][0]["wo"].as_subclass(torch.Tensor), ref["q"].as_subclass(torch.Tensor))
    assert torch.equal(r["layers"][0]["wo"].scales, ref["qs"]) and r["final_norm_w"].dtype == torch.bfloat16
    assert len(r["layers"]) == 2


def _tp_serialize(rank, world, path):
    transformers = pytest.importorskip("transformers")
    from transformers import AutoConfig, AutoModelForCausalLM
    from shuffle_exchange_b200.inference.v2.config_v2 import DeepSpeedTPConfig, RaggedInferenceEngineConfig
    from shuffle_exchange_b200.inference.v2.engine_factory import build_engine_from_hf_model, build_hf_engine
    import torch.distributed as dist
    torch.manual_seed(0)
    m = AutoModelForCausalLM.from_config(AutoConfig.for_model(
        model_type="llama", vocab_size=131, hidden_size=64, num_hidden_layers=2, num_attention_heads=4,
        num_key_value_heads=2, max_position_embeddings=128, intermediate_size=128)).float().eval()
    ids = torch.randint(0, 131, (1, 8), generator=torch.Generator().manual_seed(4))
    with torch.no_grad():
        ref = m(ids).logits[0, -1]
    cfg = RaggedInferenceEngineConfig(tensor_parallel=DeepSpeedTPConfig(tp_size=world))
    eng = build_engine_from_hf_model(m, cfg)
    eng.serialize(path)
    dist.barrier()
    import os
    assert os.path.exists(os.path.join(path, f"params_rank_{rank}_of_{world}.pt"))
    eng2 = build_hf_engine(path, cfg)             # detects ds_model_config.pkl and restores the flat shards
    assert eng2.model().flattened_params is not None
    out = eng2.put([3], [ids[0]])
    assert (out[0] - ref).abs().max() < 1e-4


def test_serialize_and_reload_per_rank_flat_checkpoint(tmp_path):
    run_dist(_tp_serialize, world=2, args=(str(tmp_path),))


def test_parameter_dsl_assembles_in_any_order_and_reports_gaps():
    from types import SimpleNamespace
    from shuffle_exchange_b200.inference.v2.model_implementations.common_parameters import (
        GQAMegatronQKVParameter, UnfusedMoEGatedMLPParameter, UnfusedQKVParameter)
    from shuffle_exchange_b200.inference.v2.model_implementations.containers import (Llama2TransformerContainer,
                                                                                      MixtralTransformerContainer)
    from shuffle_exchange_b200.inference.v2.model_implementations.layer_container_base import LayerContainer
    model = SimpleNamespace(n_experts=2, n_heads=4, n_heads_kv=2, head_size=2)
    p = UnfusedQKVParameter(model)
    p.v_params = torch.full((2, 3), 3.0); p.q_params = torch.full((4, 3), 1.0)
    assert not p.is_complete
    p.k_params = torch.full((2, 3), 2.0)
    assert p.is_complete and p.value[:, 0].tolist() == [1.0] * 4 + [2.0] * 2 + [3.0] * 2
    with pytest.raises(ValueError):
        p.q_params = torch.zeros(1)
    moe = UnfusedMoEGatedMLPParameter(model)
    moe.up_experts[1] = torch.full((2, 3), 4.0); moe.gating_experts[0] = torch.full((2, 3), 1.0)
    moe.gating_experts[1] = torch.full((2, 3), 3.0); moe.up_experts[0] = torch.full((2, 3), 2.0)
    assert moe.value.shape == (2, 4, 3) and moe.value[:, :, 0].tolist() == [[1.0, 1.0, 2.0, 2.0], [3.0, 3.0, 4.0, 4.0]]
    # grouped rows [kv group][q0 q1 k v] -> q|k|v
    g = GQAMegatronQKVParameter(model)
    rows = torch.arange(2 * 4 * 2).float().view(-1, 1)          # 2 groups x (2 q + k + v) x head_size 2
    g.params = rows
    assert g.value.view(-1).tolist() == [0, 1, 2, 3, 8, 9, 10, 11, 4, 5, 12, 13, 6, 7, 14, 15]
    c = Llama2TransformerContainer(model)
    assert c.n_params == 7 and not c.is_populated and "qkv_b" not in c.unfinished() and c.qkv_w is None
    assert not c.set_dependency("self_attn.rotary_emb.inv_freq", torch.zeros(1))
    for n in ("q", "k", "v"):
        assert c.set_dependency(f"self_attn.{n}_proj.weight", torch.zeros(2, 3))
    assert c.qkv_w.shape == (6, 3) and c.unfinished() == ["attn_out_w", "mlp_1_w", "mlp_2_w", "attn_norm_gamma",
                                                          "mlp_norm_gamma"]
    mc = MixtralTransformerContainer(model)
    assert mc.set_dependency("block_sparse_moe.experts.1.w2.weight", torch.zeros(3, 2)) and mc.moe_mlp_2 is None
    mc.set_dependency("block_sparse_moe.experts.0.w2.weight", torch.zeros(3, 2))
    assert mc.moe_mlp_2.shape == (2, 3, 2) and mc.to_layer_weights()["moe_w2"] is mc.moe_mlp_2
    with pytest.raises(TypeError):
        class Bad(LayerContainer):
            qkv_w: UnfusedQKVParameter
            PARAM_MAPPING = {"x": "qkv_w.nope"}


@pytest.mark.parametrize("name", ["llama", "mixtral", "opt"])
def test_container_policy_builds_the_same_weights_as_the_rule_policy(name, tmp_path):
    transformers = pytest.importorskip("transformers")
    from transformers import AutoConfig, AutoModelForCausalLM
    from shuffle_exchange_b200.inference.v2.checkpoint import HuggingFaceCheckpointEngine
    from shuffle_exchange_b200.inference.v2.config_v2 import RaggedInferenceEngineConfig
    from shuffle_exchange_b200.inference.v2.engine_v2 import InferenceEngineV2
    from shuffle_exchange_b200.inference.v2.model_implementations.policies import CONTAINER_POLICIES, POLICIES
    common = dict(vocab_size=96, hidden_size=32, num_hidden_layers=2, num_attention_heads=4, max_position_embeddings=64)
    cfgs = {"llama": dict(model_type="llama", intermediate_size=64, num_key_value_heads=2, **common),
            "mixtral": dict(model_type="mixtral", intermediate_size=64, num_key_value_heads=2, num_local_experts=4,
                            num_experts_per_tok=2, **common),
            "opt": dict(model_type="opt", ffn_dim=64, word_embed_proj_dim=32, **common)}
    torch.manual_seed(0)
    m = AutoModelForCausalLM.from_config(AutoConfig.for_model(**cfgs[name])).float().eval()
    m.save_pretrained(str(tmp_path), safe_serialization=True)       # per-expert tensors on disk
    a = POLICIES[name](m.config, checkpoint_engine=HuggingFaceCheckpointEngine(str(tmp_path)))
    b = CONTAINER_POLICIES[name](m.config, checkpoint_engine=HuggingFaceCheckpointEngine(str(tmp_path)))
    a.dtype_override = b.dtype_override = torch.float32
    wa, wb = a.load_weights(), b.load_weights()
    assert set(k for k in wa if k != "layers") == set(k for k in wb if k != "layers")
    for k in wa:
        if k != "layers":
            assert torch.equal(wa[k], wb[k]), k
    for la, lb in zip(wa["layers"], wb["layers"]):
        assert set(la) == set(lb), (sorted(la), sorted(lb))
        for k in la:
            assert torch.equal(la[k], lb[k]), k
    b._engine_config = RaggedInferenceEngineConfig()
    eng = InferenceEngineV2(b, b._engine_config)
    ids = torch.randint(0, 96, (1, 7), generator=torch.Generator().manual_seed(3))
    with torch.no_grad():
        ref = m(ids).logits[0, -1]
    assert (eng.put([1], [ids[0]])[0] - ref).abs().max() < 1e-4
