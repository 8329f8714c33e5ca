"""``init_inference(..., replace_with_kernel_inject=True)``: every supported architecture's layers are replaced by the
fused inference layer and the model still produces the HuggingFace outputs (CPU, fp32, random tiny models)."""
import pytest
import torch

transformers = pytest.importorskip("transformers")
from transformers import AutoConfig, AutoModel, AutoModelForCausalLM  # noqa: E402

import shuffle_exchange_b200 as sxb  # noqa: E402
from shuffle_exchange_b200.module_inject.containers import HFLayerAdapter  # noqa: E402

_C = dict(vocab_size=96, hidden_size=32, num_hidden_layers=2, num_attention_heads=4, max_position_embeddings=64)
CAUSAL = {
    "gpt2": dict(model_type="gpt2", vocab_size=96, n_embd=32, n_layer=2, n_head=4, n_positions=64),
    "opt": dict(model_type="opt", vocab_size=96, hidden_size=32, num_hidden_layers=2, num_attention_heads=4, ffn_dim=64,
                max_position_embeddings=64, word_embed_proj_dim=32),
    "llama": dict(model_type="llama", intermediate_size=48, num_key_value_heads=2, **_C),
    "gptj": dict(model_type="gptj", vocab_size=96, n_embd=32, n_layer=2, n_head=4, n_positions=64, rotary_dim=4),
    "gpt_neox": dict(model_type="gpt_neox", intermediate_size=64, rotary_pct=0.5, use_parallel_residual=True, **_C),
    "gpt_neox_seq": dict(model_type="gpt_neox", intermediate_size=64, rotary_pct=1.0, use_parallel_residual=False, **_C),
    "gpt_neo": dict(model_type="gpt_neo", vocab_size=96, hidden_size=32, num_layers=2, num_heads=4,
                    attention_types=[[["global"], 2]], max_position_embeddings=64),
    "bloom": dict(model_type="bloom", vocab_size=96, hidden_size=32, n_layer=2, n_head=4),
}
ENCODERS = {
    "bert": dict(model_type="bert", intermediate_size=64, **_C),
    "distilbert": dict(model_type="distilbert", vocab_size=96, dim=32, n_layers=2, n_heads=4, hidden_dim=64,
                       max_position_embeddings=64),
}


def _randomize_biases(model):
    """HF initialises every bias (and LayerNorm shift) to zero, which would hide bias-placement mistakes."""
    g = torch.Generator().manual_seed(11)
    with torch.no_grad():
        for n, p in model.named_parameters():
            if n.endswith("bias"):
                p.copy_(torch.randn(p.shape, generator=g) * 0.2)
            elif p.dim() == 1 and n.endswith("weight"):          # norm scales are initialised to exactly 1
                p.copy_(1 + torch.randn(p.shape, generator=g) * 0.1)


def _inject(model):
    eng = sxb.init_inference(model, dtype=torch.float32, replace_with_kernel_inject=True)
    n = sum(isinstance(m, HFLayerAdapter) for m in eng.module.modules())
    assert n == 2, f"expected both layers to be replaced, got {n}"
    return eng


@pytest.mark.parametrize("name", sorted(CAUSAL))
def test_causal_lm_logits_survive_kernel_injection(name):
    torch.manual_seed(0)
    m = AutoModelForCausalLM.from_config(AutoConfig.for_model(**CAUSAL[name])).float().eval()
    _randomize_biases(m)
    ids = torch.randint(0, 96, (2, 11), generator=torch.Generator().manual_seed(1))
    with torch.no_grad():
        ref = m(ids, use_cache=False).logits.clone()
        eng = _inject(m)
        out = eng.module(ids, use_cache=False).logits
    tol = 2e-3                        # the fused layer uses tanh-GELU where some clients use erf-GELU
    assert (out - ref).abs().max() < tol, (name, float((out - ref).abs().max()))


@pytest.mark.parametrize("name", sorted(ENCODERS))
def test_encoder_hidden_states_survive_kernel_injection(name):
    torch.manual_seed(0)
    m = AutoModel.from_config(AutoConfig.for_model(**ENCODERS[name])).float().eval()
    _randomize_biases(m)
    ids = torch.randint(0, 96, (2, 9), generator=torch.Generator().manual_seed(2))
    mask = torch.ones(2, 9, dtype=torch.long)
    mask[1, 6:] = 0                   # padding on the second row exercises the additive mask path
    with torch.no_grad():
        ref = m(ids, attention_mask=mask).last_hidden_state.clone()
        eng = _inject(m)
        out = eng.module(ids, attention_mask=mask).last_hidden_state
    keep = mask.bool()
    assert (out - ref)[keep].abs().max() < 5e-3, (name, float((out - ref)[keep].abs().max()))


def test_revert_restores_client_layers_and_unknown_model_is_rejected():
    from shuffle_exchange_b200.module_inject.replace_module import revert_transformer_layer
    torch.manual_seed(0)
    m = AutoModelForCausalLM.from_config(AutoConfig.for_model(**CAUSAL["gpt2"])).float().eval()
    eng = _inject(m)
    revert_transformer_layer(None, eng.module, None)
    assert not any(isinstance(x, HFLayerAdapter) for x in eng.module.modules())
    assert sum(type(x).__name__ == "GPT2Block" for x in eng.module.modules()) == 2
    with pytest.raises(ValueError):
        sxb.init_inference(torch.nn.Sequential(torch.nn.Linear(4, 4)), dtype=torch.float32,
                           replace_with_kernel_inject=True)


def test_container_features_meta_load_and_hybrid_lora_refresh():
    from shuffle_exchange_b200.inference.config import DeepSpeedInferenceConfig
    from shuffle_exchange_b200.module_inject.containers import DS_LLAMAContainer, LLAMALayerPolicy
    from shuffle_exchange_b200.module_inject.containers.features import HybridGatedMLPContainer, MetaTensorContainer
    torch.manual_seed(0)
    cfg = AutoConfig.for_model(**CAUSAL["llama"])
    m = AutoModelForCausalLM.from_config(cfg).float().eval()
    layer = m.model.layers[0]
    sd = {"model.layers.0." + k: v.clone() for k, v in layer.state_dict().items()}
    with torch.device("meta"):
        meta_layer = type(layer)(cfg, 0)

    class MetaLlama(MetaTensorContainer, DS_LLAMAContainer):
        pass

    icfg = DeepSpeedInferenceConfig(dtype=torch.float32)
    c = MetaLlama(LLAMALayerPolicy(meta_layer), icfg, cfg, 0, meta_layer)
    assert c.is_meta
    with pytest.raises(RuntimeError):
        c.initialize_tensors()
    c.load_params(sd, prefix="model.layers.0.")
    assert not c.is_meta
    c.create_ds_model_config(); c.initialize_tensors(); fused = c.create_module()
    assert torch.allclose(fused.attention.attn_ow.t(), layer.self_attn.o_proj.weight)

    class HybridLlama(HybridGatedMLPContainer, DS_LLAMAContainer):
        pass

    # LoRA on the client's down projection: fold -> refresh -> fused layer carries the merged weight; unfold restores
    down = layer.mlp.down_proj
    down.lora_right_weight = torch.nn.Parameter(torch.randn(down.in_features, 2) * 0.1)
    down.lora_left_weight = torch.nn.Parameter(torch.randn(2, down.out_features) * 0.1)
    down.lora_scaling = 0.5
    h = HybridLlama(LLAMALayerPolicy(layer), icfg, cfg, 0, layer)
    h.create_ds_model_config(); h.initialize_tensors(); fused = h.create_module()
    assert len(h.lora_params) == 1 and len(h.get_lora_params()) == 2
    before = down.weight.data.clone()
    h.fuse_lora(); h.refresh()
    merged = before + 0.5 * (down.lora_left_weight.t() @ down.lora_right_weight.t())
    assert torch.allclose(fused.mlp.output_w.t(), merged, atol=1e-6)
    h.unfuse_lora(); h.transform_for_inference()
    assert torch.allclose(fused.mlp.output_w.t(), before, atol=1e-6)
    up, gate = h.set_mlp_gate()
    assert torch.equal(fused.mlp.inter_w.t(), torch.cat([up, gate], 0))


def _inject_tp(rank, world, name):
    torch.manual_seed(0)
    m = AutoModelForCausalLM.from_config(AutoConfig.for_model(**CAUSAL[name])).float().eval()
    _randomize_biases(m)
    ids = torch.randint(0, 96, (2, 10), generator=torch.Generator().manual_seed(5))
    with torch.no_grad():
        ref = m(ids, use_cache=False).logits.clone()
        eng = sxb.init_inference(m, dtype=torch.float32, replace_with_kernel_inject=True,
                                 tensor_parallel={"tp_size": world})
        layer = next(x for x in eng.module.modules() if isinstance(x, HFLayerAdapter)).ds_layer
        assert layer.config.mp_size == world
        assert layer.attention.attn_ow.shape[0] == layer.config.hidden_size // world      # row-parallel slice
        out = eng.module(ids, use_cache=False).logits
    assert (out - ref).abs().max() < 2e-3, (name, rank, float((out - ref).abs().max()))


@pytest.mark.parametrize("name", ["gpt2", "llama", "gptj"])
def test_kernel_injection_with_two_way_tensor_parallelism(name):
    from conftest import run_dist
    run_dist(_inject_tp, world=2, args=(name,))


@pytest.mark.parametrize("name", ["gpt2", "llama", "gpt_neox"])
def test_int8_injection_keeps_one_byte_weights_and_close_logits(name):
    """``dtype=torch.int8``: the four GEMM weights of every injected layer are resident as int8 with one step size per
    row, expanded per call; logits stay within the int8 quantisation error of the fp32 model."""
    from shuffle_exchange_b200.module_inject.replace_module import GroupQuantizer
    from shuffle_exchange_b200.ops.transformer.inference.op_binding import dequantize_int8
    torch.manual_seed(0)
    m = AutoModelForCausalLM.from_config(AutoConfig.for_model(**CAUSAL[name])).float().eval()
    _randomize_biases(m)
    ids = torch.randint(0, 96, (2, 11), generator=torch.Generator().manual_seed(1))
    with torch.no_grad():
        ref = m(ids, use_cache=False).logits.clone()
        eng = sxb.init_inference(m, dtype=torch.int8, replace_with_kernel_inject=True)
        out = eng.module(ids, use_cache=False).logits
    layers = [mod for mod in eng.module.modules() if hasattr(mod, "attn_qkvw")]
    assert len(layers) == 2
    for att in layers:
        for w in (att.attn_qkvw, att.attn_ow):
            assert w.dtype == torch.int8 and w.scale.numel() == w.shape[0] and not w.requires_grad
    mlps = [mod for mod in eng.module.modules() if hasattr(mod, "output_w") and hasattr(mod, "inter_w")]
    assert all(mod.inter_w.dtype == torch.int8 and mod.output_w.dtype == torch.int8 for mod in mlps)
    assert (out.float() - ref).abs().max() < 0.03 * ref.abs().max()
    # the quantiser itself: round trip within half a step per element; non-GEMM tensors pass through
    q = GroupQuantizer(q_int8=True, num_groups=4)
    w = torch.randn(16, 24)
    qw = q.quantize(w)
    assert qw.dtype == torch.int8 and qw.scale.shape == (4,)
    back = dequantize_int8(qw, torch.float32)
    step = qw.scale.repeat_interleave(4)[:, None]
    err = (back - w).abs()
    assert (err <= step + 1e-6).all() and (err <= step * 0.5 + 1e-6).float().mean() > 0.98   # only +absmax clamps (127)
    passthrough = q.quantize(torch.randn(8))
    assert passthrough.dtype == torch.float32 and passthrough.scale.numel() == 1
    assert GroupQuantizer(q_int8=False).quantize(w).dtype == torch.float32


def _mp_checkpoint_worker(rank, world, name, root, dtype):
    """save_mp_checkpoint_path writes this rank's slices; a second engine built from a DIFFERENT random model and the
    descriptor reproduces the first engine's logits without seeing the original weights."""
    import json
    import os
    torch.manual_seed(0)
    cfg = AutoConfig.for_model(**CAUSAL[name])
    m = AutoModelForCausalLM.from_config(cfg).float().eval()
    _randomize_biases(m)
    ids = torch.randint(0, 96, (2, 9), generator=torch.Generator().manual_seed(1))
    tp = {"tp_size": world} if world > 1 else {}
    with torch.no_grad():
        eng = sxb.init_inference(m, dtype=dtype, replace_with_kernel_inject=True, tensor_parallel=tp,
                                 save_mp_checkpoint_path=root)
        want = eng.module(ids, use_cache=False).logits.clone()
    desc = json.load(open(os.path.join(root, "ds_inference_config.json")))
    assert desc["tp_size"] == world and desc["parallelization"] == "tp" and desc["checkpoints"]["non_tp"] == ["non-tp.pt"]
    assert len(desc["checkpoints"]["tp"]) == 8 * world and all(os.path.exists(os.path.join(root, f)) for f in desc["checkpoints"]["tp"])
    eng.destroy()
    torch.manual_seed(123)                                   # other weights: everything must come from the files
    m2 = AutoModelForCausalLM.from_config(cfg).float().eval()
    with torch.no_grad():
        eng2 = sxb.init_inference(m2, dtype=dtype, replace_with_kernel_inject=True, tensor_parallel=tp,
                                  checkpoint=os.path.join(root, "ds_inference_config.json"))
        got = eng2.module(ids, use_cache=False).logits
    assert torch.equal(got, want), float((got - want).abs().max())
    if dtype == torch.int8:
        att = [mod for mod in eng2.module.modules() if hasattr(mod, "attn_qkvw")][0]
        assert att.attn_qkvw.dtype == torch.int8 and att.attn_qkvw.scale.numel() == att.attn_qkvw.shape[0]
    if world > 1:
        try:
            sxb.init_inference(AutoModelForCausalLM.from_config(cfg).float().eval(), dtype=dtype,
                               replace_with_kernel_inject=True,
                               checkpoint=os.path.join(root, "ds_inference_config.json"))
            raised = False
        except ValueError:
            raised = True
        assert raised, "a tp_size=2 checkpoint must not load into a tp_size=1 engine"


def test_sharded_inference_checkpoint_round_trip_tp2(tmp_path):
    from conftest import run_dist
    run_dist(_mp_checkpoint_worker, world=2, args=("llama", str(tmp_path), torch.float32))


def test_sharded_inference_checkpoint_round_trip_int8(tmp_path):
    from conftest import run_dist
    run_dist(_mp_checkpoint_worker, world=1, args=("gpt2", str(tmp_path), torch.int8))
