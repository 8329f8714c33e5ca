"""GPU numerics of the fused softmax / dropout kernels and of the ops built on them (training transformer layer,
Evoformer attention, block-sparse attention): CUDA path vs the module's own PyTorch branch on CPU in fp32."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def _cuda():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    return torch.device("cuda")


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_biased_softmax_fwd_bwd(dtype):
    from shuffle_exchange_b200.ops.fused_softmax import biased_softmax
    dev = _cuda()
    torch.manual_seed(0)
    B, H, Sq, Sk = 3, 4, 37, 301
    s = torch.randn(B, H, Sq, Sk)
    b1 = torch.randn(B, 1, 1, Sk)
    b1[..., 250:] = float("-inf")
    b2 = torch.randn(1, H, Sq, Sk)
    g = torch.randn(B, H, Sq, Sk)
    sc, b1c, b2c = (t.clone().requires_grad_(True) for t in (s, b1, b2))
    ref = biased_softmax(sc, b1c, b2c, 0.3)
    ref.backward(g)
    sg, b1g, b2g = (t.to(dev, dtype).requires_grad_(True) for t in (s, b1, b2))
    out = biased_softmax(sg, b1g, b2g, 0.3)
    out.backward(g.to(dev, dtype))
    tol = 2e-5 if dtype == torch.float32 else 2e-2
    assert (out.float().cpu() - ref).abs().max() < tol
    assert (sg.grad.float().cpu() - sc.grad).abs().max() < tol
    assert (b2g.grad.float().cpu() - b2c.grad).abs().max() < tol * 4
    fin = torch.isfinite(b1c.grad)
    assert (b1g.grad.float().cpu()[fin] - b1c.grad[fin]).abs().max() < tol * 20


def test_bias_dropout_residual_statistics_and_grad():
    from shuffle_exchange_b200.ops.fused_softmax import bias_dropout_residual
    dev = _cuda()
    torch.manual_seed(0)
    x = torch.randn(512, 1024, device=dev, dtype=torch.bfloat16, requires_grad=True)
    bias = torch.randn(1024, device=dev, dtype=torch.bfloat16, requires_grad=True)
    res = torch.randn(512, 1024, device=dev, dtype=torch.bfloat16, requires_grad=True)
    # p = 0 / eval: exact
    y0 = bias_dropout_residual(x, bias, res, 0.25, training=False)
    assert torch.allclose(y0.float(), (x.float() + bias.float() + res.float()), atol=3e-2)
    y = bias_dropout_residual(x, bias, res, 0.25, training=True)
    g = torch.randn_like(y)
    g = torch.where(g == 0, torch.ones_like(g), g)
    y.backward(g)
    keep = x.grad != 0                                      # the keep mask, read back through the gradient
    assert abs(keep.float().mean().item() - 0.75) < 0.01
    # dx is g/(1-p) where kept and exactly 0 where dropped; dres = g; dbias = column sums of dx
    assert torch.allclose(x.grad[keep].float(), (g[keep].float() / 0.75), rtol=2e-2, atol=2e-2)
    assert torch.equal(res.grad, g)
    assert torch.allclose(bias.grad.float(), x.grad.float().sum(0), rtol=3e-2, atol=0.5)
    # forward agrees with the same mask: res + (x + bias)/(1-p) where kept, res where dropped
    exp = res.float() + torch.where(keep, (x.float() + bias.float()) / 0.75, torch.zeros_like(res.float()))
    assert (y.float() - exp).abs().max() < 0.1
    assert torch.equal(y[~keep], res[~keep])
    # reproducible under the same seed
    torch.manual_seed(5)
    a = bias_dropout_residual(x.detach(), None, None, 0.5, True)
    torch.manual_seed(5)
    b = bias_dropout_residual(x.detach(), None, None, 0.5, True)
    assert torch.equal(a, b)


@pytest.mark.parametrize("modes", [("add", "add"), ("mul", "mul")])
def test_bsr_softmax_matches_cpu(modes):
    from shuffle_exchange_b200.ops.fused_softmax import BsrLayout, bsr_softmax
    dev = _cuda()
    torch.manual_seed(0)
    H, nb, blk, B = 3, 6, 16, 2
    layout = (torch.rand(H, nb, nb) > 0.5).long()
    layout[:, torch.arange(nb), torch.arange(nb)] = 1
    lay = BsrLayout(layout, blk)
    S = nb * blk
    x = torch.randn(B, lay.nnz, blk, blk)
    rpe = torch.randn(H, S, S)
    if modes[0] == "add":
        kpm = torch.zeros(B, S)
        kpm[:, -9:] = -1e4
        am = torch.zeros(S, S).masked_fill(torch.ones(S, S).triu(1).bool(), -1e4)
    else:
        kpm = torch.ones(B, S)
        kpm[:, -9:] = 0
        am = torch.ones(S, S).tril()
    xc = x.clone().requires_grad_(True)
    ref = bsr_softmax(xc, lay, 0.25, rpe, kpm, am, modes[0], modes[1])
    g = torch.randn_like(ref)
    ref.backward(g)
    xg = x.to(dev).requires_grad_(True)
    out = bsr_softmax(xg, lay, 0.25, rpe.to(dev), kpm.to(dev), am.to(dev), modes[0], modes[1])
    out.backward(g.to(dev))
    assert (out.cpu() - ref).abs().max() < 2e-5
    assert (xg.grad.cpu() - xc.grad).abs().max() < 2e-5


def test_sparse_self_attention_gpu_bf16():
    from shuffle_exchange_b200.ops.sparse_attention import BigBirdSparsityConfig, SparseSelfAttention
    dev = _cuda()
    torch.manual_seed(0)
    cfg = BigBirdSparsityConfig(num_heads=4, block=16, num_random_blocks=1, num_sliding_window_blocks=3,
                                num_global_blocks=1)
    att = SparseSelfAttention(cfg, max_seq_length=256)
    q, k, v = (torch.randn(2, 4, 256, 64) for _ in range(3))
    ref = att(q, k, v)
    out = att(q.to(dev, torch.bfloat16), k.to(dev, torch.bfloat16), v.to(dev, torch.bfloat16))
    assert (out.float().cpu() - ref).abs().max() < 3e-2


def test_evoformer_gpu_matches_cpu():
    from shuffle_exchange_b200.ops.deepspeed4science import DS4Sci_EvoformerAttention
    dev = _cuda()
    torch.manual_seed(0)
    B, N, L, H, D = 1, 4, 96, 4, 32
    Q, K, V = (torch.randn(B, N, L, H, D) for _ in range(3))
    b1 = torch.randn(B, N, 1, 1, L)
    b2 = torch.randn(B, 1, H, L, L)
    args_c = [t.clone().requires_grad_(True) for t in (Q, K, V, b1, b2)]
    ref = DS4Sci_EvoformerAttention(*args_c[:3], args_c[3:])
    g = torch.randn_like(ref)
    ref.backward(g)
    args_g = [t.to(dev, torch.bfloat16).requires_grad_(True) for t in (Q, K, V, b1, b2)]
    out = DS4Sci_EvoformerAttention(*args_g[:3], args_g[3:])
    out.backward(g.to(dev, torch.bfloat16))
    assert (out.float().cpu() - ref).abs().max() < 5e-2
    for a, b in zip(args_g, args_c):
        scale = b.grad.abs().max().item()
        assert (a.grad.float().cpu() - b.grad).abs().max() < 0.05 * scale + 0.05
    # no biases, D = 64: the flash-attention path
    q2, k2, v2 = (torch.randn(1, 2, 128, 2, 64) for _ in range(3))
    ref2 = DS4Sci_EvoformerAttention(q2, k2, v2, [])
    out2 = DS4Sci_EvoformerAttention(q2.to(dev, torch.bfloat16), k2.to(dev, torch.bfloat16),
                                     v2.to(dev, torch.bfloat16), [])
    assert (out2.float().cpu() - ref2).abs().max() < 3e-2
    # sequence mask only (no gradient wanted for it), D = 64: rides along as the flash kernels' per-key bias
    from shuffle_exchange_b200.native import lib
    m = torch.zeros(1, 2, 1, 1, 128)
    m[0, 1, 0, 0, 100:] = -1e9
    qc, kc, vc = (t.clone().requires_grad_(True) for t in (q2, k2, v2))
    ref3 = DS4Sci_EvoformerAttention(qc, kc, vc, [m])
    g3 = torch.randn_like(ref3)
    ref3.backward(g3)
    qg, kg, vg = (t.to(dev, torch.bfloat16).requires_grad_(True) for t in (q2, k2, v2))
    before = lib.launch_count
    out3 = DS4Sci_EvoformerAttention(qg, kg, vg, [m.to(dev)])
    out3.backward(g3.to(dev, torch.bfloat16))
    assert lib.launch_count >= before + 2          # one forward and one backward entry point of the native library
    assert (out3.float().cpu() - ref3).abs().max() < 3e-2
    for a, b in ((qg, qc), (kg, kc), (vg, vc)):
        assert (a.grad.float().cpu() - b.grad).abs().max() < 0.05 * b.grad.abs().max().item() + 0.02


@pytest.mark.parametrize("masked", [False, True])
def test_transformer_layer_gpu_matches_cpu(masked):
    from shuffle_exchange_b200.ops.transformer import DeepSpeedTransformerConfig, DeepSpeedTransformerLayer
    dev = _cuda()
    torch.manual_seed(0)
    cfg = DeepSpeedTransformerConfig(batch_size=2, hidden_size=256, heads=4, attn_dropout_ratio=0.0,
                                     hidden_dropout_ratio=0.0, num_hidden_layers=2, initializer_range=0.05,
                                     pre_layer_norm=True, layer_norm_eps=1e-5)
    cpu = DeepSpeedTransformerLayer(cfg)
    import copy
    gpu = copy.deepcopy(cpu).to(dev, torch.bfloat16)
    x = torch.randn(2, 128, 256)
    mask = None
    if masked:
        mask = torch.zeros(2, 1, 1, 128)
        mask[..., 100:] = -1e4
    xc = x.clone().requires_grad_(True)
    ref = cpu(xc, mask)
    g = torch.randn_like(ref)
    ref.backward(g)
    xg = x.to(dev, torch.bfloat16).requires_grad_(True)
    out = gpu(xg, mask.to(dev, torch.bfloat16) if masked else None)
    out.backward(g.to(dev, torch.bfloat16))
    assert (out.float().cpu() - ref).abs().max() < 0.08
    assert (xg.grad.float().cpu() - xc.grad).abs().max() < 0.05 * xc.grad.abs().max() + 0.05
    gw = gpu.inter_w.grad.float().cpu()
    assert (gw - cpu.inter_w.grad).abs().max() < 0.05 * cpu.inter_w.grad.abs().max() + 0.05
    # dropout on: runs, keeps shape, differs between calls
    gpu.config.hidden_dropout_ratio = 0.1
    gpu.config.attn_dropout_ratio = 0.1
    gpu.train()
    y1, y2 = gpu(xg, None), gpu(xg, None)
    assert y1.shape == y2.shape and not torch.equal(y1, y2)


@pytest.mark.parametrize("arch", ["llama", "mixtral", "opt"])
def test_fastgen_policy_engine_gpu_bf16(arch):
    """v2 engine on the CUDA kernels (ragged embed, blocked-KV rotary, paged attention, tcgen05 GEMMs) vs HF in fp32."""
    transformers = pytest.importorskip("transformers")
    from transformers import AutoConfig, AutoModelForCausalLM
    from shuffle_exchange_b200.inference.v2.config_v2 import RaggedInferenceEngineConfig
    from shuffle_exchange_b200.inference.v2.engine_factory import build_engine_from_hf_model
    dev = _cuda()
    torch.manual_seed(0)
    common = dict(vocab_size=512, hidden_size=256, num_hidden_layers=2, num_attention_heads=4,
                  max_position_embeddings=512)
    kw = {"llama": dict(model_type="llama", intermediate_size=512, num_key_value_heads=2, **common),
          "mixtral": dict(model_type="mixtral", intermediate_size=256, num_key_value_heads=2, num_local_experts=4,
                          num_experts_per_tok=2, **common),
          "opt": dict(model_type="opt", vocab_size=512, hidden_size=256, num_hidden_layers=2, num_attention_heads=4,
                      ffn_dim=512, max_position_embeddings=512, word_embed_proj_dim=256)}[arch]
    ref_model = AutoModelForCausalLM.from_config(AutoConfig.for_model(**kw)).float().eval()
    ids = torch.randint(0, 512, (1, 40), generator=torch.Generator().manual_seed(1))
    with torch.no_grad():
        ref = ref_model(ids).logits[0]
    import copy
    gpu_model = copy.deepcopy(ref_model).to(dev, torch.bfloat16)
    eng = build_engine_from_hf_model(gpu_model, RaggedInferenceEngineConfig())
    out = eng.put([1], [ids[0, :33]]).float().cpu()
    scale = ref.abs().max().item()
    assert (out[0] - ref[32]).abs().max() < 0.06 * scale + 0.05
    out = eng.put([1], [ids[0, 33:]]).float().cpu()
    assert (out[0] - ref[39]).abs().max() < 0.06 * scale + 0.05
    assert out[0].argmax() == ref[39].argmax() or (ref[39].topk(2).values.diff().abs() < 0.05 * scale)


@pytest.mark.gpu
def test_multicast_kernels_only_for_local_kv():
    """K/V that are not provably this GPU's own memory must select the single-CTA (non-multicast) attention kernels:
    a range registered as foreign (what comm/symm.py does for every peer mapping), a host-pinned buffer, or an extent
    that runs past its allocation."""
    import ctypes
    from shuffle_exchange_b200.native import lib as nlib
    from shuffle_exchange_b200.comm import symm  # noqa: F401  (registers the signatures)
    L = nlib.load()
    B, S, Hq, Hk, D = 1, 256, 8, 2, 128
    k = torch.zeros(B, S, Hk, D, dtype=torch.bfloat16, device="cuda")
    v = torch.zeros_like(k)
    st = Hk * D
    assert L.sxb_attn_would_multicast(k.data_ptr(), v.data_ptr(), B, S, Hq, Hk, D, st, st) == 1
    assert L.sxb_range_is_local(k.data_ptr(), k.numel() * 2) == 1
    # extent beyond the allocation
    assert L.sxb_range_is_local(k.data_ptr(), 1 << 40) == 0
    # host memory
    h = torch.zeros(1 << 16, dtype=torch.bfloat16).pin_memory()
    assert L.sxb_range_is_local(h.data_ptr(), 1 << 16) == 0
    # foreign registration (stands for a peer mapping): a separate allocation so the other tests keep multicast
    big = torch.zeros(1 << 22, dtype=torch.bfloat16, device="cuda")
    fk = big[: k.numel()].view_as(k)
    assert L.sxb_attn_would_multicast(fk.data_ptr(), v.data_ptr(), B, S, Hq, Hk, D, st, st) == 1
    assert L.sxb_register_foreign_range(fk.data_ptr(), fk.numel() * 2) == 0
    assert L.sxb_attn_would_multicast(fk.data_ptr(), v.data_ptr(), B, S, Hq, Hk, D, st, st) == 0
    assert L.sxb_attn_would_multicast(k.data_ptr(), fk.data_ptr(), B, S, Hq, Hk, D, st, st) == 0
    test_multicast_kernels_only_for_local_kv._keep = big       # never freed: the range stays registered
