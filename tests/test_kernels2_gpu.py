"""GPU numerics of the second kernel batch (quantisers, sign packing, random-LTD movers, MoE routing, ragged
inference ops): every CUDA path is compared with the module's own PyTorch branch run on CPU (fp32)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _cuda():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    return torch.device("cuda")


@pytest.mark.parametrize("bits", [8, 4])
@pytest.mark.parametrize("sym", [True, False])
def test_int_quantizer_matches_cpu(bits, sym):
    from shuffle_exchange_b200.ops.quantizer import dequantize, quantize
    dev = _cuda()
    torch.manual_seed(0)
    x = torch.randn(64, 1024, dtype=torch.bfloat16)
    qc, sc = quantize(x, 64, bits, sym)
    qg, sg = quantize(x.to(dev), 64, bits, sym)
    assert torch.allclose(sg.cpu(), sc, rtol=1e-5, atol=1e-7)
    # identical codes except where x/scale sits on a rounding tie after fp32 reciprocal
    mism = (qg.cpu() != qc).float().mean().item()
    assert mism < 2e-3, mism
    y = dequantize(qg, sg, 64, bits, sym, torch.float32).cpu()
    step = (sc if sym else sc[:, 0]).max().item()
    assert (y - x.float().view(64, -1)).abs().max().item() <= 0.51 * step + 1e-6


def test_swizzle_and_quantized_reduction():
    from shuffle_exchange_b200.ops.quantizer import dequantize, quantize, quantized_reduction
    dev = _cuda()
    torch.manual_seed(1)
    x = torch.randn(32, 512, dtype=torch.bfloat16)
    qg, sg = quantize(x.to(dev), 32, 4, True, nodes=2, devs=4)
    qc, sc = quantize(x, 32, 4, True, nodes=2, devs=4)
    assert torch.allclose(sg.cpu(), sc, rtol=1e-5)
    rq, rs = quantized_reduction(qg, sg, 32, 8, 4, True, 4)
    assert (qg.cpu() != qc).float().mean().item() < 5e-3                 # same swizzled layout, same codes (up to ties)
    # reference built from the GPU's own first-hop codes so that only the reduce + requantise step is under test
    ref = dequantize(qg, sg, 32, 4, True, torch.float32).cpu().view(4, 8, 512).sum(0)
    got = dequantize(rq, rs, 8, 4, True, torch.float32).cpu()
    assert ((got - ref).abs().amax(dim=1) <= 0.51 * rs.cpu() + 1e-5).all()


@pytest.mark.parametrize("q_bits", [8, 6, 4, 12])
def test_fp_quantizer_matches_cpu(q_bits):
    from shuffle_exchange_b200.ops.fp_quantizer import fp_dequantize, fp_quantize
    dev = _cuda()
    torch.manual_seed(2)
    x = torch.randn(40, 512, dtype=torch.bfloat16)
    qc = fp_quantize(x, 512, q_bits)
    qg = fp_quantize(x.to(dev), 512, q_bits).cpu()
    assert torch.equal(qg[:, -4:], qc[:, -4:])                      # scales bit-identical
    assert (qg != qc).float().mean().item() < 2e-3
    yg = fp_dequantize(qg.to(dev), 512, q_bits, torch.float32).cpu()
    yc = fp_dequantize(qc, 512, q_bits, torch.float32)
    assert torch.allclose(yg, yc, rtol=0, atol=float(x.abs().max()) * 0.2)
    rows = torch.tensor([3, 0], dtype=torch.int32)
    sel = fp_dequantize(qg.to(dev), 512, q_bits, torch.float32, rows=rows.to(dev), groups_per_row=4).cpu()
    assert torch.equal(sel, torch.cat([yg[12:16], yg[0:4]]))


def test_packbits_roundtrip():
    from shuffle_exchange_b200.runtime.comm.compressed import packbits, unpackbits
    dev = _cuda()
    x = torch.randn(100003, device=dev)
    p = packbits(x)
    assert torch.equal(p.cpu(), packbits(x.cpu()))
    y = unpackbits(p, x.numel(), torch.float32, 0.5)
    assert torch.equal(y.cpu(), torch.where(x.cpu() >= 0, torch.tensor(0.5), torch.tensor(-0.5)))


def test_token_sort_gather_scatter():
    from shuffle_exchange_b200.ops.random_ltd.dropping_utils import token_gather, token_scatter_, token_sort_
    dev = _cuda()
    torch.manual_seed(3)
    idx = torch.stack([torch.randperm(512)[:200] for _ in range(6)]).to(torch.int32)
    got = token_sort_(idx.clone().to(dev)).cpu()
    assert torch.equal(got, torch.sort(idx, dim=-1).values)
    for batch_first in (True, False):
        full = torch.randn(6, 512, 64, dtype=torch.bfloat16) if batch_first else torch.randn(512, 6, 64, dtype=torch.bfloat16)
        part_ref = token_gather(full, got, batch_first)
        part = token_gather(full.to(dev), got.to(dev), batch_first)
        assert torch.equal(part.cpu(), part_ref)
        new = torch.randn_like(part_ref)
        a = token_scatter_(full.clone(), new, got, batch_first)
        b = token_scatter_(full.clone().to(dev), new.to(dev), got.to(dev), batch_first)
        assert torch.equal(b.cpu(), a)


def test_moe_routing_ops():
    from shuffle_exchange_b200.ops import moe_ops
    dev = _cuda()
    torch.manual_seed(4)
    T, E, K, H = 300, 8, 2, 256
    logits = torch.randn(T, E)
    x = torch.randn(T, H, dtype=torch.bfloat16)
    s_c, a_c, _, c_c = moe_ops.top_k_gating(logits, K)
    s_g, a_g, o_g, c_g = moe_ops.top_k_gating(logits.to(dev), K)
    assert torch.equal(a_g.cpu(), a_c) and torch.equal(c_g.cpu(), c_c)
    assert torch.allclose(s_g.cpu(), s_c, atol=1e-5)
    # offsets: a permutation of range(count) inside every expert
    for e in range(E):
        got = torch.sort(o_g.cpu()[a_c == e]).values
        assert torch.equal(got, torch.arange(int(c_c[e]), dtype=torch.int32))
    xg = x.to(dev).requires_grad_(True)
    sg = s_g.clone().requires_grad_(True)
    moe_in, cumsum, slots = moe_ops.moe_scatter(xg, a_g, o_g, c_g)
    assert torch.equal(cumsum.cpu(), torch.cumsum(c_c.long(), 0))
    assert torch.equal(moe_in[slots.long().reshape(-1)].cpu(), x.repeat_interleave(K, 0))
    y = moe_ops.moe_gather(moe_in * 2, sg, slots)
    ref = (2 * x.float()) * s_c.sum(1, keepdim=True)
    assert torch.allclose(y.float().cpu(), ref, atol=2e-2, rtol=2e-2)
    y.float().sum().backward()
    assert torch.allclose(xg.grad.float().cpu(), 2 * s_c.sum(1, keepdim=True).expand(T, H), atol=2e-2, rtol=2e-2)
    assert torch.allclose(sg.grad.cpu(), (2 * x.float()).sum(1, keepdim=True).expand(T, K), atol=0.5, rtol=2e-2)


def test_ragged_ops_and_engine_match_cpu():
    from shuffle_exchange_b200.inference.v2.config_v2 import RaggedInferenceEngineConfig
    from shuffle_exchange_b200.inference.v2.engine_factory import build_engine_from_model
    from shuffle_exchange_b200.models.llama import LlamaForCausalLM, llama_config
    dev = _cuda()
    torch.manual_seed(0)
    cfg = llama_config("llama-tiny", activation_checkpointing=False)
    lm = LlamaForCausalLM(cfg, device=dev, dtype=torch.bfloat16).eval()
    ec = RaggedInferenceEngineConfig(kv_block_size=16)
    ec.state_manager.num_kv_blocks = 64
    ec.state_manager.max_context = 256
    ec.state_manager.max_ragged_batch_size = 256
    ec.state_manager.max_ragged_sequence_count = 8
    eng = build_engine_from_model(lm, ec, device=dev)
    a = torch.randint(0, 1000, (70,))
    b = torch.randint(0, 1000, (33,))
    eng.put([1], [a[:40]])
    out = eng.put([1, 2], [a[40:], b]).float().cpu()
    with torch.no_grad():
        ref_a = lm(a[None].to(dev))[0, -1].float().cpu()
        ref_b = lm(b[None].to(dev))[0, -1].float().cpu()
    for got, ref in ((out[0], ref_a), (out[1], ref_b)):
        assert torch.nn.functional.cosine_similarity(got, ref, dim=0) > 0.995
        assert got.argmax() == ref.argmax() or (got - ref).abs().max() < 0.1


@pytest.mark.gpu
@pytest.mark.parametrize("q_bits", [4, 6, 8])
@pytest.mark.parametrize("rows", [1, 3, 8])
def test_weight_only_minifloat_gemv_matches_dequantised_matmul(q_bits, rows):
    """Fused weight-only GEMV (codes decoded in registers) vs dequantise + fp32 matmul of the SAME quantised weight."""
    import torch
    from shuffle_exchange_b200.ops.fp_quantizer import fp_dequantize, fp_gemv, fp_gemv_ok, fp_quantize
    torch.manual_seed(0)
    N, K, gs = 1000, 1024, 512                      # N not a multiple of the rows-per-block
    w = (torch.randn(N, K, device="cuda") * 0.1).to(torch.bfloat16)
    x = torch.randn(rows, K, device="cuda", dtype=torch.bfloat16)
    wq = fp_quantize(w, gs, q_bits)
    assert fp_gemv_ok(x[:1], N, K, gs, q_bits) and fp_gemv_ok(x, N, K, gs, q_bits) == (rows <= 4)
    y = fp_gemv(x, wq, N, K, q_bits, gs)
    wd = fp_dequantize(wq, gs, q_bits, dtype=torch.float32).view(N, K)
    ref = x.float() @ wd.t()
    err = ((y.float() - ref).norm() / ref.norm()).item()
    assert err < 5e-3, err                          # bf16 rounding of the output only


@pytest.mark.gpu
def test_w6a16_linear_uses_gemv_for_decode_and_gemm_for_prefill():
    import torch
    from shuffle_exchange_b200.inference.v2.kernels.core_ops import CUDAWf6Af16Linear
    torch.manual_seed(1)
    lin = CUDAWf6Af16Linear(group_size=512)
    w = (torch.randn(512, 1024, device="cuda") * 0.05).to(torch.bfloat16)
    wq = lin.quantize(w)
    for rows in (2, 64):
        x = torch.randn(rows, 1024, device="cuda", dtype=torch.bfloat16)
        y = lin(None, x, wq, (512, 1024))
        ref = x.float() @ w.float().t()
        assert ((y.float() - ref).norm() / ref.norm()).item() < 8e-2     # FP6 quantisation error of the weight


@pytest.mark.gpu
@pytest.mark.parametrize("num_bits", [4, 8])
@pytest.mark.parametrize("rows", [1, 3, 16])
def test_quantized_linear_decode_gemv_matches_dequantised_matmul(num_bits, rows):
    """inference/quantization: at decode sizes the layer multiplies straight from the codes; the result must equal the
    fp32 matmul with the layer's own dequantised weight (larger batches take the dequantise + GEMM path)."""
    import torch
    from shuffle_exchange_b200.inference.quantization import layers as ql
    torch.manual_seed(3)
    ql.quantized_weight_registry.clear()            # keyed by storage address: entries of freed layers must not be hit
    lin = torch.nn.Linear(1024, 1000, bias=True, device="cuda", dtype=torch.bfloat16)
    cfg = {"num_bits": num_bits, "group_size": 64, "group_dim": 1, "symmetric": False}
    q = ql.QuantizedLinear(cfg, lin)
    x = torch.randn(rows, 1024, device="cuda", dtype=torch.bfloat16)
    codes, scale, mn = q._deconcat(q.weight.data)
    assert ql.int_gemv_ok(x, codes, scale, cfg, 1024) == (rows <= ql.GEMV_MAX_ROWS)
    y = q(x)
    ref = x.float() @ q.dequantized_weight().float().t() + lin.bias.float()
    err = ((y.float() - ref).norm() / ref.norm()).item()
    # the layer's dequantised weight is rounded to bf16, the GEMV keeps code / scale + min in fp32
    assert err < 1e-2, err
