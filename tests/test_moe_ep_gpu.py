"""Expert-parallel exchange kernels (csrc/moe_ep.cu) and the fused MoE layer on ONE GPU.

* plan / push / pull with 2 and 4 logical ranks emulated as cooperative grids, against an index-level PyTorch
  statement of "token row t goes to the owning rank's expert-contiguous row, comes back weighted";
* the layer (``DroplessMoE``, single rank: same kernels with world = 1) against its unfused path, forward and
  backward, bf16 and fp8 expert GEMMs."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    return torch.device("cuda")


@pytest.mark.parametrize("world,L", [(2, 2), (4, 1), (4, 2)])
def test_emulated_dispatch_and_combine(world, L):
    from shuffle_exchange_b200.comm.fake_peer import FakeEP
    from shuffle_exchange_b200.ops import moe_ops
    dev = _dev()
    torch.manual_seed(world * 10 + L)
    E, T, K, H = world * L, 300, 2, 512
    cap = ((T * K * world) // 256 + L + 1) * 256
    ep = FakeEP(world, E, H, cap, dev)
    xs, assigns, offs, cnts, scores = [], [], [], [], []
    for r in range(world):
        logits = torch.randn(T, E, device=dev)
        sc, a, o, c = moe_ops.top_k_gating(logits, K)
        xs.append(torch.randn(T, H, device=dev).to(torch.bfloat16))
        assigns.append(a); offs.append(o); cnts.append(c); scores.append(sc.contiguous())
    ep.plan(cnts)
    ep.push(xs, assigns, offs, T, K)
    torch.cuda.synchronize()
    # oracle: expert e lives on rank e // L; its rows are ordered by source rank, then arrival order
    cm = torch.stack(cnts).cpu()                                   # [src, E]
    tot = cm.sum(0)
    for d in range(world):
        tab = ep.tables[d].cpu()
        goff = tab[E: E + L + 1].tolist()
        assert tab[E + 2 * L + 1] == 0                             # no overflow
        off = 0
        for le in range(L):
            e = d * L + le
            assert goff[le] == off and tab[E + L + 1 + le] == tot[e]
            off += (int(tot[e]) + 255) // 256 * 256
        assert goff[L] == off
        recv = ep.recv.tensors[d].view(cap, H)
        for le in range(L):
            e = d * L + le
            base = goff[le]
            for s in range(world):
                sel = (assigns[s] == e)
                rows_t = sel.nonzero()[:, 0]
                slot = offs[s][sel].long()
                want = torch.zeros(int(cm[s, e]), H, device=dev, dtype=torch.bfloat16)
                want[slot] = xs[s][rows_t]
                got = recv[base: base + int(cm[s, e])]
                assert torch.equal(got, want), (d, e, s)
                base += int(cm[s, e])
            assert not bool(recv[base: goff[le + 1]].any()), "padding rows must be zero"
    # expert compute stand-in: out row = 2 * in row  -> return buffers; combine with the routing weights
    for d in range(world):
        ep.ret.tensors[d].copy_(ep.recv.tensors[d] * 2)
    ys = [torch.empty(T, H, device=dev, dtype=torch.bfloat16) for _ in range(world)]
    keeps = [torch.empty(T * K, H, device=dev, dtype=torch.bfloat16) for _ in range(world)]
    ep.pull(ys, assigns, offs, T, K, weights=scores, keep_rows=keeps)
    torch.cuda.synchronize()
    for r in range(world):
        want = (2 * xs[r].float()) * scores[r].sum(1, keepdim=True)
        assert torch.allclose(ys[r].float(), want, rtol=2e-2, atol=2e-2)
        assert torch.equal(keeps[r].view(T, K, H)[:, 0], (xs[r] * 2))
    # backward of the combine: push score * dy, d_score = <dy, kept rows>
    ep.plan(None)
    dys = [torch.randn(T, H, device=dev).to(torch.bfloat16) for _ in range(world)]
    dscores = [torch.empty(T, K, device=dev) for _ in range(world)]
    ep.push(dys, assigns, offs, T, K, weights=scores, pulled=keeps, d_scores=dscores)
    torch.cuda.synchronize()
    for r in range(world):
        want = (dys[r].float()[:, None, :] * keeps[r].view(T, K, H).float()).sum(-1)
        assert torch.allclose(dscores[r], want, rtol=1e-3, atol=1e-1)
    # rows that arrived at rank 0 for its first expert from source 0: score * dy
    tab = ep.tables[0].cpu()
    sel = (assigns[0] == 0)
    if bool(sel.any()):
        rows_t = sel.nonzero()[:, 0]
        slot = offs[0][sel].long()
        got = ep.recv.tensors[0].view(cap, H)[int(tab[E]): int(tab[E]) + int(cm[0, 0])]
        want = torch.zeros_like(got)
        want[slot] = (dys[0][rows_t].float() * scores[0][sel][:, None]).to(torch.bfloat16)
        assert torch.allclose(got.float(), want.float(), rtol=1e-2, atol=1e-2)


@pytest.mark.parametrize("num_experts,fp8", [(4, False), (1, False), (4, True)])
def test_fused_layer_matches_unfused_single_rank(num_experts, fp8):
    from shuffle_exchange_b200.moe.dropless import DroplessMoE
    dev = _dev()
    torch.manual_seed(7)
    H, F_, T = 512, 1024, 600
    ref = DroplessMoE(H, F_, num_experts, ep_size=1, k=min(2, num_experts), device=dev, dtype=torch.bfloat16,
                      fused_exchange=False)
    ref.reset_parameters(0.05)
    fus = DroplessMoE(H, F_, num_experts, ep_size=1, k=min(2, num_experts), device=dev, dtype=torch.bfloat16,
                      fused_exchange=True, fp8_experts=fp8)
    fus.load_state_dict(ref.state_dict())
    x = (torch.randn(T, H, device=dev) * 0.5).to(torch.bfloat16)
    xr, xf = x.clone().requires_grad_(), x.clone().requires_grad_()
    yr, _, cr = ref(xr)
    yf, _, cf = fus(xf)
    assert torch.equal(cr, cf)
    rel = lambda a, b: ((a.float() - b.float()).norm() / b.float().norm().clamp_min(1e-6)).item()
    if not fp8:
        assert torch.allclose(yf.float(), yr.float(), rtol=5e-2, atol=5e-3), (yf.float() - yr.float()).abs().max()
    else:
        assert rel(yf, yr) < 0.08, rel(yf, yr)          # two chained e4m3 products
    g = torch.randn_like(x)
    yr.backward(g)
    yf.backward(g)
    fus.check_overflow()
    assert rel(xf.grad, xr.grad) < (0.02 if not fp8 else 0.1)
    for (n, pr), (_, pf) in zip(ref.named_parameters(), fus.named_parameters()):
        assert rel(pf.grad, pr.grad) < (0.02 if not fp8 else 0.1), n


@pytest.mark.gpu
@pytest.mark.parametrize("T,E,K", [(300, 8, 2), (17, 4, 1), (1024, 8, 2)])
def test_grouped_inference_moe_matches_per_expert_loop(T, E, K):
    """FastGen MoE body: plan / push / grouped GEMMs / pull on one rank, no host synchronisation, vs the loop over
    experts in fp32."""
    import torch
    from shuffle_exchange_b200.moe import fused_ep
    from shuffle_exchange_b200.ops import functional as F, moe_ops
    torch.manual_seed(0)
    H, Fi = 256, 512
    dev = "cuda"
    x = torch.randn(T, H, device=dev, dtype=torch.bfloat16)
    w1 = (torch.randn(E, 2 * Fi, H, device=dev) * 0.05).to(torch.bfloat16)
    w2 = (torch.randn(E, H, Fi, device=dev) * 0.05).to(torch.bfloat16)
    logits = torch.randn(T, E, device=dev)
    scores, assign, offsets, counts = moe_ops.top_k_gating(logits, K, renormalize=True)
    assert fused_ep.grouped_inference_ok(x, w1, w2)
    for _ in range(2):                                    # second call reuses the cached buffers
        y = fused_ep.grouped_experts_inference(x, scores, assign, offsets, counts, w1, w2, F.swiglu, K)
    ref = torch.zeros(T, H, device=dev)
    for t in range(T):
        for j in range(K):
            e = int(assign[t, j])
            h = x[t].float() @ w1[e].float().t()
            a = torch.nn.functional.silu(h[:Fi]) * h[Fi:]
            ref[t] += float(scores[t, j]) * (a.to(torch.bfloat16).float() @ w2[e].float().t())
    err = ((y.float() - ref).norm() / ref.norm()).item()
    assert err < 2e-2, err
