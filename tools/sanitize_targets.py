"""One small launch of every kernel family in ONE process, for a single compute-sanitizer invocation:

    python tools/sanitize_targets.py                                           # must exit 0 first
    compute-sanitizer --tool racecheck --print-limit 50 python tools/sanitize_targets.py > gpurun_out/racecheck.log

Shapes are the smallest each kernel takes (a sanitizer pass slows kernels 10-100x).  Results are checked against
PyTorch so that a run under the sanitizer that silently computes garbage is noticed too."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def rel(a, b):
    return ((a.float() - b.float()).norm() / b.float().norm().clamp_min(1e-9)).item()


def main():
    from shuffle_exchange_b200.comm.fake_peer import FakeEP, FakePeerBuffer, FakePeerGroup
    from shuffle_exchange_b200.inference.v2.kernels import ragged_ops as R
    from shuffle_exchange_b200.inference.v2.ragged.ragged_wrapper import RaggedBatchWrapper
    from shuffle_exchange_b200.ops import functional as F, moe_ops
    from shuffle_exchange_b200.ops.gemm_ext import (Gate, aligned_offsets, fp8_gemm_nt, fp8_rowwise_quant, gemm_gated,
                                                    grouped_gemm_nt)
    dev = torch.device("cuda")
    torch.manual_seed(0)
    bf = dict(device=dev, dtype=torch.bfloat16)
    done = []

    def ok(name, err, tol=2e-2):
        assert err == err and err < tol, (name, err)
        done.append(name)

    # ---- GEMMs: pair (512-multiples) and single-CTA kernels, three layouts, fp32 accumulate
    a, b = torch.randn(512, 256, **bf), torch.randn(512, 256, **bf)
    ok("gemm_nt", rel(F.gemm(a, b, 0), a.float() @ b.float().t()))
    ok("gemm_nn", rel(F.gemm(a, b.t().contiguous(), 1), a.float() @ b.float().t()))
    ok("gemm_tn", rel(F.gemm(a, b, 2), a.float().t() @ b.float()))
    s_a, s_b = torch.randn(128, 64, **bf), torch.randn(256, 64, **bf)
    ok("gemm_single", rel(F.gemm(s_a, s_b, 0), s_a.float() @ s_b.float().t()))
    acc = torch.ones(256, 256, device=dev)
    F.gemm(a, b, 2, out=acc, accumulate=True)
    ok("gemm_acc_fp32", rel(acc, a.float().t() @ b.float() + 1))
    # ---- extended GEMM: gated, grouped, fp8
    bn = torch.randn(256, 512, **bf)
    flags = torch.full(((bn.numel() * 2 + 65535) >> 16,), 3, dtype=torch.int32, device=dev)
    ok("gemm_gated", rel(gemm_gated(a, bn, 1, Gate(flags, 3, 1 << 16, bn.data_ptr())), a.float() @ bn.float()))
    counts = torch.tensor([300, 200], device=dev)
    off = aligned_offsets(counts)
    xg = torch.randn(int(off[-1]), 256, **bf)
    wg = torch.randn(2, 256, 256, **bf)
    yg = grouped_gemm_nt(xg, wg, off)
    o = off.tolist()
    ok("gemm_grouped", rel(yg[o[0]:o[0] + 300], xg[o[0]:o[0] + 300].float() @ wg[0].float().t()))
    qa, sa = fp8_rowwise_quant(a)
    qb, sb = fp8_rowwise_quant(b)
    ok("gemm_fp8", rel(fp8_gemm_nt(qa, sa, qb, sb), a.float() @ b.float().t()), 8e-2)
    # ---- attention: dense fwd + bwd, packed sequences, paged atoms (plain and GQA-packed)
    Hq, Hk, D, S = 4, 2, 128, 256
    qkv = torch.randn(1, S, (Hq + 2 * Hk) * D, **bf).requires_grad_(True)
    out = F.flash_attention_qkv(qkv, Hq, Hk, D, causal=True)
    q, k, v = F._split_qkv(qkv.detach(), Hq, Hk, D)
    ok("attn_fwd", rel(out, F.attention_ref(q.float(), k.float(), v.float(), True)))
    out.backward(torch.randn_like(out))
    assert torch.isfinite(qkv.grad).all()
    done.append("attn_bwd")
    cu = torch.tensor([0, 100, 256], dtype=torch.int32, device=dev)
    q3, k3, v3 = (t[0].contiguous().requires_grad_(True) for t in (q, k, v))
    ov = F.flash_attention_varlen(q3, k3, v3, cu, 156)
    ok("attn_varlen", rel(ov, F.attention_varlen_ref(q3.detach().float(), k3.detach().float(), v3.detach().float(), cu.cpu())))
    ov.backward(torch.randn_like(ov))
    assert torch.isfinite(q3.grad).all()

    class Seq:
        def __init__(self, seen, blocks):
            self.seen_tokens, self._b = seen, torch.tensor(blocks, dtype=torch.int32)

        def kv_blocks(self):
            return self._b
    cache = torch.randn(16, 64, 2, Hk, D, **bf)
    w = RaggedBatchWrapper(256, 4, 8, device=dev)
    w.insert_sequence(Seq(0, [3, 4, 5]), torch.zeros(150, dtype=torch.int64))
    w.insert_sequence(Seq(100, [7, 8]), torch.zeros(1, dtype=torch.int64))
    w.finalize()
    qr = torch.randn(w.current_tokens, Hq, D, **bf)
    ok("attn_atoms", rel(R.ragged_attention(qr, cache, w),
                         R.paged_attention(qr, cache, w.token_to_seq(), w.token_positions(), w.block_tables())))
    # ---- bandwidth kernels
    x = torch.randn(64, 512, **bf)
    wn = torch.randn(512, **bf)
    ok("rmsnorm", rel(F.rms_norm(x, wn, 1e-6)[0], F.rms_norm_ref(x.float(), wn.float(), 1e-6)[0]))
    g2 = torch.randn(64, 1024, **bf)
    ok("swiglu", rel(F.swiglu(g2), torch.nn.functional.silu(g2[:, :512].float()) * g2[:, 512:].float()))
    # ---- emulated multi-rank kernels (4 logical ranks, one cooperative launch each)
    world, part = 4, 1 << 12
    grp = FakePeerGroup(world, dev)
    gbuf = FakePeerBuffer(world, part * world, torch.bfloat16, dev)
    for t in gbuf.tensors:
        t.copy_(torch.randn(part * world, device=dev).to(torch.bfloat16))
    mean = torch.stack([t.float() for t in gbuf.tensors]).mean(0)
    outs = [torch.empty(part, **bf) for _ in range(world)]
    grp.rs_adam(gbuf, part, None, None, None, None, {"lr": 1e-3, "betas": (0.9, 0.95), "eps": 1e-8, "weight_decay": 0.0},
                step=1, grad_out=outs)
    torch.cuda.synchronize()
    ok("symm_reduce_scatter", max(rel(outs[r], mean[r * part:(r + 1) * part]) for r in range(world)))
    src = FakePeerBuffer(world, part, torch.bfloat16, dev)
    for t in src.tensors:
        t.copy_(torch.randn(part, device=dev).to(torch.bfloat16))
    ref = torch.stack([t.float() for t in src.tensors]).mean(0)
    sm = [torch.empty(part, **bf) for _ in range(world)]
    grp.subset_mean(src, [list(range(world))], sm, part)
    torch.cuda.synchronize()
    ok("symm_subset_mean", max(rel(t, ref) for t in sm))
    grp.subset_mean(src, [list(range(world))], sm, part, shared=True)
    torch.cuda.synchronize()
    ok("symm_subset_mean_two_phase", max(rel(t, ref) for t in sm))
    full = [torch.empty(part * world, **bf) for _ in range(world)]
    grp.all_gather(src, part, full)
    torch.cuda.synchronize()
    ok("symm_all_gather", rel(full[1], torch.cat(src.tensors)))
    E, T, Kt, H = 8, 256, 2, 256
    ep = FakeEP(world, E, H, (T * Kt * 2 // 256 + 4) * 256, dev)
    xs, asg, ofs, cnt, sc = [], [], [], [], []
    for r in range(world):
        s_, a_, o_, c_ = moe_ops.top_k_gating(torch.randn(T, E, device=dev), Kt)
        xs.append(torch.randn(T, H, **bf)); asg.append(a_); ofs.append(o_); cnt.append(c_); sc.append(s_.contiguous())
    ys = [torch.empty(T, H, **bf) for _ in range(world)]
    ep.plan(cnt)
    ep.push(xs, asg, ofs, T, Kt)
    for d in range(world):
        ep.ret.tensors[d].copy_(ep.recv.tensors[d])
    ep.pull(ys, asg, ofs, T, Kt, weights=sc)
    torch.cuda.synchronize()
    ok("moe_ep_roundtrip", max(rel(ys[r], xs[r].float() * sc[r].sum(-1, keepdim=True)) for r in range(world)))
    print("sanitize targets ok:", ", ".join(done))


if __name__ == "__main__":
    main()
