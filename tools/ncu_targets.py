"""Short single-GPU workload that launches each kernel family once per iteration, for ``ncu`` (one capture per call):

    python tools/ncu_targets.py && ncu --set full --clock-control none --import-source on \
        -k regex:'gemm_pair_x_kernel|gemm_pair_kernel|rs_adam_emu|push_emu|pull_emu|attn_fwd_kernel|attn_bwd|all_gather_emu|fp_gemv' \
        -s 36 -c 18 -o gpurun_out/prof_r2 python tools/ncu_targets.py

Kernels: base pair GEMM (qkv fwd shape), extended GEMM gated NN / grouped NT / fp8 NT, emulated (4 logical ranks) fused
reduce-scatter+AdamW, all-gather with flags, MoE push / pull, attention forward with the row-scatter epilogue."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    from shuffle_exchange_b200.comm.fake_peer import FakeEP, FakePeerBuffer, FakePeerGroup
    from shuffle_exchange_b200.ops import functional as F, moe_ops
    from shuffle_exchange_b200.ops.gemm_ext import (Gate, aligned_offsets, fp8_gemm_nt, fp8_rowwise_quant, gemm_gated,
                                                    grouped_gemm_nt)
    from shuffle_exchange_b200.sequence.fused_ulysses import UlyssesPeers, _Stage
    dev = torch.device("cuda")
    torch.manual_seed(0)
    M, N, K = 8192, 6144, 4096
    a = torch.randn(M, K, device=dev).to(torch.bfloat16)
    b = torch.randn(N, K, device=dev).to(torch.bfloat16)
    bn = torch.randn(K, N, device=dev).to(torch.bfloat16)
    flags = torch.full(((bn.numel() * 2 + (1 << 20) - 1) >> 20,), 2, dtype=torch.int32, device=dev)
    gate = Gate(flags, 2, 1 << 20, bn.data_ptr())
    counts = torch.tensor([5000, 3000, 4000, 4384], device=dev)
    off = aligned_offsets(counts)
    xg = torch.randn(int(off[-1]), K, device=dev).to(torch.bfloat16)
    wg = torch.randn(4, 2048, K, device=dev).to(torch.bfloat16)
    qa, sa = fp8_rowwise_quant(a)
    qb, sb = fp8_rowwise_quant(b)
    world = 4
    grp = FakePeerGroup(world, dev)
    part = 1 << 22
    g = FakePeerBuffer(world, part * world, torch.bfloat16, dev)
    hp = [torch.randn(part, device=dev) for _ in range(world)]
    m = [torch.zeros(part, device=dev) for _ in range(world)]
    v = [torch.zeros(part, device=dev) for _ in range(world)]
    lp = [torch.empty(part, device=dev, dtype=torch.bfloat16) for _ in range(world)]
    cfg = {"lr": 1e-3, "betas": (0.9, 0.95), "eps": 1e-8, "weight_decay": 0.1}
    sh = FakePeerBuffer(world, part, torch.bfloat16, dev)
    full = [torch.empty(part * world, device=dev, dtype=torch.bfloat16) for _ in range(world)]
    agf = [torch.zeros((part * world * 2) >> 20, dtype=torch.int32, device=dev) for _ in range(world)]
    E, T, Kt, H = 8, 4096, 2, 4096
    ep = FakeEP(world, E, H, (T * Kt * 2 // 256 + 4) * 256, dev)
    xs, asg, ofs, cnt, sc = [], [], [], [], []
    for r in range(world):
        s_, a_, o_, c_ = moe_ops.top_k_gating(torch.randn(T, E, device=dev), Kt)
        xs.append(torch.randn(T, H, device=dev).to(torch.bfloat16)); asg.append(a_); ofs.append(o_); cnt.append(c_); sc.append(s_.contiguous())
    ys = [torch.empty(T, H, device=dev, dtype=torch.bfloat16) for _ in range(world)]
    Hq, Hk, D, Sl = 8, 2, 128, 4096
    up = UlyssesPeers(None, Sl, Hq, Hk, D, emulate_world=2)
    wqkv = (torch.randn((Hq + 2 * Hk) * D, 4096, device=dev) / 64).to(torch.bfloat16)
    cos, sin = F.rope_tables(2 * Sl, D, 10000.0, dev)
    stages = [_Stage(up, r, True) for r in range(2)]
    xu = [torch.randn(Sl, 4096, device=dev).to(torch.bfloat16) for _ in range(2)]
    qkv_d = (torch.randn(1, 8192, (32 + 16) * 128, device=dev) / 4).to(torch.bfloat16).requires_grad_(True)
    # ragged inference: one 2048-token prefill + 32 decodes at 1024 (atoms; short atoms GQA-packed), FP6 weight-only GEMV
    from shuffle_exchange_b200.inference.v2.kernels import ragged_ops as R
    from shuffle_exchange_b200.inference.v2.ragged.ragged_wrapper import RaggedBatchWrapper
    from shuffle_exchange_b200.ops.fp_quantizer import fp_gemv, fp_quantize

    class _Seq:
        def __init__(self, seen, blocks):
            self.seen_tokens, self._b = seen, torch.tensor(blocks, dtype=torch.int32)

        def kv_blocks(self):
            return self._b
    plan_rb = [(0, 2048)] + [(1024, 1)] * 32
    cache = torch.randn(sum((s_ + n_ + 63) // 64 for s_, n_ in plan_rb) + 1, 64, 2, 8, 128, device=dev, dtype=torch.bfloat16)
    rb = RaggedBatchWrapper(sum(n_ for _, n_ in plan_rb), len(plan_rb), 33, device=dev)
    nxt = 0
    for seen_, n_ in plan_rb:
        need = (seen_ + n_ + 63) // 64
        rb.insert_sequence(_Seq(seen_, list(range(nxt, nxt + need))), torch.zeros(n_, dtype=torch.int64))
        nxt += need
    rb.finalize()
    q_rb = torch.randn(rb.current_tokens, 32, 128, device=dev, dtype=torch.bfloat16)
    w6 = fp_quantize((torch.randn(28672, 4096, device=dev) * 0.05).to(torch.bfloat16), 512, 6)
    x6 = torch.randn(1, 4096, device=dev, dtype=torch.bfloat16)
    for it in range(4):
        R.ragged_attention(q_rb, cache, rb)
        fp_gemv(x6, w6, 28672, 4096, 6, 512)
        # dense causal attention at the Llama-3-8B layer shape: forward, dQ and dK/dV kernels
        o_d = F.flash_attention_qkv(qkv_d, 32, 8, 128, causal=True)
        o_d.backward(o_d.detach())
        qkv_d.grad = None
        F.gemm(a, b, 0)
        gemm_gated(a, bn, 1, gate)
        grouped_gemm_nt(xg, wg, off)
        fp8_gemm_nt(qa, sa, qb, sb)
        grp.rs_adam(g, part, hp, m, v, lp, cfg, step=it + 1)
        grp.all_gather(sh, part, full, chunk_bytes=1 << 20, chunk_flags=agf, flag_value=it + 1)
        ep.plan(cnt)
        ep.push(xs, asg, ofs, T, Kt)
        for d in range(world):
            ep.ret.tensors[d].copy_(ep.recv.tensors[d])
        ep.pull(ys, asg, ofs, T, Kt, weights=sc)
        for r in range(2):
            stages[r].fwd1(xu[r], wqkv)
        for r in range(2):
            stages[r].fwd2(cos, sin)
        torch.cuda.synchronize()
    print("ncu targets ok")


if __name__ == "__main__":
    main()
