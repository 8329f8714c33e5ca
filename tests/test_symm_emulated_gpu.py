"""Peer-memory collectives on ONE GPU: ``world`` logical ranks emulated as one cooperative grid per collective
(comm/fake_peer.py).  Each kernel body is the one the multi-GPU path runs; here it is compared with its PyTorch
statement: reduce-scatter + AdamW (K1), plain reduce-scatter, subset mean (K9), all-gather with arrival flags (K2
gather half) and the row all-to-all (K4-K7 transport)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _close(a, b, rtol, atol):
    a, b = a.float(), b.float()
    bad = (a - b).abs() > atol + rtol * b.abs()
    assert not bool(bad.any()), f"{int(bad.sum())} of {a.numel()} elements off; max abs err {(a - b).abs().max().item():.3e}"


@pytest.mark.parametrize("world", [2, 4, 8])
def test_emulated_collectives_match_pytorch(world):
    from shuffle_exchange_b200.comm.fake_peer import FakePeerBuffer, FakePeerGroup
    from shuffle_exchange_b200.native import lib as n
    torch.manual_seed(world)
    dev = torch.device("cuda", 0)
    grp = FakePeerGroup(world, dev)
    part = (1 << 15) + 64            # not a multiple of the grid stride: exercises the loop tails
    total = part * world
    # ---- K1: reduce-scatter + AdamW (two steps, second with non-zero moments)
    g = FakePeerBuffer(world, total, torch.bfloat16, dev)
    cfg = {"lr": 1e-2, "betas": (0.9, 0.95), "eps": 1e-8, "weight_decay": 0.1}
    hp = [torch.randn(part, device=dev) for _ in range(world)]
    m = [torch.zeros(part, device=dev) for _ in range(world)]
    v = [torch.zeros(part, device=dev) for _ in range(world)]
    lp = [torch.empty(part, device=dev, dtype=torch.bfloat16) for _ in range(world)]
    ref = [torch.nn.Parameter(h.clone()) for h in hp]
    opts = [torch.optim.AdamW([r], lr=1e-2, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1) for r in ref]
    for step in (1, 2):
        for t in g.tensors:
            t.copy_(torch.randn(total, device=dev).to(torch.bfloat16))
        mean = torch.stack([t.float() for t in g.tensors]).mean(0)
        for r in range(world):
            ref[r].grad = mean[r * part:(r + 1) * part].clone()
            opts[r].step()
        grp.rs_adam(g, part, hp, m, v, lp, cfg, step=step)
        torch.cuda.synchronize()
        for r in range(world):
            _close(hp[r], ref[r].data, 1e-5, 1e-6)
            _close(lp[r], ref[r].data, 8e-3, 1e-6)          # bf16 rounding of the written shard
    # ---- plain reduce-scatter (mean), bf16 output
    out = [torch.empty(part, device=dev, dtype=torch.bfloat16) for _ in range(world)]
    grp.rs_adam(g, part, None, None, None, None, cfg, step=1, grad_out=out)
    torch.cuda.synchronize()
    for r in range(world):
        _close(out[r], mean[r * part:(r + 1) * part], 8e-3, 1e-6)
    # ---- K9: two rings (even / odd ranks); every rank averages over its own ring, as in the shuffle method
    rings = [list(range(0, world, 2)), list(range(1, world, 2))]
    src = FakePeerBuffer(world, part, torch.bfloat16, dev)
    for t in src.tensors:
        t.copy_(torch.randn(part, device=dev).to(torch.bfloat16))
    sm_out = [torch.empty(part, device=dev, dtype=torch.bfloat16) for _ in range(world)]
    sm_refs = [torch.stack([src.tensors[r].float() for r in ring]).mean(0) for ring in rings]
    grp.subset_mean(src, rings, sm_out, part)
    torch.cuda.synchronize()
    for ring, sm_ref in zip(rings, sm_refs):
        for r in ring:
            _close(sm_out[r], sm_ref, 8e-3, 1e-6)
    # two-phase form (every member of a ring has the same list): each member reduces one slice into its own buffer,
    # the reduced slices are exchanged after an in-kernel barrier; all members end with bit-identical results
    for o in sm_out:
        o.zero_()
    grp.subset_mean(src, rings, sm_out, part, shared=True)
    torch.cuda.synchronize()
    for ring, sm_ref in zip(rings, sm_refs):
        for r in ring:
            _close(sm_out[r], sm_ref, 8e-3, 1e-6)
            assert torch.equal(sm_out[r], sm_out[ring[0]])
    # ---- K2 gather half: exact copy + every chunk flag published
    sh = FakePeerBuffer(world, part, torch.bfloat16, dev)
    for t in sh.tensors:
        t.copy_(torch.randn(part, device=dev).to(torch.bfloat16))
    full = [torch.empty(total, device=dev, dtype=torch.bfloat16) for _ in range(world)]
    chunk = 1 << 14
    nchunks = (total * 2 + chunk - 1) // chunk
    flags = [torch.zeros(nchunks, dtype=torch.int32, device=dev) for _ in range(world)]
    grp.all_gather(sh, part, full, chunk_bytes=chunk, chunk_flags=flags, flag_value=7)
    torch.cuda.synchronize()
    ag_ref = torch.cat(sh.tensors)
    for r in range(world):
        assert torch.equal(full[r], ag_ref)
        assert bool((flags[r] == 7).all())
    # ---- row all-to-all with uneven counts
    H = 256
    gen = torch.Generator().manual_seed(world)
    counts = torch.randint(0, 40, (world, world), generator=gen)
    srcs, send_off = [], []
    for r in range(world):
        send_off.append(torch.cumsum(counts[r], 0) - counts[r])
        srcs.append(torch.randn(int(counts[r].sum()) + 1, H, device=dev).to(torch.bfloat16))
    recv_off = [[int(counts[:r, d].sum()) for d in range(world)] for r in range(world)]     # rank r's rows inside d
    cap = int(counts.sum(0).max()) + 1
    dst = FakePeerBuffer(world, cap * H, torch.bfloat16, dev)
    grp.a2a_rows(dst, srcs, [s.tolist() for s in send_off], recv_off, counts.tolist(), H * 2)
    torch.cuda.synchronize()
    for d in range(world):
        want = torch.cat([srcs[r][int(send_off[r][d]): int(send_off[r][d]) + int(counts[r][d])] for r in range(world)])
        got = dst.tensors[d].view(cap, H)[: want.shape[0]]
        assert torch.equal(got, want), d
    assert n.hang_info("symm") is None
