"""One-GPU check of the peer-memory collectives in fake-peer mode (N logical ranks on one device).

    python tools/symm_fake_peer_check.py [world]          # default 4; writes gpurun_out/symm_fake_peer.json

Compares reduce-scatter+AdamW, plain reduce-scatter, subset mean and all-gather with their PyTorch statements."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def rel(a, b):
    return float((a.float() - b.float()).abs().max() / b.float().abs().max().clamp_min(1e-6))


def main():
    from shuffle_exchange_b200.comm.fake_peer import FakePeerBuffer, FakePeerGroup
    from shuffle_exchange_b200.native import lib as n
    world = int(sys.argv[1]) if len(sys.argv) > 1 else 4
    torch.manual_seed(0)
    dev = torch.device("cuda", 0)
    grp = FakePeerGroup(world, dev)
    part = 1 << 16
    total = part * world
    res = {"world": world}
    # ---- reduce-scatter + AdamW
    g = FakePeerBuffer(world, total, torch.bfloat16, dev)
    for t in g.tensors:
        t.copy_(torch.randn(total, device=dev).to(torch.bfloat16))
    mean = torch.stack([t.float() for t in g.tensors]).mean(0)
    cfg = {"lr": 1e-2, "betas": (0.9, 0.95), "eps": 1e-8, "weight_decay": 0.1}
    hp = [torch.randn(part, device=dev) for _ in range(world)]
    m = [torch.zeros(part, device=dev) for _ in range(world)]
    v = [torch.zeros(part, device=dev) for _ in range(world)]
    lp = [torch.empty(part, device=dev, dtype=torch.bfloat16) for _ in range(world)]
    ref_hp = []
    for r in range(world):
        p, gr = hp[r].clone(), mean[r * part:(r + 1) * part]
        mm, vv = 0.1 * gr, 0.05 * gr * gr
        upd = (mm / (1 - 0.9)) / ((vv / (1 - 0.95)).sqrt() + 1e-8) + 0.1 * p
        ref_hp.append(p - 1e-2 * upd)
    grp.rs_adam(g, part, hp, m, v, lp, cfg, step=1)
    torch.cuda.synchronize()
    res["rs_adam_err"] = max(rel(hp[r], ref_hp[r]) for r in range(world))
    res["rs_adam_lp_err"] = max(rel(lp[r], ref_hp[r]) for r in range(world))
    # ---- plain reduce-scatter (mean)
    out = [torch.empty(part, device=dev, dtype=torch.bfloat16) for _ in range(world)]
    grp.rs_adam(g, part, None, None, None, None, cfg, step=1, grad_out=out)
    torch.cuda.synchronize()
    res["rs_only_err"] = max(rel(out[r], mean[r * part:(r + 1) * part]) for r in range(world))
    # ---- subset mean over the even ranks
    members = list(range(0, world, 2))
    src = FakePeerBuffer(world, part, torch.bfloat16, dev)
    for t in src.tensors:
        t.copy_(torch.randn(part, device=dev).to(torch.bfloat16))
    sm_out = [torch.empty(part, device=dev, dtype=torch.bfloat16) for _ in range(world)]
    grp.subset_mean(src, members, sm_out, part)
    torch.cuda.synchronize()
    sm_ref = torch.stack([src.tensors[r].float() for r in members]).mean(0)
    res["subset_mean_err"] = max(rel(sm_out[r], sm_ref) for r in members)
    # ---- all-gather
    sh = FakePeerBuffer(world, part, torch.bfloat16, dev)
    for t in sh.tensors:
        t.copy_(torch.randn(part, device=dev).to(torch.bfloat16))
    full = [torch.empty(total, device=dev, dtype=torch.bfloat16) for _ in range(world)]
    grp.all_gather(sh, part, full)
    torch.cuda.synchronize()
    ag_ref = torch.cat(sh.tensors)
    res["all_gather_err"] = max(rel(full[r], ag_ref) for r in range(world))
    hi = getattr(n, "hang_info", lambda *_: None)("symm")
    res["watchdog"] = hi
    res["ok"] = bool(hi is None and res["rs_adam_err"] < 1e-5 and res["rs_only_err"] < 1e-2
                     and res["subset_mean_err"] < 1e-2 and res["all_gather_err"] == 0.0)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(res, open(os.path.join(ROOT, "gpurun_out", "symm_fake_peer.json"), "w"), indent=1)
    print(json.dumps(res))
    return 0 if res["ok"] else 1


if __name__ == "__main__":
    sys.exit(main())
