"""2+ GPU check of the fused expert-parallel MoE path (symmetric-memory dispatch / combine + grouped GEMM, moe/fused_ep.py)
against the NCCL all_to_all_single path of the same layer, forward and backward, several layers in a row (buffers and
epochs are shared between layers).

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29559 tools/moe_ep_check.py
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch.distributed as td
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    import shuffle_exchange_b200 as sxb
    sxb.init_distributed(dist_backend="nccl")
    from shuffle_exchange_b200.moe.dropless import DroplessMoE
    dev = torch.device("cuda", local)
    H, Fi, T, n_layers = 1024, 2048, 3000, 3
    res = {"world": world}
    for E, fp8 in ((world * 2, False), (world, False), (world, True)):
        torch.manual_seed(1)
        ref = [DroplessMoE(H, Fi, E, ep_size=world, k=2, device=dev, dtype=torch.bfloat16, fused_exchange=False)
               for _ in range(n_layers)]
        fus = [DroplessMoE(H, Fi, E, ep_size=world, k=2, device=dev, dtype=torch.bfloat16, fused_exchange=True,
                           fp8_experts=fp8) for _ in range(n_layers)]
        for i, (a, b) in enumerate(zip(ref, fus)):
            torch.manual_seed(100 * i + rank)            # expert weights differ per rank (each rank owns other experts)
            a.reset_parameters(0.03)
            with torch.no_grad():
                td.broadcast(a.wg, 0)                     # the router is replicated
            b.load_state_dict(a.state_dict())
            a._create_process_groups()
            b._create_process_groups()
        torch.manual_seed(7 + rank)
        x = (torch.randn(T, H, device=dev) * 0.5).to(torch.bfloat16)
        g = torch.randn(T, H, device=dev).to(torch.bfloat16)
        outs = []
        for layers in (ref, fus):
            xi = x.clone().requires_grad_()
            h = xi
            for m in layers:
                y, _, _ = m(h)
                h = (h + y).to(torch.bfloat16)
            h.backward(g)
            outs.append((h.detach(), xi.grad, [p.grad for m in layers for p in m.parameters()]))
        fus[0].check_overflow()
        rel = lambda a, b: float(((a.float() - b.float()).norm() / b.float().norm().clamp_min(1e-6)))
        key = f"E{E}_fp8{int(fp8)}"
        errs = {"y": rel(outs[1][0], outs[0][0]), "dx": rel(outs[1][1], outs[0][1]),
                "dw_max": max(rel(a, b) for a, b in zip(outs[1][2], outs[0][2]))}
        t = torch.tensor([errs["y"], errs["dx"], errs["dw_max"]], device=dev)
        td.all_reduce(t, op=td.ReduceOp.MAX)
        res[key] = dict(zip(("y", "dx", "dw_max"), [float(v) for v in t]))
        lim = 0.02 if not fp8 else 0.12
        res[key]["ok"] = all(v < lim for v in res[key].values() if isinstance(v, float))
    # timing of one fused layer vs the NCCL layer (E = world, Mixtral-like shapes scaled down)
    from shuffle_exchange_b200.native import lib as n
    import shuffle_exchange_b200.moe.fused_ep  # noqa: F401
    res["watchdog"] = n.hang_info("moe_ep")
    res["ok"] = all(v.get("ok", True) for v in res.values() if isinstance(v, dict)) and res["watchdog"] is None
    if rank == 0:
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        json.dump(res, open(os.path.join(ROOT, "gpurun_out", "moe_ep_check.json"), "w"), indent=1)
        print(json.dumps(res))
    td.barrier()
    td.destroy_process_group()


if __name__ == "__main__":
    main()
