"""Multi-GPU check of the symmetric-memory fused collectives against their NCCL statements.

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29555 tools/symm_test.py

Each rank prints nothing; rank 0 writes gpurun_out/symm_test.json (numerics + device-timed ms, max over ranks).
"""
import json
import os
import sys

import torch
import torch.distributed as td

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def timed(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    td.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda")
    td.all_reduce(t, op=td.ReduceOp.MAX)
    return float(t)


def rel(a, b):
    return float((a.float() - b.float()).abs().max() / b.float().abs().max().clamp_min(1e-6))


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    td.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
    from shuffle_exchange_b200.comm import symm
    from shuffle_exchange_b200.native import lib as n
    res = {"world": world}
    sg = symm.SymmGroup.get(None)
    torch.manual_seed(100 + rank)

    # ---- K1: reduce-scatter + AdamW ------------------------------------------------------------------
    part = 1 << 22                      # 4M elements per shard
    total = part * world
    gbuf = sg.buffer("grads", total, torch.bfloat16)
    res["multicast"] = bool(gbuf.multicast_ptr)
    g = torch.randn(total, device="cuda").to(torch.bfloat16)
    gbuf.tensor.copy_(g)
    hp = torch.randn(part, device="cuda")
    m, v = torch.zeros_like(hp), torch.zeros_like(hp)
    lp = torch.empty(part, device="cuda", dtype=torch.bfloat16)
    cfg = dict(lr=1e-2, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1, bias_correction=True)
    # NCCL statement
    ref_g = torch.empty(part, device="cuda", dtype=torch.bfloat16)
    td.reduce_scatter_tensor(ref_g, g, op=td.ReduceOp.AVG)
    hp_ref = hp.clone().requires_grad_()
    opt = torch.optim.AdamW([hp_ref], lr=1e-2, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1)
    hp_ref.grad = ref_g.float()
    opt.step()
    for use_mc in ([False, True] if gbuf.multicast_ptr else [False]):
        h2, m2, v2 = hp.clone(), m.clone(), v.clone()
        sg.rs_adam(gbuf, part, h2, m2, v2, lp, cfg, step=1, use_multicast=use_mc)
        torch.cuda.synchronize()
        td.barrier()
        res[f"rs_adam_err_mc{int(use_mc)}"] = rel(h2, hp_ref.detach())
        go = torch.empty(part, device="cuda", dtype=torch.bfloat16)
        sg.rs_adam(gbuf, part, None, None, None, None, cfg, step=1, grad_out=go, use_multicast=use_mc)
        torch.cuda.synchronize()
        res[f"rs_only_err_mc{int(use_mc)}"] = rel(go, ref_g)
        res[f"rs_adam_ms_mc{int(use_mc)}"] = timed(lambda: sg.rs_adam(gbuf, part, h2, m2, v2, lp, cfg, step=2, use_multicast=use_mc))

    def nccl_path():
        td.reduce_scatter_tensor(ref_g, g, op=td.ReduceOp.AVG)
        n.call("sxb_adam_flat", hp.data_ptr(), ref_g.data_ptr(), m.data_ptr(), v.data_ptr(), lp.data_ptr(), part,
               n.BF16, n.BF16, 1e-2, 0.9, 0.95, 1e-8, 0.1, 2, 1, 1, 1.0, None, None, n.stream_ptr())
    res["nccl_rs_plus_adam_ms"] = timed(nccl_path)

    # ---- K9: subset mean ------------------------------------------------------------------------------
    nn_ = 1 << 22
    sb = sg.buffer("avg", nn_, torch.bfloat16)
    x = torch.randn(nn_, device="cuda").to(torch.bfloat16)
    sb.tensor.copy_(x)
    out = torch.empty_like(x)
    members = list(range(world))
    sg.subset_mean(sb, members, out, nn_)
    ref = x.float().clone()
    td.all_reduce(ref)
    ref /= world
    torch.cuda.synchronize()
    res["subset_mean_err"] = rel(out, ref)
    res["subset_mean_ms"] = timed(lambda: sg.subset_mean(sb, members, out, nn_))
    xr = x.clone()
    res["nccl_allreduce_ms"] = timed(lambda: td.all_reduce(xr))

    # ---- K2: all-gather with arrival flags -------------------------------------------------------------
    shard = sg.buffer("ag", part, torch.bfloat16)
    sh = torch.randn(part, device="cuda").to(torch.bfloat16)
    shard.tensor.copy_(sh)
    dst = torch.empty(total, device="cuda", dtype=torch.bfloat16)
    flags = torch.zeros((total * 2 + (1 << 20) - 1) >> 20, dtype=torch.int32, device="cuda")
    ep = sg.all_gather(shard, part, dst, 1 << 20, flags)
    ref_ag = torch.empty_like(dst)
    td.all_gather_into_tensor(ref_ag, sh)
    torch.cuda.synchronize()
    res["all_gather_err"] = rel(dst, ref_ag)
    res["all_gather_flags_ok"] = bool((flags == ep).all())
    res["all_gather_ms"] = timed(lambda: sg.all_gather(shard, part, dst, 1 << 20, flags))
    res["nccl_all_gather_ms"] = timed(lambda: td.all_gather_into_tensor(ref_ag, sh))

    # ---- a2a rows ---------------------------------------------------------------------------------------
    rows_per, M = 4096, 4096
    recv = sg.buffer("a2a", rows_per * world * M, torch.bfloat16)
    src = torch.randn(rows_per * world, M, device="cuda").to(torch.bfloat16)
    send_off = [d * rows_per for d in range(world)]
    recv_off = [rank * rows_per for _ in range(world)]
    cnt = [rows_per] * world
    ep = sg.a2a_rows(recv, src, send_off, recv_off, cnt, M * 2)
    sg.wait_landed(recv, ep)
    ref_a2a = torch.empty_like(src)
    td.all_to_all_single(ref_a2a, src)
    torch.cuda.synchronize()
    td.barrier()
    res["a2a_err"] = rel(recv.tensor[: rows_per * world * M].view(-1, M), ref_a2a)

    def a2a():
        e = sg.a2a_rows(recv, src, send_off, recv_off, cnt, M * 2)
        sg.wait_landed(recv, e)
    res["a2a_ms"] = timed(a2a)
    res["nccl_a2a_ms"] = timed(lambda: td.all_to_all_single(ref_a2a, src))
    res["a2a_bytes_per_rank_out"] = rows_per * (world - 1) * M * 2
    res["watchdog"] = n.hang_info("symm")
    if rank == 0:
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        with open(os.path.join(ROOT, "gpurun_out", "symm_test.json"), "w") as f:
            json.dump(res, f, indent=1)
        print(json.dumps(res))
    td.barrier()
    td.destroy_process_group()


if __name__ == "__main__":
    main()
