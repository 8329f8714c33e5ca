"""Replica-averaging micro-benchmark (the fork's namesake), 8 GPUs: RR / shuffle / H-RR / Gossip on the peer-memory
kernel vs the NCCL / p2p path of this framework, and the cost of the REFERENCE's way of changing rings — destroying and
re-creating NCCL process groups (deepspeed/runtime/zero/stage_1_and_2.py:698-711, 713-721).

    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29565 tools/replica_avg_bench.py

Device-timed (CUDA events), max over ranks; writes gpurun_out/replica_avg_bench.json."""
import json
import os
import sys
import time

import torch
import torch.distributed as td

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def timed(fn, iters=8, warm=3):
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


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    import shuffle_exchange_b200 as sxb
    sxb.init_distributed(dist_backend="nccl")
    from shuffle_exchange_b200.runtime.zero.replica_averaging import ReplicaAverager
    numel = int(os.environ.get("RA_NUMEL", str(256 << 20)))         # bf16 elements of this rank's partition (512 MB)
    slice_count = 2 if world >= 4 else 1
    rings = 2
    res = {"world": world, "slice_count": slice_count, "replicas": world // slice_count, "partition_bytes": numel * 2, "methods": {}}
    torch.manual_seed(rank)
    base = torch.randn(numel, device="cuda").to(torch.bfloat16)
    for method in ("RR", "shuffle", "H-RR", "Gossip"):
        entry = {}
        for fused in (True, False):
            if method == "Gossip" and fused:
                continue                                    # push-sum runs on batched isend/irecv in both settings
            ra = ReplicaAverager(method, rings=rings, shuffle_step=1, slice_count=slice_count, seed=7, fused=fused)
            shard = base.clone()

            def step():
                ra.shuffle_exchange()                       # re-draws the rings every call (shuffle_step = 1)
                ra.average(shard)
            try:
                entry["kernel_ms" if fused else "nccl_ms"] = timed(step)
            except Exception as e:
                entry["kernel_error" if fused else "nccl_error"] = f"{type(e).__name__}: {e}"[:200]
            if fused and method in ("RR", "shuffle", "H-RR"):
                # numerics against the NCCL statement on fresh data
                shard = base.clone()
                ra2 = ReplicaAverager(method, rings=rings, shuffle_step=10 ** 9, slice_count=slice_count, seed=7, fused=False)
                ra3 = ReplicaAverager(method, rings=rings, shuffle_step=10 ** 9, slice_count=slice_count, seed=7, fused=True)
                a, b = base.clone(), base.clone()
                ra2.average(a)
                ra3.average(b)
                entry["kernel_vs_nccl_max_abs"] = float((a.float() - b.float()).abs().max())
        n_rep = world // slice_count
        members = n_rep if method in ("RR", "H-RR") else (n_rep // rings if method == "shuffle" else 2)
        # two-phase kernel (members share their result): 2 (m-1)/m shard sizes arrive over NVLink per rank — the
        # traffic of a bandwidth-optimal all-reduce; Gossip p2p: one shard
        entry["nvlink_bytes_in"] = int(numel * 2 * 2 * (members - 1) / members) if method != "Gossip" else numel * 2
        if "kernel_ms" in entry:
            entry["kernel_gbps_in"] = entry["nvlink_bytes_in"] / entry["kernel_ms"] / 1e6
            entry["roofline_fraction_770"] = entry["kernel_gbps_in"] / 770.0
        res["methods"][method] = entry
    # the reference's reshuffle: destroy + new_group for every ring x slice partition, then the first collective
    n_rep = world // slice_count
    groups = []
    torch.cuda.synchronize()
    td.barrier()
    t0 = time.perf_counter()
    for rnd in range(2):
        for g in groups:
            td.destroy_process_group(g)
        groups = []
        perm = torch.randperm(n_rep, generator=torch.Generator().manual_seed(rnd)).view(rings, -1).tolist()
        mine = None
        for ring in perm:
            for off in range(slice_count):
                g = td.new_group([r * slice_count + off for r in ring])
                groups.append(g)
                if rank in [r * slice_count + off for r in ring]:
                    mine = g
        x = torch.ones(1024, device="cuda")
        td.all_reduce(x, group=mine)                        # communicators are built lazily: include the first use
        torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 2
    t = torch.tensor([dt * 1e3], device="cuda")
    td.all_reduce(t, op=td.ReduceOp.MAX)
    res["reference_style_group_rebuild_ms"] = float(t)
    res["table_reshuffle_ms"] = 0.0
    from shuffle_exchange_b200.runtime.zero.replica_averaging import ShuffleSchedule
    sch = ShuffleSchedule(n_rep, rings, 7)
    t0 = time.perf_counter()
    for _ in range(100):
        sch.reshuffle()
    res["table_reshuffle_ms"] = (time.perf_counter() - t0) * 10
    if rank == 0:
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        json.dump(res, open(os.path.join(ROOT, "gpurun_out", "replica_avg_bench.json"), "w"), indent=1)
        print(json.dumps(res))
    td.barrier()
    td.destroy_process_group()


if __name__ == "__main__":
    main()
