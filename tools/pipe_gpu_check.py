"""Pipeline parallelism on real GPUs (torchrun, 2 or 4 ranks): PP=2 over NCCL p2p, DP = world / 2 with ZeRO-`--zero`
inside each stage.  Compares the first batches' losses with a sequential fp32-master oracle on the same data and writes
gpurun_out/pipe_gpu_check.json.

    torchrun --nproc-per-node 2 tools/pipe_gpu_check.py --zero 1
    torchrun --nproc-per-node 4 tools/pipe_gpu_check.py --zero 3"""
import argparse
import copy
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn as nn


class Block(nn.Module):
    def __init__(self, d):
        super().__init__()
        self.lin = nn.Linear(d, d)

    def forward(self, x):
        return torch.tanh(self.lin(x))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--zero", type=int, default=1)
    ap.add_argument("--dim", type=int, default=256)
    ap.add_argument("--batches", type=int, default=3)
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    import shuffle_exchange_b200 as ds
    from shuffle_exchange_b200.runtime.pipe.module import LayerSpec, PipelineModule
    torch.manual_seed(0)
    d, micro, gas, dp = a.dim, 8, 4, world // 2
    ref = nn.Sequential(*[Block(d) for _ in range(4)])
    ref_state = copy.deepcopy(ref.state_dict())
    net = PipelineModule([LayerSpec(Block, d) for _ in range(4)], num_stages=2, loss_fn=nn.MSELoss(),
                         partition_method="uniform")
    own = {k: v for k, v in ref_state.items() if int(k.split(".")[0]) in range(net._local_start, net._local_stop)}
    net.load_state_dict(own, strict=True)
    net.to(dev).to(torch.bfloat16)
    engine, _, _, _ = ds.initialize(model=net, model_parameters=[p for p in net.parameters()], config={
        "train_micro_batch_size_per_gpu": micro, "gradient_accumulation_steps": gas, "steps_per_print": 10 ** 9,
        "bf16": {"enabled": True}, "optimizer": {"type": "AdamW", "params": {"lr": 1e-3, "weight_decay": 0.0}},
        "zero_optimization": {"stage": a.zero, "stage3_param_persistence_threshold": 0}})
    dp_rank = engine.grid.data_parallel_id
    g = torch.Generator().manual_seed(11)
    data = [[[(torch.randn(micro, d, generator=g), torch.randn(micro, d, generator=g)) for _ in range(gas)]
             for _ in range(dp)] for _ in range(a.batches)]
    losses = []
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for b in range(a.batches):
        it = iter([(x.to(dev, torch.bfloat16), y.to(dev, torch.bfloat16)) for x, y in data[b][dp_rank]])
        losses.append(float(engine.train_batch(data_iter=it)))
    e1.record()
    torch.cuda.synchronize()
    # oracle
    master = [p.detach().float().clone().requires_grad_() for p in ref.parameters()]
    opt = torch.optim.AdamW(master, lr=1e-3, weight_decay=0.0)
    want = []
    for b in range(a.batches):
        for p, m in zip(ref.parameters(), master):
            p.data = m.detach().to(torch.bfloat16).float()
        ref.zero_grad()
        tot = 0.0
        for r in range(dp):
            for x, y in data[b][r]:
                l = nn.functional.mse_loss(ref(x.to(torch.bfloat16).float()), y.to(torch.bfloat16).float()) / (dp * gas)
                l.backward()
                tot += float(l)
        for p, m in zip(ref.parameters(), master):
            m.grad = p.grad.clone()
        opt.step()
        want.append(tot)
    ok = all(abs(x - y) < 3e-2 * max(1.0, abs(y)) for x, y in zip(losses, want))
    if rank == world - 1 or rank == 0:
        out = {"rank": rank, "world": world, "pp": 2, "dp": dp, "zero_stage": a.zero, "losses": losses, "oracle": want,
               "ok": ok, "ms_total": e0.elapsed_time(e1), "backend": torch.distributed.get_backend()}
        print(json.dumps(out))
        if rank == 0:
            os.makedirs("gpurun_out", exist_ok=True)
            json.dump(out, open(f"gpurun_out/pipe_gpu_check_zero{a.zero}.json", "w"), indent=1)
    assert ok, (losses, want)


if __name__ == "__main__":
    main()
