"""Pipeline parallelism on gloo: topology math, 1F1B schedule invariants, 2-stage training parity."""
import copy

import pytest
import torch
import torch.nn as nn

from conftest import run_dist
from shuffle_exchange_b200.runtime.pipe import schedule as S
from shuffle_exchange_b200.runtime.pipe.topology import PipeModelDataParallelTopology, ProcessTopology


def test_topology_rank_coord_roundtrip():
    t = ProcessTopology(axes=["pipe", "data", "model"], dims=[2, 2, 2])
    assert t.world_size() == 8
    for r in range(8):
        c = t.get_coord(r)
        assert t.get_rank(pipe=c.pipe, data=c.data, model=c.model) == r
    assert t.get_axis_comm_lists("pipe") == [[0, 4], [1, 5], [2, 6], [3, 7]]
    assert t.get_axis_comm_lists("model") == [[0, 1], [2, 3], [4, 5], [6, 7]]
    assert t.filter_match(pipe=1, model=0) == [4, 6]
    assert PipeModelDataParallelTopology(2, 2, 2).get_rank_repr(5) == "model_01"


def test_train_schedule_is_consistent():
    """Every send has a matching receive in the same step; each micro-batch does one fwd and one bwd per stage."""
    M, stages = 5, 3
    per_stage = [list(S.TrainSchedule(M, stages, s).steps()) for s in range(stages)]
    assert all(len(p) == 2 * (M + stages - 1) for p in per_stage)
    for step in range(len(per_stage[0])):
        for s in range(stages - 1):
            sends = sum(isinstance(c, S.SendActivation) for c in per_stage[s][step])
            recvs = sum(isinstance(c, S.RecvActivation) for c in per_stage[s + 1][step])
            assert sends == recvs
            gs = sum(isinstance(c, S.SendGrad) for c in per_stage[s + 1][step])
            gr = sum(isinstance(c, S.RecvGrad) for c in per_stage[s][step])
            assert gs == gr
    for s in range(stages):
        flat = [c for st in per_stage[s] for c in st]
        assert sum(isinstance(c, S.ForwardPass) for c in flat) == M
        assert sum(isinstance(c, S.BackwardPass) for c in flat) == M
        assert isinstance(flat[-1], S.OptimizerStep)
        assert S.TrainSchedule(M, stages, s).num_pipe_buffers() == max(2, min(stages - s, M))


class _Block(nn.Module):
    def __init__(self, d):
        super().__init__()
        self.lin = nn.Linear(d, d)

    def forward(self, x):
        return torch.tanh(self.lin(x))


def _pipe_worker(rank, world, pipeline_cfg=None):
    import shuffle_exchange_b200 as ds
    from shuffle_exchange_b200.runtime.pipe.module import LayerSpec, PipelineModule
    torch.manual_seed(0)
    d, micro, gas = 8, 4, 2
    ref = nn.Sequential(*[_Block(d) for _ in range(4)])
    ref_state = copy.deepcopy(ref.state_dict())
    data = [(torch.randn(micro, d), torch.randn(micro, d)) for _ in range(gas * 3)]
    torch.manual_seed(0)
    net = PipelineModule([LayerSpec(_Block, d) for _ in range(4)], num_stages=2, loss_fn=nn.MSELoss(),
                         partition_method="uniform")
    # load identical weights: stage s owns layers [2s, 2s+2)
    own = {k: v for k, v in ref_state.items() if int(k.split(".")[0]) in range(net._local_start, net._local_stop)}
    net.load_state_dict(own, strict=True)
    engine, _, _, _ = ds.initialize(model=net, model_parameters=[p for p in net.parameters()], config={
        "train_micro_batch_size_per_gpu": micro, "gradient_accumulation_steps": gas, "steps_per_print": 10 ** 9,
        "optimizer": {"type": "SGD", "params": {"lr": 0.1}}, **({"pipeline": pipeline_cfg} if pipeline_cfg else {})})
    if pipeline_cfg:
        from shuffle_exchange_b200.runtime.activation_checkpointing import checkpointing as ck
        assert net.activation_checkpoint_interval == pipeline_cfg["activation_checkpoint_interval"]
        want_fn = ck.checkpoint if pipeline_cfg.get("use_reentrant", True) else ck.non_reentrant_checkpoint
        assert net.activation_checkpoint_func is want_fn
        calls = []
        net.activation_checkpoint_func = lambda f, *a: (calls.append(1), want_fn(f, *a))[1]
    it = iter(data)
    losses = [float(engine.train_batch(data_iter=it)) for _ in range(3)]
    if pipeline_cfg:
        assert len(calls) >= 3 * gas, "activation checkpointing was configured but never used"
    # reference: plain SGD with gradient accumulation
    opt = torch.optim.SGD(ref.parameters(), lr=0.1)
    it = iter(data)
    want = []
    for _ in range(3):
        opt.zero_grad()
        tot = 0.0
        for _ in range(gas):
            x, y = next(it)
            l = nn.functional.mse_loss(ref(x), y) / gas
            l.backward()
            tot += float(l)
        opt.step()
        want.append(tot)
    assert all(abs(a - b) < 1e-5 for a, b in zip(losses, want)), (losses, want)
    for k, v in net.state_dict().items():
        assert torch.allclose(v, ref.state_dict()[k], atol=1e-5), k


def test_two_stage_training_matches_sequential():
    run_dist(_pipe_worker, 2)


@pytest.mark.parametrize("use_reentrant", [True, False])
def test_two_stage_training_with_activation_checkpointing(use_reentrant):
    """``pipeline.activation_checkpoint_interval`` recomputes every layer in backward (re-entrant or hook-based per
    ``pipeline.use_reentrant``); training must be bit-for-bit the run without recomputation."""
    run_dist(_pipe_worker, 2, ({"activation_checkpoint_interval": 1, "use_reentrant": use_reentrant},))


def _pipe_zero_worker(rank, world, zero_stage, offload=False):
    """PP=2 x DP=2 with the stage's data-parallel pair running ZeRO-``zero_stage`` (the reference asserts stage < 2 under
    pipeline parallelism, pipe/engine.py:77-78): two batches must match unsharded AdamW on the same data."""
    import shuffle_exchange_b200 as ds
    from shuffle_exchange_b200.runtime.pipe.module import LayerSpec, PipelineModule
    torch.manual_seed(0)
    d, micro, gas = 16, 4, 2
    ref = nn.Sequential(*[_Block(d) for _ in range(4)])
    ref_state = copy.deepcopy(ref.state_dict())
    net = PipelineModule([LayerSpec(_Block, d) for _ in range(4)], num_stages=2, loss_fn=nn.MSELoss(),
                         partition_method="uniform")
    own = {k: v for k, v in ref_state.items() if int(k.split(".")[0]) in range(net._local_start, net._local_stop)}
    net.load_state_dict(own, strict=True)
    engine, _, _, _ = ds.initialize(model=net, model_parameters=[p for p in net.parameters()], config={
        "train_micro_batch_size_per_gpu": micro, "gradient_accumulation_steps": gas, "steps_per_print": 10 ** 9,
        "bf16": {"enabled": True}, "optimizer": {"type": "AdamW", "params": {"lr": 1e-2, "weight_decay": 0.0}},
        "zero_optimization": dict({"stage": zero_stage, "stage3_param_persistence_threshold": 0},
                                  **({"offload_optimizer": {"device": "cpu"}, "offload_param": {"device": "cpu"}}
                                     if offload else {}))})
    if offload:
        assert all(fg.hp_on_cpu and fg.lp_shard.device.type == "cpu" for fg in engine.optimizer.groups)
    assert engine.zero_optimization_stage() == zero_stage
    dp_rank = engine.grid.data_parallel_id
    g = torch.Generator().manual_seed(11)
    # data[batch][dp][micro] ; both stages of a replica draw the same stream (inputs enter at stage 0, labels at stage 1)
    data = [[[(torch.randn(micro, d, generator=g), torch.randn(micro, d, generator=g)) for _ in range(gas)]
             for _ in range(2)] for _ in range(2)]
    losses = []
    for b in range(2):
        it = iter([(x.to(torch.bfloat16), y.to(torch.bfloat16)) for x, y in data[b][dp_rank]])
        losses.append(float(engine.train_batch(data_iter=it)))
    # oracle: fp32 master, bf16 compute weights, mean gradient over 2 replicas x gas micro-batches
    master = [p.detach().float().clone().requires_grad_() for p in ref.parameters()]
    opt = torch.optim.AdamW(master, lr=1e-2, weight_decay=0.0)
    want = []
    for b in range(2):
        for p, m in zip(ref.parameters(), master):
            p.data = m.detach().to(torch.bfloat16).float()
        ref.zero_grad()
        tot = 0.0
        for dp in range(2):
            for x, y in data[b][dp]:
                l = nn.functional.mse_loss(ref(x.to(torch.bfloat16).float()), y.to(torch.bfloat16).float()) / (2 * gas)
                l.backward()
                tot += float(l)
        for p, m in zip(ref.parameters(), master):
            m.grad = p.grad.clone()
        opt.step()
        want.append(tot)
    assert all(abs(a - b) < 3e-2 * max(1.0, abs(b)) for a, b in zip(losses, want)), (zero_stage, losses, want)
    assert losses[1] < losses[0] + 1e-3 or want[1] >= want[0], (losses, want)


@pytest.mark.parametrize("zero_stage", [2, 3])
def test_zero_sharding_inside_pipeline_stages(zero_stage):
    run_dist(_pipe_zero_worker, 4, (zero_stage,))


def test_zero3_with_host_offload_inside_pipeline_stages():
    """BASELINE config 5 in miniature: ZeRO-3 + pipeline (PP=2) with optimizer state and parameter shards in host
    memory, DP=2 per stage."""
    run_dist(_pipe_zero_worker, 4, (3, True))


class _Embed(nn.Module):
    def __init__(self, vocab, d):
        super().__init__()
        self.weight = nn.Parameter(torch.randn(vocab, d) * 0.1)

    def forward(self, ids):
        return nn.functional.embedding(ids, self.weight)


def _unembed(module, x):
    return x @ module.weight.t()


def _ce(logits, labels):
    return nn.functional.cross_entropy(logits.reshape(-1, logits.shape[-1]), labels.reshape(-1))


def _tied_pipe_worker(rank, world, root):
    """Embedding on the first stage tied to the LM head on the last: both copies start equal, their gradients are
    summed every step, training equals a single-process model that really shares the tensor; eval_batch returns the
    same loss everywhere; a checkpoint written by the two stages restores a fresh engine bit for bit."""
    import shuffle_exchange_b200 as ds
    from shuffle_exchange_b200.runtime.pipe.module import LayerSpec, PipelineModule, TiedLayerSpec
    V, d, micro, gas = 32, 8, 4, 2

    def build():
        torch.manual_seed(0)
        return PipelineModule([TiedLayerSpec("embed", _Embed, V, d), LayerSpec(_Block, d), LayerSpec(_Block, d),
                               TiedLayerSpec("embed", _Embed, V, d, forward_fn=_unembed)],
                              num_stages=2, loss_fn=_ce, partition_method="uniform")

    cfg = {"train_micro_batch_size_per_gpu": micro, "gradient_accumulation_steps": gas, "steps_per_print": 10 ** 9,
           "optimizer": {"type": "SGD", "params": {"lr": 0.5}}}
    # single-process oracle with ONE shared embedding
    torch.manual_seed(0)
    emb = _Embed(V, d)
    blocks = nn.Sequential(_Block(d), _Block(d))
    net = build()
    tied = net.tied_modules["embed"] if hasattr(net, "tied_modules") else None
    assert tied is not None
    with torch.no_grad():                            # identical starting point on both sides
        tied.weight.copy_(emb.weight)
        for i, blk in enumerate(blocks):
            idx = i + 1
            if net._local_start <= idx < net._local_stop:
                net.forward_funcs[idx - net._local_start].load_state_dict(blk.state_dict())
    engine, _, _, _ = ds.initialize(model=net, model_parameters=[p for p in net.parameters()], config=cfg)
    g = torch.Generator().manual_seed(3)
    data = [(torch.randint(0, V, (micro, 6), generator=g),) * 2 for _ in range(gas * 4)]
    it = iter(data)
    losses = [float(engine.train_batch(data_iter=it)) for _ in range(3)]
    opt = torch.optim.SGD(list(emb.parameters()) + list(blocks.parameters()), lr=0.5)
    it2 = iter(data)
    want = []
    for _ in range(3):
        opt.zero_grad()
        tot = 0.0
        for _ in range(gas):
            ids, labels = next(it2)
            l = _ce(blocks(emb(ids)) @ emb.weight.t(), labels) / gas
            l.backward()
            tot += float(l)
        opt.step()
        want.append(tot)
    assert all(abs(a - b) < 1e-5 for a, b in zip(losses, want)), (losses, want)
    assert torch.allclose(tied.weight, emb.weight, atol=1e-5)                 # both stages hold the SAME updated tensor
    # evaluation: forward only, loss known on every stage
    ev = float(engine.eval_batch(iter(data[-gas:])))
    with torch.no_grad():
        want_ev = sum(float(_ce(blocks(emb(i)) @ emb.weight.t(), l)) for i, l in data[-gas:]) / gas
    assert abs(ev - want_ev) < 1e-5, (ev, want_ev)
    assert engine.is_first_stage() == (rank == 0) and engine.is_last_stage() == (rank == world - 1)
    # checkpoint: per-layer files from both stages, restored into a fresh engine that continues identically
    engine.save_checkpoint(root, tag="t")
    import os
    files = os.listdir(os.path.join(root, "t"))
    assert sum(f.startswith("layer_") for f in files) == 4, files
    net2 = build()
    e2, _, _, _ = ds.initialize(model=net2, model_parameters=[p for p in net2.parameters()], config=cfg)
    e2.load_checkpoint(root, tag="t")
    assert e2.global_steps == 3
    nxt = iter(data[-gas:])
    nxt2 = iter(data[-gas:])
    assert abs(float(engine.train_batch(data_iter=nxt)) - float(e2.train_batch(data_iter=nxt2))) < 1e-6
    for (n, p), (_, q) in zip(net.named_parameters(), net2.named_parameters()):
        assert torch.allclose(p, q, atol=1e-6), n


def test_tied_embedding_eval_and_checkpoint_across_two_stages(tmp_path):
    run_dist(_tied_pipe_worker, 2, (str(tmp_path),))
