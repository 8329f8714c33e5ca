"""ZeRO-3 configuration switches that used to be accepted and ignored: every one of them must now change what runs and
still train like an unsharded AdamW — qwZ (int8 weight gather), hpZ (secondary partition for the backward gathers),
qgZ with and without LoCo (int4 gradient all-to-all), partial optimizer offload (``ratio``), ``offload_param: cpu``
and NVMe-swapped optimizer state (ZeRO-Infinity).  CPU / gloo."""
import copy

import pytest
import torch
import torch.nn as nn

from conftest import run_dist


class _Net(nn.Module):
    def __init__(self):
        super().__init__()
        self.layers = nn.ModuleList([nn.Linear(64, 64) for _ in range(3)])
        self.head = nn.Linear(64, 8)

    def forward(self, x):
        for l in self.layers:
            x = torch.tanh(l(x))
        return (self.head(x).float() ** 2).mean()


def _worker(rank, world, extra, tmp):
    import shuffle_exchange_b200 as ds
    from shuffle_exchange_b200.utils.tensor_fragment import safe_get_full_fp32_param
    torch.manual_seed(0)
    model = _Net().to(torch.bfloat16)
    oracle = copy.deepcopy(model).float()
    start = [p.detach().clone() for p in oracle.parameters()]
    z = {"stage": 3, "stage3_param_persistence_threshold": 0}
    z.update(copy.deepcopy(extra))
    if z.get("offload_optimizer", {}).get("device") == "nvme":
        z["offload_optimizer"]["nvme_path"] = tmp
    if z.get("offload_param", {}).get("device") == "nvme":
        z["offload_param"]["nvme_path"] = tmp
    cfg = {"train_micro_batch_size_per_gpu": 4, "steps_per_print": 10 ** 9, "bf16": {"enabled": True},
           "optimizer": {"type": "AdamW", "params": {"lr": 1e-2, "weight_decay": 0.0}}, "zero_optimization": z}
    engine, opt, _, _ = ds.initialize(model=model, config=cfg, model_parameters=model.parameters())
    # the switch reached the optimizer
    if extra.get("zero_quantized_weights"):
        assert all(fg.quantized_weights for fg in opt.groups)
    if extra.get("zero_hpz_partition_size", 1) > 1:
        assert all(fg.hpz_group is not None for fg in opt.groups)
    if extra.get("zero_quantized_gradients"):
        assert all(fg.quantized_gradients is not None for fg in opt.groups)
    if "offload_param" in extra:
        assert all(fg.lp_shard.device.type == "cpu" for fg in opt.groups)
    if extra.get("offload_param", {}).get("device") == "nvme":
        import os
        store = opt._param_store
        assert store is not None and all(os.path.exists(store.path(fg._lp_key)) for fg in opt.groups)
        assert all(fg._lp_shard is None for fg in opt.groups), "shards must live in the store, not in the group"
    if extra.get("offload_optimizer", {}).get("ratio", 1.0) < 1.0:
        on_cpu = [fg.hp_on_cpu for fg in opt.groups]
        assert any(on_cpu) and not all(on_cpu), on_cpu
    g = torch.Generator().manual_seed(9)
    o = torch.optim.AdamW(oracle.parameters(), lr=1e-2, weight_decay=0.0)
    for _ in range(3):
        xs = [torch.randn(4, 64, generator=g) for _ in range(world)]
        loss = engine(xs[rank].to(torch.bfloat16))
        engine.backward(loss)
        engine.step()
        o.zero_grad()
        for x in xs:
            (oracle(x.to(torch.bfloat16).float()) / world).backward()
        o.step()
    if extra.get("offload_param", {}).get("device") == "nvme":
        store = opt._param_store
        assert store.resident_count <= 2, store.resident_count         # buffer_count of the config
        assert store.reads > len(opt.groups) and store.writes > len(opt.groups)   # shards really went to disk and back
    if extra.get("offload_optimizer", {}).get("device") == "nvme":
        assert opt.optimizer_swapper is not None and len(opt._swapped) == len(opt.groups)
        assert all(fg.hp_param.data.numel() == 0 for fg in opt.groups), "masters must be on NVMe between steps"
        opt.swap_in_all()
    num = da = db = 0.0
    for (n, p), q, q0 in zip(engine.module.named_parameters(), oracle.parameters(), start):
        got = safe_get_full_fp32_param(p).cpu()
        a, b = (got - q0).double().flatten(), (q.detach() - q0).double().flatten()
        num += float(a @ b)
        da += float(a @ a)
        db += float(b @ b)
    cos = num / (da ** 0.5 * db ** 0.5)
    assert cos > 0.9, (extra, cos)


CASES = [
    (2, {"zero_quantized_weights": True}),
    (4, {"zero_hpz_partition_size": 2}),
    (4, {"zero_quantized_gradients": True}),
    (4, {"zero_quantized_gradients": True, "zeropp_loco_param": {"err_beta": 0.8, "reset_T": 1024}}),
    (2, {"offload_optimizer": {"device": "cpu", "ratio": 0.5}}),
    (2, {"offload_param": {"device": "cpu"}}),
    (2, {"offload_optimizer": {"device": "nvme", "buffer_count": 16}}),
    (2, {"offload_param": {"device": "nvme", "buffer_count": 2, "max_in_cpu": 0}}),   # nothing resident beyond 2 shards
]


@pytest.mark.parametrize("world,extra", CASES, ids=[",".join(c[1].keys()) + ("+loco" if "zeropp_loco_param" in c[1] else "")
                                                    + str(list(c[1].values())[0].get("device", "") if isinstance(list(c[1].values())[0], dict) else "")
                                                    for c in CASES])
def test_zero3_switch_acts_and_trains(world, extra, tmp_path):
    run_dist(_worker, world, (extra, str(tmp_path)))


def test_unsupported_switches_raise():
    from shuffle_exchange_b200.runtime.zero.stage3 import DeepSpeedZeroOptimizer_Stage3

    class _Cfg:
        device = "nvme"
        nvme_path = None
        pin_memory = False

    m = _Net().to(torch.bfloat16)
    base = torch.optim.AdamW(m.parameters(), lr=1e-3)
    with pytest.raises(ValueError):                       # NVMe parameter offload without a path
        DeepSpeedZeroOptimizer_Stage3(m, base, offload_param_config=_Cfg())
    with pytest.raises(NotImplementedError):
        DeepSpeedZeroOptimizer_Stage3(m, base, zero_quantized_nontrainable_weights=True)


# ------------------------------------------------------------------ gradient buffers that are not zero-filled
class _PartlyUsed(nn.Module):
    """One layer's bias is frozen and one whole linear is never used in forward: their ranges of the flat gradient
    buffer are never written by a gradient producer."""

    def __init__(self):
        super().__init__()
        self.layers = nn.ModuleList([nn.Linear(64, 64) for _ in range(2)])
        self.unused = nn.ModuleList([nn.Linear(64, 64)])
        self.layers[0].bias.requires_grad_(False)

    def forward(self, x):
        for l in self.layers:
            x = torch.tanh(l(x))
        return (x.float() ** 2).mean()


def _lazy_zero_worker(rank, world):
    import shuffle_exchange_b200 as ds
    from shuffle_exchange_b200.utils.tensor_fragment import safe_get_full_fp32_param
    torch.manual_seed(0)
    model = _PartlyUsed().to(torch.bfloat16)
    before = {n: p.detach().float().clone() for n, p in model.named_parameters()}
    cfg = {"train_micro_batch_size_per_gpu": 4, "steps_per_print": 10 ** 9, "bf16": {"enabled": True},
           "optimizer": {"type": "AdamW", "params": {"lr": 1e-2, "weight_decay": 0.0}},
           "zero_optimization": {"stage": 3, "stage3_param_persistence_threshold": 0}}
    engine, opt, _, _ = ds.initialize(model=model, config=cfg, model_parameters=model.parameters())
    assert all(fg.lazy_grad_zero for fg in opt.groups)
    # poison what the allocator hands out next: a buffer that is read without having been written shows up as NaN
    for fg in opt.groups:
        assert fg.dead_ranges() is not None
    real_empty = torch.empty

    def poisoned(*a, **k):
        t = real_empty(*a, **k)
        if t.is_floating_point() and t.numel():
            t.fill_(float("nan"))
        return t
    torch.empty = poisoned
    try:
        for _ in range(2):
            loss = engine(torch.randn(4, 64).to(torch.bfloat16))
            engine.backward(loss)
            engine.step()
    finally:
        torch.empty = real_empty
    after = {n: safe_get_full_fp32_param(p) for n, p in model.named_parameters()}
    for n, a in after.items():
        assert torch.isfinite(a).all(), n
        moved = not torch.equal(a.cpu().float(), before[n])
        if n.startswith("unused") or n == "layers.0.bias":
            assert not moved, f"{n} has no gradient and no weight decay: it must not move"
        else:
            assert moved, n


def test_unwritten_gradient_ranges_read_as_zero():
    run_dist(_lazy_zero_worker, 2)
