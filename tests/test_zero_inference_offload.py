"""ZeRO-3 without an optimizer (``DeepSpeedZeRoOffload``: ZeRO-Inference, frozen sub-models): parameters sharded over
the group, optionally resident in host memory (``offload_param: cpu``), gathered per unit around its forward.
Reference: ``runtime/zero/parameter_offload.py:86-510``.  CPU / gloo, 2 ranks."""
import copy

import torch
import torch.nn as nn

from conftest import run_dist


class _Net(nn.Module):
    def __init__(self):
        super().__init__()
        self.layers = nn.ModuleList([nn.Linear(32, 32) for _ in range(3)])
        self.head = nn.Linear(32, 4)

    def forward(self, x):
        for l in self.layers:
            x = torch.relu(l(x))
        return self.head(x)


def _worker(rank, world, offload, nvme_dir=None):
    import shuffle_exchange_b200 as ds  # noqa: F401  (initialises the comm layer)
    from shuffle_exchange_b200 import comm as dist
    from shuffle_exchange_b200.runtime.zero.offload_config import DeepSpeedZeroOffloadParamConfig
    from shuffle_exchange_b200.runtime.zero.parameter_offload import DeepSpeedZeRoOffload
    dist.init_distributed(dist_backend="gloo")
    torch.manual_seed(0)
    model = _Net()
    for p in model.layers[1].parameters():
        p.requires_grad_(False)                               # a frozen block, as in ZeRO-Inference
    dense = copy.deepcopy(model)
    cfg = None
    if offload == "nvme":
        cfg = DeepSpeedZeroOffloadParamConfig(device="nvme", nvme_path=nvme_dir, buffer_count=2, max_in_cpu=0, pin_memory=False)
    elif offload:
        cfg = DeepSpeedZeroOffloadParamConfig(device="cpu", pin_memory=False)
    z = DeepSpeedZeRoOffload(model, param_persistence_threshold=0, offload_param_config=cfg)
    eng = z._engine
    # sharded: every unit's flat buffer is split over the two ranks and released outside its forward
    assert all(fg.lp_shard.numel() * world == fg.total for fg in eng.groups)
    if offload:
        assert all(fg.lp_shard.device.type == "cpu" for fg in eng.groups)
    if offload == "nvme":
        assert eng._param_store is not None and eng._param_store.resident_count <= 2
    assert all(p.numel() == 0 for u in eng.units if not u.persistent for fg in u.groups for p in fg.params)
    x = torch.randn(5, 32)
    with torch.no_grad():
        got = model(x)
        want = dense(x)
    torch.testing.assert_close(got, want)
    assert [p.requires_grad for p in model.layers[1].parameters()] == [False, False]
    z.partition_all_parameters()
    with torch.no_grad():
        torch.testing.assert_close(model(x), want)            # a second pass re-gathers
    assert z.get_param_coordinator() is not None and z.mark_persistent_parameters(0, 0) == z.persistent_parameters


def test_optimizerless_zero3_matches_dense():
    run_dist(_worker, 2, (False,))


def test_optimizerless_zero3_with_host_resident_shards():
    run_dist(_worker, 2, (True,))


def test_optimizerless_zero3_with_nvme_resident_shards(tmp_path):
    run_dist(_worker, 2, ("nvme", str(tmp_path)))
