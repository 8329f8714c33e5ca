"""Small runtime components that had no test: progressive layer drop schedule, communication bandwidth accounting,
gather/release profiler, memory-efficient linear, contiguous allocator, post-training int8 layer quantisation."""
import math
import types

import pytest
import torch


def test_progressive_layer_drop_schedule():
    from shuffle_exchange_b200.runtime.progressive_layer_drop import ProgressiveLayerDrop
    pld = ProgressiveLayerDrop(theta=0.5, gamma=0.01)
    assert pld.get_theta() == 1.0 and pld.get_state() == {"progressive_layer_drop": True, "pld_theta": 1.0}
    pld.update_state(100)
    assert abs(pld.get_theta() - (0.5 * math.exp(-1.0) + 0.5)) < 1e-12
    pld.update_state(10 ** 6)
    assert abs(pld.get_theta() - 0.5) < 1e-9                           # decays to the keep-probability floor


def test_comms_logger_bandwidth_formulas(capsys):
    from shuffle_exchange_b200.utils import comms_logging as cl
    n, size, ms = 8, 1 << 20, 2.0
    alg, bus, total = cl.calc_bw_log("all_reduce", size, ms, n)
    assert total == size and abs(alg - 2 * size / 2e-3 * 8 / 1e9) < 1e-9
    assert abs(bus - size / 2e-3 * (2 * (n - 1) / n) * 8 / 1e9) < 1e-9
    alg, bus, total = cl.calc_bw_log("all_gather_into_tensor", size, ms, n)
    assert total == size * n and abs(bus / alg - (n - 1) / n) < 1e-12
    alg, bus, _ = cl.calc_bw_log("all_to_all_single", size, ms, n)
    assert abs(bus / alg - (n - 1) / n) < 1e-12
    alg, bus, _ = cl.calc_bw_log("send", size, ms, n)
    assert alg == bus
    assert cl.convert_size(0) == "0B" and cl.convert_size(1536) == "1.5 KB" and cl.convert_size(1 << 30) == "1.0 GB"
    lg = cl.CommsLogger()
    lg.configure(types.SimpleNamespace(enabled=True, verbose=False, debug=False, prof_ops=["all_reduce"], prof_all=False))
    lg.append("all_reduce", "all_reduce", 2.0, size, n)
    lg.append("all_reduce", "all_reduce", 4.0, size, n)
    lg.append("broadcast", "broadcast", 1.0, 4096, n)
    rows = lg.log_all(print_log=True)
    ar = [r for r in rows if r[0] == "all_reduce"][0]
    assert ar[1] == size and ar[2] == 2 and ar[3] == 6.0 and ar[4] == 3.0
    assert "all_reduce" in capsys.readouterr().out
    lg.start_profiling_op(["all_gather"])
    assert "all_gather" in lg.prof_ops
    lg.stop_profiling_op(["all_gather", "all_reduce"])
    assert lg.prof_ops == []


def test_partitioned_parameter_profiler_counts_events():
    from shuffle_exchange_b200.runtime.zero.partitioned_param_profiler import PartitionedParameterProfiler
    from shuffle_exchange_b200.utils.timer import SynchronizedWallClockTimer
    off = PartitionedParameterProfiler(None)
    off.start_event("x")
    off.stop_event("x", 5)
    assert not off.enabled and off.event_counters == {}
    prof = PartitionedParameterProfiler(SynchronizedWallClockTimer())
    for numel in (10, 20):
        prof.start_event("fetch")
        prof.stop_event("fetch", numel)
    t = prof.event_counters["fetch"]
    assert (t.count, t.num_elem) == (2, 30)
    with pytest.raises(KeyError):
        prof.stop_event("never_started", 1)
    prof.log_events()
    prof.reset_events()
    assert (t.count, t.num_elem) == (0, 0)


def test_memory_efficient_linear_matches_torch_linear():
    from shuffle_exchange_b200.runtime.zero.linear import LinearModuleForZeroStage3, zero3_linear_wrap
    torch.manual_seed(0)
    m = LinearModuleForZeroStage3(12, 7, bias=True)
    ref = torch.nn.Linear(12, 7)
    ref.load_state_dict(m.state_dict())
    x = torch.randn(3, 5, 12, requires_grad=True)
    xr = x.detach().clone().requires_grad_()
    gy = torch.randn(3, 5, 7)
    m(x).backward(gy)
    ref(xr).backward(gy)
    assert torch.allclose(x.grad, xr.grad, atol=1e-5) and torch.allclose(m.weight.grad, ref.weight.grad, atol=1e-5)
    assert torch.allclose(m.bias.grad, ref.bias.grad, atol=1e-5)
    y = zero3_linear_wrap(x.detach(), m.weight)                         # no bias
    assert torch.allclose(y, torch.nn.functional.linear(x.detach(), m.weight), atol=1e-6)


def test_contiguous_allocator_defragments_and_keeps_contents():
    from shuffle_exchange_b200.runtime.zero.contiguous_memory_allocator import ContiguousMemoryAllocator
    mem = ContiguousMemoryAllocator(64, torch.float32, "cpu")
    a, b, c = mem.allocate_tensor(16), mem.allocate_tensor(16), mem.allocate_tensor(16)
    for t, v in ((a, 1.0), (b, 2.0), (c, 3.0)):
        t.fill_(v)
    p = torch.nn.Parameter(torch.empty(0))
    mem.assign_to_param(c, p, 16, (4, 4))
    mem.release_tensor(b)                                               # leaves a 16-element hole + 16 at the end
    assert mem.total_free == 32 and mem.largest_contiguous == 16
    d = mem.allocate_tensor(32)                                         # only fits after compaction
    d.fill_(4.0)
    assert mem.total_free == 0
    assert torch.all(a == 1.0) and torch.all(p.data == 3.0) and p.shape == (4, 4)   # survivors moved with their data
    with pytest.raises(AssertionError):
        mem.allocate_tensor(1)
    mem.release_tensor(a)
    assert mem.total_free == 16


def test_int8_layer_quantisation_in_place():
    from shuffle_exchange_b200.module_inject.module_quantize import quantize_transformer_layer

    class Block(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.q = torch.nn.Linear(16, 16)
            self.fc = torch.nn.Linear(16, 32)

        def forward(self, x):
            return self.fc(self.q(x))

    torch.manual_seed(0)
    model = torch.nn.Sequential(Block(), torch.nn.Linear(32, 4))
    x = torch.randn(5, 16)
    want = model(x)
    quantize_transformer_layer(Block, model, groups=4)
    blk = model[0]
    assert blk.q.weight.dtype == torch.int8 and not blk.q.weight.requires_grad and blk.q.weight_scale.numel() == 4
    assert model[1].weight.dtype == torch.float32                       # other layers untouched
    got = model(x)
    assert (got - want).abs().max() < 0.05 * want.abs().max()           # int8, 4 groups: ~1 % weight error
    w = blk.fc.dequantized_weight()
    assert w.shape == (32, 16) and (w.reshape(4, -1).abs().amax(1) > 0).all()
