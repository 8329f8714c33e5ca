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


def test_random_ltd_end_to_end_on_cpu():
    """Random layer-token drop: convert layers, schedule the kept length, run a step where the wrapped layers see only
    the kept tokens (others pass through unchanged), gradients flow to both; checkpoints load into the plain model."""
    from shuffle_exchange_b200.runtime.data_pipeline.constants import (
        RANDOM_LTD_GLOBAL_BATCH_SIZE, RANDOM_LTD_HIDDEN_STATE_ORDER, RANDOM_LTD_INCREASE_STEP, RANDOM_LTD_LAYER_NUM,
        RANDOM_LTD_MAX_VALUE, RANDOM_LTD_MICRO_BATCH_SIZE, RANDOM_LTD_MIN_VALUE, RANDOM_LTD_MODEL_MASK_NAME,
        RANDOM_LTD_MODEL_TYPE, RANDOM_LTD_REQUIRE_STEP, RANDOM_LTD_SCHEDULE_CONFIG, RANDOM_LTD_SCHEDULER,
        RANDOM_LTD_SCHEDULER_TYPE, RANDOM_LTD_TOTAL_LAYER_NUM)
    from shuffle_exchange_b200.runtime.data_pipeline.data_routing.basic_layer import RandomLayerTokenDrop
    from shuffle_exchange_b200.runtime.data_pipeline.data_routing.helper import (convert_to_random_ltd,
                                                                               remove_random_ltd_state_dict)
    from shuffle_exchange_b200.runtime.data_pipeline.data_routing.scheduler import RandomLTDScheduler

    class Layer(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.fc = torch.nn.Linear(8, 8)
            self.seen = None

        def forward(self, x, attention_mask=None):
            self.seen = x.shape
            return x + self.fc(x)

    class Net(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.layers = torch.nn.ModuleList([Layer() for _ in range(3)])
            self.head = torch.nn.Linear(8, 1)

        def forward(self, x):
            for l in self.layers:
                x = l(x)
            return self.head(x).sum()

    torch.manual_seed(0)
    net = Net()
    plain_keys = set(net.state_dict())
    convert_to_random_ltd(net, Layer)
    wrapped = [m for m in net.modules() if isinstance(m, RandomLayerTokenDrop)]
    assert len(wrapped) == 3 and set(remove_random_ltd_state_dict(net.state_dict())) == plain_keys
    cfg = {RANDOM_LTD_TOTAL_LAYER_NUM: 3, RANDOM_LTD_LAYER_NUM: 3, RANDOM_LTD_GLOBAL_BATCH_SIZE: 2,
           RANDOM_LTD_MICRO_BATCH_SIZE: 2, RANDOM_LTD_MODEL_MASK_NAME: None, RANDOM_LTD_MODEL_TYPE: "decoder",
           RANDOM_LTD_HIDDEN_STATE_ORDER: "batch_seq_dim",
           RANDOM_LTD_SCHEDULER: {RANDOM_LTD_MIN_VALUE: 4, RANDOM_LTD_MAX_VALUE: 16, RANDOM_LTD_SCHEDULER_TYPE: "fixed_linear",
                                  RANDOM_LTD_SCHEDULE_CONFIG: {RANDOM_LTD_REQUIRE_STEP: 10, RANDOM_LTD_INCREASE_STEP: 4}}}
    sched = RandomLTDScheduler(cfg)
    for i, w in enumerate(wrapped):
        w.init_config(cfg, sched, i)
    assert sched.get_current_seq() == 4
    sched.update_seq(5)                                   # halfway: 4 + 0.5 * 12 = 10 -> multiple of 4 -> 8
    assert sched.get_current_seq() == 8
    assert sched.state_dict()["consumed_layer_tokens"] == 2 * 8 * 3
    net.train()
    x = torch.randn(2, 16, 8, requires_grad=True)
    net(x).backward()
    assert all(tuple(w.random_ltd_layer.seen) == (2, 8, 8) for w in wrapped)      # layers ran on the 8 kept tokens
    assert x.grad is not None and x.grad.abs().sum() > 0
    assert all(w.random_ltd_layer.fc.weight.grad is not None for w in wrapped)
    sched.update_seq(10)
    assert sched.get_current_seq() == 16
    net(torch.randn(2, 16, 8))
    assert tuple(wrapped[0].random_ltd_layer.seen) == (2, 16, 8)                    # schedule complete: nothing dropped
    net.eval()
    sched.set_current_seq(4)
    net(torch.randn(2, 16, 8))
    assert tuple(wrapped[0].random_ltd_layer.seen) == (2, 16, 8)                    # never drops in eval
    s2 = RandomLTDScheduler(cfg)
    s2.load_state_dict(sched.state_dict())
    assert s2.get_current_seq() == 4 and s2.state_dict() == sched.state_dict()


def test_zero_and_checkpoint_helpers():
    import collections
    from shuffle_exchange_b200.checkpoint.utils import (clone_tensors_for_torch_save, get_layer_ckpt_name_for_rank,
                                                        get_model_ckpt_name_for_rank, get_zero_ckpt_name_for_rank)
    from shuffle_exchange_b200.io.utils import bytes_to_tensor, obj_serialization_details, serialize_details, tensor_to_bytes
    from shuffle_exchange_b200.runtime.zero import utils as zu
    from shuffle_exchange_b200.runtime.zero.mics_utils import _generate_mics_config
    P = collections.namedtuple("P", "a b")
    nested = {"x": [torch.ones(2), (torch.zeros(1), 3)], "y": P(torch.ones(1), "s")}
    doubled = zu.apply_to_tensors_only(lambda t: t * 2, nested)
    assert doubled["x"][0].tolist() == [2.0, 2.0] and doubled["x"][1][1] == 3
    assert isinstance(doubled["y"], P) and doubled["y"].a.item() == 2.0 and doubled["y"].b == "s"
    ts = [torch.zeros(3), torch.zeros(2, 2)]
    assert [(o, n) for _, o, n in zu.get_mapping_to_flat_buffer(ts)] == [(0, 3), (3, 4)]
    assert zu.is_zero_supported_optimizer(torch.optim.AdamW([torch.nn.Parameter(torch.zeros(1))]))
    assert not zu.is_zero_supported_optimizer(torch.optim.LBFGS([torch.nn.Parameter(torch.zeros(1))]))

    cfg = _generate_mics_config(world_size=16, ndev_per_node=4, shard_size=8)
    assert cfg["shard_groups"] == [list(range(8)), list(range(8, 16))]
    assert cfg["replicate_groups"] == [[i, i + 8] for i in range(8)] and cfg["span_nodes"] == 2
    with pytest.raises(AssertionError):
        _generate_mics_config(world_size=12, ndev_per_node=4, shard_size=8)

    flat = torch.arange(1000.0)
    view = {"w": flat[10:20], "l": [flat[:3]], "n": 5}
    cloned = clone_tensors_for_torch_save(view)
    assert cloned["w"].untyped_storage().nbytes() == 40 and torch.equal(cloned["w"], flat[10:20]) and cloned["n"] == 5
    assert get_model_ckpt_name_for_rank("/c", "03").endswith("mp_rank_03_model_states.pt")
    assert get_zero_ckpt_name_for_rank("/c", 2, 1).endswith("zero_pp_rank_2_mp_rank_01_optim_states.pt")
    # (sic) the reference's helper joins with "_" although PipelineModule writes "layer_XX-model_YY-model_states.pt"
    assert get_layer_ckpt_name_for_rank("/c", "layer_04", 1).endswith("layer_04-model_01_model_states.pt")
    t = torch.randn(7).to(torch.bfloat16)
    assert torch.equal(bytes_to_tensor(tensor_to_bytes(t)).view(torch.bfloat16), t)
    assert obj_serialization_details({"a": [torch.zeros(4), torch.zeros(2, dtype=torch.float64)], "b": 1}) == (2, 32)
    assert serialize_details is obj_serialization_details


def _ints_worker(rank, world):
    from shuffle_exchange_b200.runtime.zero.utils import assert_ints_same_as_other_ranks, get_lst_from_rank0
    assert get_lst_from_rank0([rank + 5, 9]) == [5, 9]
    assert_ints_same_as_other_ranks([1, 2, 3])
    try:
        assert_ints_same_as_other_ranks([1, 2, 3 + rank])
        raised = False
    except RuntimeError:
        raised = True
    assert raised == (rank != 0)


def test_ranks_disagreeing_on_a_fetch_order_are_caught():
    from conftest import run_dist
    run_dist(_ints_worker, 2)


def _curriculum_pld_worker(rank, world):
    """Legacy sequence-length curriculum and progressive layer drop reach the model's forward as keyword arguments, step
    by step, and only while training."""
    import math
    import shuffle_exchange_b200 as ds

    class Net(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.emb = torch.nn.Embedding(32, 8)
            self.head = torch.nn.Linear(8, 32)
            self.calls = []

        def forward(self, ids, curriculum_seqlen=None, progressive_layer_drop=None, pld_theta=None):
            self.calls.append((curriculum_seqlen, progressive_layer_drop, pld_theta))
            if curriculum_seqlen is not None:
                ids = ids[:, :curriculum_seqlen]
            logits = self.head(self.emb(ids))
            return torch.nn.functional.cross_entropy(logits.reshape(-1, 32), ids.reshape(-1))

    net = Net()
    engine, _, _, _ = ds.initialize(model=net, model_parameters=net.parameters(), config={
        "train_micro_batch_size_per_gpu": 2, "steps_per_print": 10 ** 9,
        "optimizer": {"type": "SGD", "params": {"lr": 0.1}},
        "progressive_layer_drop": {"enabled": True, "theta": 0.5, "gamma": 0.1},
        "curriculum_learning": {"enabled": True, "curriculum_type": "seqlen", "min_difficulty": 4, "max_difficulty": 16,
                                "schedule_type": "fixed_linear",
                                "schedule_config": {"total_curriculum_step": 4, "difficulty_step": 4}}})
    ids = torch.randint(0, 32, (2, 16))
    for _ in range(6):
        engine.backward(engine(ids))
        engine.step()
    seqlens = [c[0] for c in net.calls]
    assert seqlens == [4, 8, 12, 16, 16, 16], seqlens                       # multiples of difficulty_step, capped
    assert all(c[1] is True for c in net.calls)
    thetas = [c[2] for c in net.calls]
    assert thetas[0] == 1.0 and all(a >= b for a, b in zip(thetas, thetas[1:]))
    # the schedule is advanced with the step count BEFORE it is incremented (as in the reference): 6th forward sees t = 4
    assert abs(thetas[-1] - (0.5 * math.exp(-0.1 * 4) + 0.5)) < 1e-9
    net.calls.clear()
    engine.eval()
    with torch.no_grad():
        engine(ids)
    assert net.calls == [(None, None, None)]                                 # evaluation sees the plain signature


def test_legacy_curriculum_and_pld_reach_the_forward():
    from conftest import run_dist
    run_dist(_curriculum_pld_worker, 1)
