"""Config sections under their reference import paths, data-efficiency config normalisation, comm call introspection,
IO staging buffers, launcher helper environment mapping."""
import os

import pytest
import torch


def test_section_configs_parse_from_the_whole_dictionary():
    from shuffle_exchange_b200.monitor.config import get_monitor_config
    from shuffle_exchange_b200.nebula.config import DeepSpeedNebulaConfig
    from shuffle_exchange_b200.profiling.config import DeepSpeedFlopsProfilerConfig
    from shuffle_exchange_b200.runtime.activation_checkpointing.config import DeepSpeedActivationCheckpointingConfig
    from shuffle_exchange_b200.runtime.precision_config import get_bfloat16_config, get_float16_config
    from shuffle_exchange_b200.utils.config import get_timers_config
    pd = {"fp16": {"enabled": True, "initial_scale_power": 8, "hysteresis": 3}, "bfloat16": {"enabled": True},
          "nebula": {"enabled": True, "persistent_storage_path": "/x", "num_of_version_in_retention": 5},
          "flops_profiler": {"enabled": True, "profile_step": 7},
          "activation_checkpointing": {"partition_activations": True, "number_checkpoints": 4},
          "csv_monitor": {"enabled": True, "output_path": "/tmp/o"}, "timers": {"throughput": {"synchronized": False}}}
    f = get_float16_config(pd)
    assert f.initial_dynamic_scale() == 256 and f.dynamic_loss_scale_args()["delayed_shift"] == 3
    assert get_bfloat16_config(pd).enabled                       # the old "bfloat16" spelling is honoured
    n = DeepSpeedNebulaConfig(pd)
    assert n.enabled and n.persistent_storage_path == "/x" and n.num_of_version_in_retention == 5 \
        and n.persistent_time_interval == 100
    assert DeepSpeedFlopsProfilerConfig(pd).profile_step == 7
    a = DeepSpeedActivationCheckpointingConfig(pd)
    assert a.partition_activations and a.number_checkpoints == 4 and not a.cpu_checkpointing
    m = get_monitor_config(pd)
    assert m.enabled and m.csv_monitor.output_path == "/tmp/o" and not m.tensorboard.enabled
    assert not get_timers_config(pd).synchronized and get_timers_config({}).enabled
    import shuffle_exchange_b200.pipe as pipe
    from shuffle_exchange_b200 import constants
    assert pipe.PipelineModule and pipe.LayerSpec and constants.TORCH_DISTRIBUTED_DEFAULT_PORT == 29500


def test_data_efficiency_config_is_fully_defaulted():
    from shuffle_exchange_b200.runtime.data_pipeline import config as dc
    full = dc.get_data_efficiency_config({})
    assert full == {"enabled": False, "seed": 1234,
                    "data_sampling": {"enabled": False, "num_epochs": 1000, "num_workers": 0, "pin_memory": False,
                                      "curriculum_learning": {"enabled": False},
                                      "dynamic_batching": {"enabled": False, "lr_scaling_method": "linear",
                                                           "min_batch_size": 1, "max_batch_size": None,
                                                           "sequence_picking_order": "dataloader", "verbose": False}},
                    "data_routing": {"enabled": False, "random_ltd": {"enabled": False,
                                                                       "layer_token_lr_schedule": {"enabled": False}}}}
    pd = {"data_efficiency": {"enabled": True, "seed": 7, "data_sampling": {
        "enabled": True, "num_workers": 2,
        "curriculum_learning": {"enabled": True, "curriculum_metrics": {"seqlen": {"min_difficulty": 8}}}},
        "data_routing": {"enabled": True, "random_ltd": {"enabled": True, "total_layer_num": 12,
                                                          "layer_token_lr_schedule": {"enabled": True}}}}}
    c = dc.get_data_efficiency_config(pd)
    assert c["seed"] == 7 and c["data_sampling"]["num_workers"] == 2
    assert c["data_sampling"]["curriculum_learning"]["curriculum_metrics"]["seqlen"]["min_difficulty"] == 8
    assert c["data_routing"]["random_ltd"]["total_layer_num"] == 12
    assert c["data_routing"]["random_ltd"]["layer_token_lr_schedule"]["enabled"]
    with pytest.raises(ValueError):
        dc.get_curriculum_learning({"curriculum_learning": {"enabled": True}})
    with pytest.raises(ValueError):
        dc.get_dynamic_batching({"dynamic_batching": {"enabled": True}})
    assert dc.get_curriculum_params_legacy({}) is False
    assert dc.get_curriculum_params_legacy({"curriculum_learning": {"enabled": True, "curriculum_type": "seqlen"}}) == \
        {"curriculum_type": "seqlen"}


def test_comm_utils_find_the_payload():
    from shuffle_exchange_b200.comm import utils as cu

    def all_reduce(tensor, op=None, group=None, async_op=False): ...
    def all_gather(tensor_list, tensor, group=None): ...
    t = torch.zeros(10, dtype=torch.float32)
    assert cu.get_tensor_position(all_reduce) == 0 and cu.get_msg_size_from_args(all_reduce, t) == 40
    assert cu.get_msg_size_from_args(all_reduce, tensor=t.half()) == 20
    assert cu.get_msg_size_from_args(all_gather, [t, t], t) == 40            # "tensor" is the payload by convention
    assert cu.get_default_args(all_reduce) == {"op": None, "group": None, "async_op": False}
    assert cu.get_debug_log_name({"log_name": "x"}, "caller") == "x | [Caller Func: caller]"
    old = dict(os.environ)
    try:
        os.environ.pop("LOCAL_RANK", None); os.environ["OMPI_COMM_WORLD_LOCAL_RANK"] = "3"
        assert cu.get_local_rank_from_launcher() == 3
    finally:
        os.environ.clear(); os.environ.update(old)


@pytest.mark.parametrize("kind", ["single", "double"])
def test_io_buffers_write_every_byte_in_order(tmp_path, kind):
    from shuffle_exchange_b200.io.double_io_buffer import Double_IO_Buffer
    from shuffle_exchange_b200.io.single_io_buffer import Single_IO_Buffer
    from shuffle_exchange_b200.ops.aio import aio_handle
    h = aio_handle(block_size=4096, queue_depth=4, single_submit=False, overlap_events=True, intra_op_parallelism=1)
    pinned = torch.empty(8192, dtype=torch.uint8).pin_memory() if torch.cuda.is_available() else torch.empty(8192, dtype=torch.uint8)
    buf = (Single_IO_Buffer if kind == "single" else Double_IO_Buffer)(pinned, h)
    data = torch.arange(0, 3000, dtype=torch.int32)               # 12000 bytes: several drains + an unaligned tail
    path = str(tmp_path / "out.bin")
    open(path, "wb").close()
    src_off, file_off = 0, 0
    total = data.numel() * 4
    while src_off < total:
        src_off += buf.fill(data, src_off)
        if buf.is_full() or src_off == total:
            n = buf.get_aligned_num_bytes()
            if n:
                buf.drain(n, path, file_off)
                file_off += n
    buf.complete_ongoing_drain()
    tail = buf.get_unaligned_num_bytes()
    assert tail == total % 1024 == buf.get_offset()
    with open(path, "r+b") as f:
        f.seek(file_off)
        f.write(buf.get_buffer()[:tail].numpy().tobytes())
    got = torch.frombuffer(bytearray(open(path, "rb").read()), dtype=torch.int32)
    assert torch.equal(got, data)
    buf.reset()
    assert buf.is_empty()


def test_launcher_helper_maps_mpi_environment():
    from shuffle_exchange_b200.launcher import launcher_helper as lh
    a = lh.parse_args(["--launcher", "mpich", "--module", "train.py", "--lr", "1"])
    assert a.launcher == "mpich" and a.module and a.user_script == "train.py" and a.user_args == ["--lr", "1"]
    env = lh.env_mapping({"PMI_RANK": "5", "MPI_LOCALRANKID": "1"}, ["PMI_RANK"], ["MPI_LOCALRANKID"])
    assert env["RANK"] == "5" and env["LOCAL_RANK"] == "1"
    with pytest.raises(EnvironmentError):
        lh.env_mapping({"A": "1", "B": "2", "L": "0"}, ["A", "B"], ["L"])
    with pytest.raises(EnvironmentError):
        lh.env_mapping({}, ["A"], ["L"])


def _backend_obj(rank, world):
    from shuffle_exchange_b200.comm.torch import TorchBackend
    cdb = TorchBackend(backend="gloo")
    assert cdb.is_initialized() and cdb.get_world_size() == world and cdb.get_rank() == rank
    t = torch.full((4,), float(rank + 1))
    cdb.all_reduce(t)
    assert t.tolist() == [3.0] * 4
    out = torch.empty(8)
    cdb.all_gather_into_tensor(out, torch.full((4,), float(rank)))
    assert out.tolist() == [0.0] * 4 + [1.0] * 4
    g = cdb.new_group([0, 1])
    assert cdb.get_all_ranks_from_group(g) == [0, 1] and cdb.has_all_gather_into_tensor()


def test_torch_backend_object_forwards_to_the_facade():
    from conftest import run_dist
    run_dist(_backend_obj, world=2)


def test_flag_protocol_checker_invariants():
    from shuffle_exchange_b200.comm.symm_check import MAX_PEERS, PadMonitor, check_pad
    world = 4
    def pad(rows):        # rows: {channel: [values per peer]}
        p = [0] * (8 * MAX_PEERS)
        for c, vals in rows.items():
            for i, v in enumerate(vals):
                p[c * MAX_PEERS + i] = v
        return p
    epochs = [5, 5, 0, 0, 2, 0, 0, 0]
    assert check_pad(pad({0: [5, 5, 6, 5], 1: [5, 5, 5, 5], 4: [2, 2, 2, 2]}), epochs, world) == []
    v = check_pad(pad({0: [5, 4, 5, 5]}), epochs, world, channels=[0])
    assert len(v) == 1 and v[0].rule == "completed waits" and v[0].peer == 1
    v = check_pad(pad({0: [5, 5, 8, 5]}), epochs, world, channels=[0])
    assert [x.rule for x in v] == ["bounded lead"] and "grads_ready" in str(v[0])
    v = check_pad(pad({0: [5, 5, 5, 5]}), epochs, world, prev=pad({0: [5, 6, 5, 5]}), channels=[0])
    assert [x.rule for x in v] == ["no regress"]
    v = check_pad(pad({4: [2, 2, 3, 2]}), epochs, world, quiescent=True, channels=[4])
    assert [x.rule for x in v] == ["quiescent agreement"]
    # wrap-around: 0xFFFFFFFF -> 0 is an advance, not a regress
    assert check_pad(pad({0: [0, 0, 0, 0]}), [0] * 8, world, prev=pad({0: [0xFFFFFFFF] * 4}), channels=[0]) == []

    from types import SimpleNamespace
    Buf = SimpleNamespace(epochs=epochs, world=world)
    mon = PadMonitor()
    assert mon.check(Buf, pad=pad({0: [5, 5, 5, 5], 1: [5] * 4, 4: [2] * 4})) == []
    with pytest.raises(AssertionError):
        mon.check(Buf, pad=pad({0: [5, 5, 4, 5], 1: [5] * 4, 4: [2] * 4}))


def test_released_parameters_are_poisoned_in_debug_mode(monkeypatch):
    from shuffle_exchange_b200.runtime.zero import flat_group as fgm
    monkeypatch.setattr(fgm, "_POISON_RELEASED", True)
    p = torch.nn.Parameter(torch.ones(8))
    fg = fgm.FlatGroup([p], ["w"], world=1, rank=0, process_group=None, device=torch.device("cpu"),
                       lp_dtype=torch.float32, keep_full_params=False)
    if fg.flat_lp is None:
        fg.allocate_full()
    stale = fg.flat_lp
    fg.release_params()
    assert torch.isnan(stale).all() and fg.flat_lp is None


def _shm_allreduce(rank, world):
    import time
    import torch.distributed as td
    from shuffle_exchange_b200 import comm as dist
    from shuffle_exchange_b200.comm.shm import get_shm_comm
    eng = get_shm_comm()
    eng.initialize(world, rank, name=f"sxb200_test_{os.environ['MASTER_PORT']}", slot_bytes=64 * 1024)
    g = torch.Generator().manual_seed(rank)
    for dtype, n, tol in ((torch.float32, 1000, 1e-6), (torch.bfloat16, 70001, 2e-2), (torch.float16, 33, 2e-3),
                          (torch.float32, 50000, 1e-5)):            # 50000 fp32 > one 64 KiB slot: several pieces
        x = torch.randn(n, generator=g).to(dtype)
        ref = x.clone().float()
        td.all_reduce(ref)
        eng.all_reduce(x)
        assert (x.float() - ref).abs().max() <= tol * max(1.0, ref.abs().max().item()), (dtype, n)
    # the facade picks the engine when the whole job is on this host
    os.environ["LOCAL_WORLD_SIZE"] = str(world)
    y = torch.full((5,), float(rank + 1))
    dist.inference_all_reduce(y)
    assert y.tolist() == [float(sum(range(1, world + 1)))] * 5
    with pytest.raises(TypeError):
        eng.all_reduce(torch.zeros(4, dtype=torch.int32))
    eng.finalize()


@pytest.mark.parametrize("world", [2, 3])
def test_shared_memory_allreduce_matches_gloo(world):
    from conftest import run_dist
    run_dist(_shm_allreduce, world=world)


def test_reference_named_helpers_lr_overrides_groups_debug_comm_switches(capsys):
    import argparse
    from shuffle_exchange_b200.comm import comm
    from shuffle_exchange_b200.comm.torch import all_reduce_comm_off, backward_comm_off
    from shuffle_exchange_b200.runtime import lr_schedules as L
    from shuffle_exchange_b200.utils import debug, groups
    from shuffle_exchange_b200.utils import tensor_fragment as tf
    p = {"warmup_max_lr": 0.1}
    L.override_params(argparse.Namespace(warmup_max_lr=0.5, cycle_min_lr=None), p)
    assert p == {"warmup_max_lr": 0.5} and L.update_lr([{"lr": 0}, {"lr": 0}], [1, 2]) == [1, 2]
    groups.set_tensor_model_parallel_world_size(4); groups.set_tensor_model_parallel_rank(3)
    try:
        assert groups.get_tensor_model_parallel_world_size() == 4 and groups.get_tensor_model_parallel_rank() == 3
    finally:
        groups.set_tensor_model_parallel_rank(None); groups.set_tensor_model_parallel_world_size(1)
    assert groups.get_tensor_model_parallel_src_rank() == 0
    m = torch.nn.Linear(2, 2)
    debug.debug_extract_module_and_param_names(m)
    assert "numel=4" in debug.debug_param2name_id_numel(m.weight) and "Linear" in debug.debug_module2name_class(m)
    debug.print_backward_tensors(m(torch.ones(1, 2)).sum())
    assert "leaf (2, 2)" in capsys.readouterr().out
    debug.debug_clear_module_and_param_names()
    assert debug.debug_param2name(m.weight) == "unknown"
    w = torch.zeros(4)
    tf.map_to_flat_opt_states(w, [], {w: {"exp_avg": torch.zeros(4)}}, ["exp_avg"])
    with pytest.raises(KeyError):
        tf.map_to_flat_opt_states(w, [], {w: {}}, ["exp_avg"])
    try:
        all_reduce_comm_off(True)
        assert comm._off("ALL_REDUCE") and isinstance(comm.all_reduce(torch.ones(1)), (comm.Noop, torch.Tensor, type(None)))
    finally:
        backward_comm_off(False)
    assert not comm._off("ALL_REDUCE") and not comm.in_aml() and comm.init_deepspeed_backend("hccl") is None
    with pytest.raises(ValueError):
        comm.init_deepspeed_backend("carrier-pigeon")


def test_numa_core_lists_and_memory_policy():
    import pytest
    from shuffle_exchange_b200.utils import numa
    assert numa.parse_range_list("0,2-4,6") == [0, 2, 3, 4, 6]
    assert list(numa.parse_range("7")) == [7] and list(numa.parse_range("2-3")) == [2, 3]
    for bad in ("3-1", "1-2-3", "a"):
        with pytest.raises(ValueError):
            numa.parse_range(bad)
    with pytest.raises(ValueError):
        numa.parse_range_list("0-3,2")                                   # overlap
    assert numa._format_cores([0, 1, 2, 5, 7, 8]) == "0-2,5,7-8"
    two = [[0, 1, 2, 3], [4, 5, 6, 7]]
    assert numa._memory_policy([4, 5], two) == ["-m", "1"]
    assert numa._memory_policy([3, 4], two) == []                        # spans both domains
    assert numa._memory_policy([4, 5], two + [[], []]) == ["-p", "3"]    # flat HBM: node 1 pairs with node 3
    assert numa._memory_policy([0], [[0, 1], [0, 1], [2, 3]]) == ["-m", "0,1"]   # fake NUMA
