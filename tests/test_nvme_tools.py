"""NVMe benchmark/sweep tools, device-tensor IO handle and the parallel checkpoint-writer factories."""
import os

import pytest
import torch


def test_single_run_engines_and_sweep_to_aio_config(tmp_path, capsys):
    from shuffle_exchange_b200.nvme import perf_generate_param as gp
    from shuffle_exchange_b200.nvme import perf_run_sweep as sw
    from shuffle_exchange_b200.nvme.test_ds_aio import ds_io_main
    d = str(tmp_path)
    for eng in ("aio_handle", "aio_basic", "torch_io", "torch_fastio"):
        r = ds_io_main(["--folder", d, "--io_size", "2M", "--engine", eng, "--loops", "1"])
        assert r[0][2] == 2 << 20 and r[0][0] > 0
    ds_io_main(["--folder", d, "--io_size", "2M", "--read", "--loops", "1", "--io_parallel", "2"])
    assert "E2E Read Speed" in capsys.readouterr().out
    with pytest.raises(SystemExit):
        ds_io_main(["--folder", os.path.join(d, "missing"), "--io_size", "1M"])
    cfg = tmp_path / "cfg.json"
    cfg.write_text('{"block_size": ["1M"], "queue_depth": [8, 16], "io_parallel": [1]}')
    logs = str(tmp_path / "logs")
    args = sw.parse_sweep_arguments(["--nvme_dir", d, "--io_size", "1M", "--log_dir", logs, "--sweep_config", str(cfg)])
    assert sw.validate_arguments(args)
    sw.sweep_main(args)
    assert sorted(os.listdir(os.path.join(logs, "_aio_bench_read_logs"))) == \
        ["read_block_overlap_p1_d16_bs1M.txt", "read_block_overlap_p1_d8_bs1M.txt"]
    block = gp.generate_main(logs)
    assert block["aio"]["block_size"] == 1 << 20 and block["aio"]["queue_depth"] in (8, 16)
    assert gp.convert_to_param(("read", "single", "sequential", "p4", "d64", "bs8M")) == \
        {"single_submit": "true", "overlap_events": "false", "thread_count": 4, "queue_depth": 64,
         "block_size": 8 << 20}
    lines = sw.get_sweep_cmd_lines({"block_size": ["1M", "2M"], "single_submit": [True, False]})
    assert len(lines) == 4 and ["--block_size 1M", "--single_submit"] in lines and ["--block_size 2M"] in lines


def test_gds_handle_host_path_roundtrip(tmp_path):
    from shuffle_exchange_b200.ops.gds import GDSBuilder
    h = GDSBuilder().load().gds_handle(block_size=1 << 16, queue_depth=4, intra_op_parallelism=2)
    src = torch.arange(100000, dtype=torch.int32)
    f = str(tmp_path / "t.bin")
    h.sync_pwrite(src, f)
    dst = torch.empty_like(src)
    h.sync_pread(dst, f)
    assert torch.equal(src, dst)
    t = h.pin_device_tensor(src)
    assert t is src
    h.unpin_device_tensor(src)


@pytest.mark.gpu
def test_gds_handle_device_tensor_roundtrip(tmp_path):
    from shuffle_exchange_b200.ops.gds import gds_handle
    h = gds_handle(block_size=1 << 20, queue_depth=8, intra_op_parallelism=4)
    h.BOUNCE_BYTES = 1 << 20
    src = torch.randn(3 * (1 << 20) // 4 + 123, device="cuda")
    f = str(tmp_path / "t.bin")
    h.sync_pwrite(src, f)
    dst = torch.zeros_like(src)
    h.sync_pread(dst, f)
    assert torch.equal(src, dst)


def _info(world, rank, machines, mpu=None, ep=None, pure_dp=True):
    from shuffle_exchange_b200.runtime.model_checkpointing import UniversalParallelInfo
    per = world // machines
    return UniversalParallelInfo(global_world_size=world, global_rank=rank, local_rank=rank % per, mpu_info=mpu,
                                 ep_info=ep, pure_dp=pure_dp, num_machines=machines, machine_rank=rank // per,
                                 num_sockets=2 * machines)


def test_data_parallel_writer_selection():
    from shuffle_exchange_b200.runtime.model_checkpointing import (CheckpointDataParallel as U,
                                                                   DataParallelWriterFactory as F, MPUInfo)
    # classic: only rank 0 writes
    assert F(_info(16, 0, 2), None).create_config(0, False).world_size == 1
    assert F(_info(16, 5, 2), None).create_config(0, False) is None
    assert F(_info(16, 5, 2), None).create_config(1, False).world_size == 1     # unique optimizer shards
    # replica: everyone writes 1/16
    c = F(_info(16, 5, 2), U.REPLICA).create_config(0, False)
    assert (c.world_size, c.rank) == (16, 5)
    # machine: one writer per node (ranks 0 and 8)
    writers = {r: F(_info(16, r, 2), U.MACHINE).create_config(0, False) for r in range(16)}
    assert [r for r, c in writers.items() if c is not None] == [0, 8]
    assert (writers[8].world_size, writers[8].rank) == (2, 1)
    # socket: 4 sockets -> ranks 0, 4, 8, 12
    writers = {r: F(_info(16, r, 2), U.SOCKET).create_config(0, False) for r in range(16)}
    assert [r for r, c in writers.items() if c is not None] == [0, 4, 8, 12]
    assert writers[12].rank == 3
    # ZeRO-3: every rank writes its own shard file whole
    assert F(_info(16, 7, 2), U.MACHINE).create_config(3, False).world_size == 1
    # TP=2 x DP=8 on 2 nodes (tp-major ranks): replicas of tp slice 1 are the odd ranks; one writer per node
    got = {}
    for r in range(16):
        peers = [p for p in range(16) if p % 2 == r % 2]
        mpu = MPUInfo(1, 0, 2, r % 2, 8, peers, peers.index(r))
        got[r] = F(_info(16, r, 2, mpu=mpu, pure_dp=False), U.MACHINE).create_config(0, False)
    assert [r for r, c in got.items() if c is not None] == [0, 1, 8, 9]
    assert (got[9].world_size, got[9].rank) == (2, 1)


def test_checkpoint_config_accessors_and_writer_factory(tmp_path):
    from shuffle_exchange_b200.runtime.model_checkpointing import (CheckpointWriterFactory, DataParallelWriterConfig,
                                                                   get_checkpoint_config)
    cfg = get_checkpoint_config({"checkpoint": {"tag_validation": "fail", "writer": {"type": "fast",
                                                                                     "data_parallel": "socket",
                                                                                     "io_buffer_size": 1 << 20}}})
    assert cfg["tag_validation"] == "FAIL" and cfg["writer"]["type"] == "FAST" and cfg["writer"]["data_parallel"] == "SOCKET"
    assert get_checkpoint_config({})["writer"] is None
    with pytest.raises(AssertionError):
        get_checkpoint_config({"checkpoint": {"writer": {"type": "carrier_pigeon"}}})
    aio = {"block_size": 1 << 20, "queue_depth": 8, "single_submit": False, "overlap_events": True,
           "intra_op_parallelism": 2, "use_gds": False}
    dp = DataParallelWriterConfig(world_size=2, rank=1, global_rank=1, local_rank=1, pure_dp=True)
    fac = CheckpointWriterFactory(cfg["writer"], aio, dp)
    w = fac.create_writer(str(tmp_path / "model_states.pt"), optimize_dp_state=True)
    assert w.file_path().endswith("model_states.pt-1.2")
    with pytest.raises(AssertionError):
        fac.create_writer(str(tmp_path / "other.pt"), False)
    torch.save({"a": torch.arange(10)}, w)
    fac.release_writer()
    for kind in ("mock", "python"):
        f2 = CheckpointWriterFactory(get_checkpoint_config({"checkpoint": {"writer": {"type": kind}}})["writer"],
                                     aio, dp)
        w2 = f2.create_writer(str(tmp_path / f"{kind}.pt"), False)
        torch.save({"a": torch.arange(10)}, w2)
        f2.release_writer()
    assert torch.equal(torch.load(str(tmp_path / "python.pt"))["a"], torch.arange(10))
