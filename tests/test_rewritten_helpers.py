"""Direct tests of helpers that were re-derived in round 2 and had only indirect coverage: data loader strategies and
the repeating loader, the write-behind tensor swapper, the sweep job runner, state merging of the checkpoint reshaper,
and the straight-through quantisers of the compression library."""
import os
import sys

import pytest
import torch
from torch.utils.data import Dataset


class _Range(Dataset):
    def __init__(self, n):
        self.n = n

    def __len__(self):
        return self.n

    def __getitem__(self, i):
        return torch.tensor(i)


def test_data_loader_strategies_and_repeating_loader():
    from shuffle_exchange_b200.runtime.dataloader import DeepSpeedDataLoader, RepeatingLoader
    ds = _Range(20)
    seen = []
    for rank in range(2):                                            # one process per device: each rank reads its shard
        ld = DeepSpeedDataLoader(ds, batch_size=4, pin_memory=False, local_rank=rank, tput_timer=None,
                                 num_local_io_workers=0, data_parallel_world_size=2, data_parallel_rank=rank)
        assert len(ld) == 3                                          # ceil(10 / 4)
        batches = list(ld)
        assert [len(b) for b in batches] == [4, 4, 2]
        seen += [int(x) for b in batches for x in b]
    assert sorted(seen) == list(range(20))                           # the two shards tile the dataset
    ld = DeepSpeedDataLoader(ds, batch_size=4, pin_memory=False, local_rank=0, tput_timer=None, num_local_io_workers=0,
                             data_parallel_world_size=2, data_parallel_rank=0, dataloader_drop_last=True)
    assert len(ld) == 2 and [len(b) for b in ld] == [4, 4]
    single = DeepSpeedDataLoader(ds, batch_size=5, pin_memory=False, local_rank=-1, tput_timer=None,
                                 num_local_io_workers=0)
    assert sum(len(b) for b in single) == 20                         # no launcher rank: one process sees everything
    rep = RepeatingLoader(ld)
    first_epoch = [next(rep) for _ in range(2)]
    second_epoch = [next(rep) for _ in range(2)]                     # wraps around instead of raising StopIteration
    assert rep.epochs_completed == 1 and ld.data_sampler.epoch == 1  # and tells the sampler (reshuffles per epoch)
    flat = lambda bs: sorted(int(x) for b in bs for x in b)          # noqa: E731
    assert len(flat(first_epoch)) == len(flat(second_epoch)) == 8


def test_write_behind_swapper_round_trips_through_files(tmp_path):
    from shuffle_exchange_b200.ops.aio import aio_handle
    from shuffle_exchange_b200.runtime.swap_tensor.async_swapper import AsyncTensorSwapper
    h = aio_handle(1 << 16, 8, False, True, 2)
    sw = AsyncTensorSwapper(h, numel_alignment=1024)
    with pytest.raises(RuntimeError):
        sw.swap_out_tensors([torch.zeros(4)], [str(tmp_path / "x")])          # no staging buffers yet
    sw.add_buffers([torch.empty(8192), torch.empty(8192)])
    with pytest.raises(RuntimeError):
        sw.add_buffers([torch.empty(8192)])
    tensors = [torch.randn(n) for n in (1000, 3000, 5000, 8000, 100, 2500)]  # forces several slot turnovers
    paths = [str(tmp_path / f"t{i}.swp") for i in range(len(tensors))]
    sw.swap_out_tensors(tensors, paths)
    sw.flush()
    assert sw.num_elements_swapped == sum((t.numel() + 1023) // 1024 * 1024 for t in tensors)
    for t, p in zip(tensors, paths):
        back = torch.empty((t.numel() + 1023) // 1024 * 1024)
        assert h.sync_pread(back, p) >= 0
        assert torch.equal(back[: t.numel()], t)
    with pytest.raises(ValueError):
        sw.swap_out_tensors([torch.zeros(9000)], [str(tmp_path / "big")])     # larger than a staging buffer
    bufs = sw.release_buffers()
    assert len(bufs) == 2 and not sw.has_buffers()


def test_sweep_job_runner(tmp_path):
    from shuffle_exchange_b200.nvme.ds_aio_job import Job, run_job
    out = tmp_path / "log.txt"
    job = Job([sys.executable, "-c 'import sys; print(sys.argv[1:])'", "--block_size 1M", "--queue_depth 8"],
              output_file=str(out), work_dir=str(tmp_path))
    run_job(job)
    assert job.returncode == 0 and out.read_text().strip() == "['--block_size', '1M', '--queue_depth', '8']"
    bad = Job([sys.executable, "-c", "import sys; sys.exit(4)"], split_pieces=False)
    with pytest.raises(RuntimeError, match="exit code 4"):
        run_job(bad)
    # no shell: metacharacters are data, not syntax
    marker = tmp_path / "pwned"
    Job([sys.executable, "-c", "pass", f"; touch {marker}"], split_pieces=False).run()
    assert not marker.exists()


def test_checkpoint_state_merging():
    from shuffle_exchange_b200.checkpoint.reshape_utils import merge_state, partition_data
    assert partition_data(list(range(6)), 3) == [[0, 1], [2, 3], [4, 5]]
    with pytest.raises(AssertionError):
        partition_data(list(range(5)), 2)
    a = {"w": torch.ones(2), "l": [torch.zeros(1), 3], "step": 5, "t": (torch.ones(1),)}
    b = {"w": torch.full((3,), 2.0), "l": [torch.ones(2), 3], "step": 5, "t": (torch.zeros(2),)}
    m = merge_state(a, b)
    assert m["w"].tolist() == [1, 1, 2, 2, 2] and m["l"][0].tolist() == [0, 1, 1] and m["step"] == 5
    assert isinstance(m["t"], tuple) and m["t"][0].tolist() == [1, 0, 0]


def test_straight_through_quantisers():
    from shuffle_exchange_b200.compression.utils import (AsymQuantizer, BinaryQuantizer, SymQuantizer,
                                                         TernaryQuantizer, TopKBinarizer)
    torch.manual_seed(0)
    x = torch.randn(4, 64, requires_grad=True)
    for q, bits in ((SymQuantizer, 4), (AsymQuantizer, 4), (SymQuantizer, 8)):
        y = q.apply(x, bits, None, None, 4)
        assert y.shape == x.shape and len(torch.unique(y.detach().reshape(4, -1)[0])) <= 2 ** bits
        assert (y - x).abs().max() < (x.abs().max() * 2 / 2 ** bits) * 1.01       # within one step
        y.sum().backward()
        assert torch.equal(x.grad, torch.ones_like(x))                             # straight-through gradient
        x.grad = None
    t = TernaryQuantizer.apply(x, 2, None, None, 4).detach().reshape(4, -1)
    assert all(len(torch.unique(row)) <= 3 for row in t)
    b = BinaryQuantizer.apply(x, 1, None, None, 4).detach().reshape(4, -1)
    assert all(len(torch.unique(row)) == 2 and torch.allclose(row.abs(), row.abs()[0].expand_as(row)) for row in b)
    scores = torch.randn(10, 10, requires_grad=True)
    mask = TopKBinarizer.apply(scores, 0.3, False)
    assert mask.sum() == 30 and mask[scores >= scores.flatten().sort(descending=True).values[29]].all()
    mask.sum().backward()
    assert scores.grad is not None
