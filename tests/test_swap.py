"""NVMe swapping on a temp directory: optimizer-state swap out/in (sync and pipelined) and parameter shards."""
import pytest
import torch

AIO = {"block_size": 1 << 16, "queue_depth": 8, "intra_op_parallelism": 2, "single_submit": False,
       "overlap_events": True, "use_gds": False}


class _Cfg:
    buffer_count = 8
    pipeline_read = True
    pipeline_write = True


def _opt():
    torch.manual_seed(0)
    ps = [torch.nn.Parameter(torch.randn(300_000)) for _ in range(3)]
    opt = torch.optim.Adam(ps, lr=1e-2)
    for p in ps:
        p.grad = torch.randn_like(p)
    opt.step()
    for p in ps:
        p.grad = None
    return ps, opt


def test_partitioned_optimizer_swapper_roundtrip(tmp_path):
    from shuffle_exchange_b200.runtime.swap_tensor import PartitionedOptimizerSwapper
    ps, opt = _opt()
    sw = PartitionedOptimizerSwapper(_Cfg(), AIO, str(tmp_path), opt, 300_000, "cpu", torch.float32, None)
    ref = [(p.data.clone(), opt.state[p]["exp_avg"].clone(), opt.state[p]["exp_avg_sq"].clone()) for p in ps]
    for p in ps:
        sw.swap_out_optimizer_state(p)
        assert p.data.numel() == 0 and opt.state[p]["exp_avg"].numel() == 0
    for p, (w, m, v) in zip(ps, ref):
        sw.swap_in_optimizer_state(p)
        assert torch.equal(p.data, w) and torch.equal(opt.state[p]["exp_avg"], m) and torch.equal(opt.state[p]["exp_avg_sq"], v)
        p.grad = torch.ones_like(p)
        opt.step()                                   # update while resident
        sw.swap_out_optimizer_state(p)
        p.grad = None


def test_pipelined_swapper_prefetch(tmp_path):
    from shuffle_exchange_b200.runtime.swap_tensor import PipelinedOptimizerSwapper
    ps, opt = _opt()
    sw = PipelinedOptimizerSwapper(_Cfg(), AIO, str(tmp_path), opt, 300_000, "cpu", torch.float32, None)
    ref = [p.data.clone() for p in ps]
    for p in ps:
        sw.swap_out_optimizer_state(p)
    for i, p in enumerate(ps):
        nxt = ps[i + 1] if i + 1 < len(ps) else None
        sw.swap_in_optimizer_state(p, async_parameter=nxt)
        assert torch.equal(p.data, ref[i])
        sw.swap_out_optimizer_state(p, async_swap=True)
    sw.flush()


def test_param_swapper(tmp_path):
    from shuffle_exchange_b200.runtime.swap_tensor import AsyncPartitionedParameterSwapper, PartitionedParamStatus
    cfg = {"aio": AIO, "offload_param": {"nvme_path": str(tmp_path), "buffer_size": 1 << 19, "buffer_count": 4}}
    sw = AsyncPartitionedParameterSwapper(cfg, torch.bfloat16)
    shards = [torch.randn(200_000).bfloat16() for _ in range(3)]
    ref = [s.clone() for s in shards]
    for i, s in enumerate(shards):
        s.ds_id = i
    sw.swap_out_and_release(shards)
    assert all(s.status == PartitionedParamStatus.NOT_AVAILABLE for s in shards) and sw.available_swap_in_buffers() == 4
    assert sw.swap_in(shards[:2], async_op=True)
    assert shards[0].status == PartitionedParamStatus.INFLIGHT
    sw.synchronize_reads()
    assert torch.equal(shards[0], ref[0]) and torch.equal(shards[1], ref[1])
    sw.remove_partition_and_release_buffers(shards[:2])
    assert sw.swap_in([shards[2]], async_op=False) and torch.equal(shards[2], ref[2])


@pytest.mark.parametrize("double_buffer", [True, False])
def test_fast_file_writer_round_trips_torch_save(tmp_path, double_buffer):
    """torch.save through the asynchronous, buffered writer: many buffer flips (tiny staging memory), aligned body +
    unaligned tail, byte-identical to a plain torch.save and loadable."""
    import io as _io
    from shuffle_exchange_b200.io import FastFileWriter, FastFileWriterConfig, MockFileWriter, PyFileWriter
    torch.manual_seed(0)
    state = {"w": torch.randn(300, 257), "b": torch.arange(10007, dtype=torch.int64), "n": {"step": 3, "name": "x"},
             "h": torch.randn(5000).to(torch.bfloat16)}
    plain = _io.BytesIO()
    torch.save(state, plain)
    path = str(tmp_path / "fast.pt")
    cfg = FastFileWriterConfig(pinned_tensor=torch.empty(6 * 4096, dtype=torch.uint8), double_buffer=double_buffer)
    w = FastFileWriter(path, cfg)
    torch.save(state, w)
    w.close()
    w.close()                                            # idempotent
    raw = open(path, "rb").read()
    assert len(raw) == plain.getbuffer().nbytes and raw == plain.getvalue()
    back = torch.load(path, weights_only=False)
    assert torch.equal(back["w"], state["w"]) and torch.equal(back["h"], state["h"]) and back["n"] == state["n"]
    stats = w.get_stats() if hasattr(w, "get_stats") else w._stats
    assert sum(v for k, v in stats.items() if "bytes" in k and "aio" in k.lower()) >= len(raw) // 4096 * 4096 - 4096 * 6
    # the two reference writers used to measure serialisation cost
    m = MockFileWriter(str(tmp_path / "mock.pt"))
    torch.save(state, m)
    m.close()
    p = PyFileWriter(str(tmp_path / "py.pt"))
    torch.save(state, p)
    p.close()
    assert open(tmp_path / "py.pt", "rb").read() == raw
