import os
import socket
import sys
import traceback

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: test needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _entry(rank, world, port, fn, args, errq):
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    os.environ["SXB_LOG_LEVEL"] = "WARNING"
    try:
        import torch.distributed as td
        td.init_process_group("gloo", rank=rank, world_size=world)
        fn(rank, world, *args)
        td.barrier()
        td.destroy_process_group()
    except Exception:
        errq.put((rank, traceback.format_exc()))
        raise


def run_dist(fn, world=2, args=()):
    """Run ``fn(rank, world, *args)`` in ``world`` gloo processes (the reference ships no test harness;
    this is the CPU multi-process fixture SURVEY.md §4 asks for)."""
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    errq = ctx.SimpleQueue()
    port = free_port()
    procs = [ctx.Process(target=_entry, args=(r, world, port, fn, args, errq)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
    errs = []
    while not errq.empty():
        errs.append(errq.get())
    for p in procs:
        if p.is_alive():
            p.terminate()
            errs.append((-1, "timeout"))
    assert not errs, "\n".join(f"[rank {r}] {e}" for r, e in errs)
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
