"""Benchmark engines + the multi-process driver.

Every engine is a class with ``prepare(args, tid, read_op) -> ctx``, ``run(ctx) -> None`` and ``finish(ctx)``;
``io_engine_multiprocessing`` forks ``--multi_process`` workers, lines them up on a barrier around the timed loop and
returns ``(task_latency, e2e_latency, bytes)`` per worker (reference ``nvme/io_engine.py:17-126`` +
``ds_aio_handle.py`` / ``ds_aio_basic.py`` / ``torch_io.py`` / ``torch_fastio_engine.py``)."""
from __future__ import annotations

import os
import time
from multiprocessing import Barrier, Pool
from typing import Dict

import torch

from .ds_aio_constants import AIO_BASIC, AIO_HANDLE, TORCH_FAST_IO, TORCH_IO
from .test_ds_aio_utils import create_buffer, create_file, create_filename, report_results, task_barrier, task_log


def _folder_for(args, tid: int) -> str:
    if args.folder is not None:
        return args.folder
    folders = list(args.folder_to_device_mapping)
    return folders[tid % len(folders)]


class _EngineBase:
    name = "base"

    def prepare(self, args, tid, read_op) -> Dict:
        folder = _folder_for(args, tid)
        fname = create_filename(folder, read_op, args.io_size, tid)
        if read_op and (not os.path.isfile(fname) or os.path.getsize(fname) != args.io_size):
            create_file(fname, args.io_size)
        buf = create_buffer(args.io_size, gpu=args.gpu and self.device_capable, pinned=self.wants_pinned(args))
        return {"file": fname, "buffer": buf, "num_bytes": args.io_size, "read": read_op, "args": args, "iter": 0}

    device_capable = False

    def wants_pinned(self, args) -> bool:
        return not args.slow_bounce_buffer

    def _file(self, ctx) -> str:
        if ctx["args"].different_file_each_iteration and not ctx["read"]:
            ctx["iter"] += 1
            return f'{ctx["file"]}.{ctx["iter"]}'
        return ctx["file"]

    def run(self, ctx) -> None:
        raise NotImplementedError

    def finish(self, ctx) -> None:
        ctx["buffer"] = None


class AioHandleEngine(_EngineBase):
    """The native thread-pool engine; with ``--use_gds`` / ``--gpu`` device tensors go through the bounce pipeline."""
    name = AIO_HANDLE
    device_capable = True

    def prepare(self, args, tid, read_op):
        ctx = super().prepare(args, tid, read_op)
        kw = dict(block_size=args.block_size, queue_depth=args.queue_depth, single_submit=args.single_submit,
                  overlap_events=not args.sequential_requests, intra_op_parallelism=args.io_parallel or 1)
        if args.use_gds or ctx["buffer"].is_cuda:
            from ..ops.gds import gds_handle
            ctx["handle"] = gds_handle(**kw)
        else:
            from ..ops.aio import aio_handle
            ctx["handle"] = aio_handle(**kw)
        return ctx

    def run(self, ctx):
        h, f = ctx["handle"], self._file(ctx)
        if ctx["read"]:
            h.async_pread(ctx["buffer"], f)
        else:
            h.async_pwrite(ctx["buffer"], f)
        h.wait()

    def finish(self, ctx):
        ctx["handle"] = None
        super().finish(ctx)


class AioBasicEngine(AioHandleEngine):
    """Synchronous calls on a throw-away handle per operation (the reference's handle-less ``aio_read/aio_write``)."""
    name = AIO_BASIC
    device_capable = False

    def run(self, ctx):
        h, f = ctx["handle"], self._file(ctx)
        (h.sync_pread if ctx["read"] else h.sync_pwrite)(ctx["buffer"], f)


class TorchIOEngine(_EngineBase):
    name = TORCH_IO

    def run(self, ctx):
        f = self._file(ctx)
        if ctx["read"]:
            ctx["loaded"] = torch.load(f, weights_only=False) if ctx.get("torch_file") else open(f, "rb").read()
        else:
            torch.save(ctx["buffer"], f, _use_new_zipfile_serialization=not ctx["args"].torch_legacy_save)


class TorchFastIOEngine(_EngineBase):
    """``torch.save`` into the framework's ``FastFileWriter`` (double-buffered pinned staging + native engine)."""
    name = TORCH_FAST_IO

    def prepare(self, args, tid, read_op):
        assert not read_op, "torch_fastio is a write engine"
        ctx = super().prepare(args, tid, read_op)
        from ..ops.aio import aio_handle
        ctx["handle"] = aio_handle(block_size=args.block_size, queue_depth=args.queue_depth,
                                   single_submit=args.single_submit, overlap_events=not args.sequential_requests,
                                   intra_op_parallelism=args.io_parallel or 1)
        ctx["pinned"] = torch.empty(args.fast_io_size, dtype=torch.uint8,
                                    pin_memory=torch.cuda.is_available() and not args.slow_bounce_buffer)
        return ctx

    def run(self, ctx):
        from ..io import FastFileWriter, FastFileWriterConfig
        w = FastFileWriter(self._file(ctx), FastFileWriterConfig(dnvme_handle=ctx["handle"],
                                                                 pinned_tensor=ctx["pinned"], double_buffer=True))
        torch.save(ctx["buffer"], w, _use_new_zipfile_serialization=not ctx["args"].torch_legacy_save)
        w.close()


ENGINES = {e.name: e for e in (AioHandleEngine, AioBasicEngine, TorchIOEngine, TorchFastIOEngine)}

_barrier = None


def _init_tasklet(b):
    global _barrier
    _barrier = b


def io_engine_tasklet(pool_params):
    args, tid, read_op = pool_params
    n = args.multi_process
    engine = ENGINES[args.engine]()
    task_log(tid, f"schedule = {engine.name} {'read' if read_op else 'write'}")
    ctx = engine.prepare(args, tid, read_op)
    task_barrier(_barrier, n)
    loops = args.warmup_loops + args.loops
    start_at = 0 if args.include_warmup_time else args.warmup_loops
    t_start = time.time()
    op_time = 0.0
    for i in range(loops):
        if i == start_at:
            task_barrier(_barrier, n) if False else None
            t_start = time.time()
        t0 = time.time()
        engine.run(ctx)
        if ctx["buffer"] is not None and ctx["buffer"].is_cuda:
            torch.cuda.synchronize()
        if i >= start_at:
            op_time += time.time() - t0
    e2e = time.time() - t_start
    task_barrier(_barrier, n)
    engine.finish(ctx)
    timed = loops - start_at
    return op_time / timed, e2e / timed, args.io_size


def io_engine_multiprocessing(args, read_op):
    b = Barrier(args.multi_process)
    params = [(args, tid, read_op) for tid in range(args.multi_process)]
    if args.multi_process == 1:
        _init_tasklet(b)
        results = [io_engine_tasklet(params[0])]
    else:
        with Pool(processes=args.multi_process, initializer=_init_tasklet, initargs=(b,)) as pool:
            results = pool.map(io_engine_tasklet, params)
    report_results(args, read_op, results)
    return results
