"""Small helpers of the IO benchmark: size parsing, report lines, buffer creation."""
from __future__ import annotations

import os
from typing import List

import torch

BYTES_PER_GB = 1024 ** 3
BYTES_PER_MB = 1024 ** 2
BYTES_PER_KB = 1024
LOG_TIDS = [0]
_SUFFIX = {"K": BYTES_PER_KB, "M": BYTES_PER_MB, "G": BYTES_PER_GB}


def task_log(tid, msg, force=False):
    if force or tid in LOG_TIDS:
        print(f"tid {tid}: {msg}")


def task_barrier(barrier, num_parties):
    assert barrier.parties == num_parties
    barrier.wait()
    assert not barrier.broken


def report_results(args, read_op, pool_results):
    """Prints the two lines ``parse_nvme_stats`` greps for: ``... Latency = X usec`` and ``... Speed = Y GB/sec``."""
    io_string = "Read" if read_op else "Write"
    if None in pool_results:
        print(f"Failure in one of {args.multi_process} {io_string} processes")
        return
    total_bytes = sum(num_bytes for _, _, num_bytes in pool_results)
    task_latency = max(e2e for e2e, _, _ in pool_results)
    task_speed = total_bytes / task_latency / BYTES_PER_GB
    print(f"Task {io_string} Latency = {task_latency * 1e6:.0f} usec")
    print(f"Task {io_string} Speed = {task_speed:.2f} GB/sec")
    e2e_latency = sum(sec for _, sec, _ in pool_results) / len(pool_results)
    print(f"E2E {io_string} Latency = {e2e_latency * 1e6:.0f} usec")
    print(f"E2E {io_string} Speed = {total_bytes / e2e_latency / BYTES_PER_GB / len(pool_results):.2f} GB/sec")


def get_block_size_and_count(io_bytes):
    """Largest power-of-two-ish (block, count) with block <= 1 MiB used to fill a test file."""
    if io_bytes > BYTES_PER_MB and io_bytes % BYTES_PER_MB == 0:
        return BYTES_PER_MB, io_bytes // BYTES_PER_MB
    if io_bytes > BYTES_PER_KB and io_bytes % BYTES_PER_KB == 0:
        return BYTES_PER_KB, io_bytes // BYTES_PER_KB
    return io_bytes, 1


def refine_integer_value(value) -> int:
    s = str(value).strip()
    if s and s[-1].upper() in _SUFFIX:
        return int(float(s[:-1]) * _SUFFIX[s[-1].upper()])
    return int(s)


def create_filename(folder, read_op, size, tid):
    return os.path.join(folder, f"_aio_{'read' if read_op else 'write'}_{size}.pt.{tid}")


def create_file(filename, num_bytes):
    block, count = get_block_size_and_count(num_bytes)
    chunk = os.urandom(block)
    with open(filename, "wb") as f:
        for _ in range(count):
            f.write(chunk)


def create_buffer(num_bytes: int, gpu: bool, pinned: bool) -> torch.Tensor:
    if gpu and torch.cuda.is_available():
        return torch.randint(0, 255, (num_bytes,), dtype=torch.uint8, device="cuda")
    t = torch.randint(0, 255, (num_bytes,), dtype=torch.uint8)
    return t.pin_memory() if pinned and torch.cuda.is_available() else t
