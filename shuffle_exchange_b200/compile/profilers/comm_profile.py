"""Collective timing tables (reference ``compile/profilers/comm_profile.py``: ``run_all_gather`` / ``run_reduce_scatter``
over a size sweep and ``create_predictor`` interpolating between measured sizes)."""
from __future__ import annotations

import bisect
import time
from typing import Callable, List, Tuple

import torch
import torch.distributed as dist


def _time(fn: Callable, iters: int = 5, warmup: int = 2) -> float:
    for _ in range(warmup):
        fn()
    if torch.cuda.is_available():
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(iters):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) * 1e3 / iters
    t = time.perf_counter()
    for _ in range(iters):
        fn()
    return (time.perf_counter() - t) * 1e6 / iters


def run_all_gather(device, dtype, max_numel: int, group=None) -> List[Tuple[int, float]]:
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    out = []
    n = 1 << 12
    while n <= max_numel:
        src = torch.empty(n, dtype=dtype, device=device)
        dst = torch.empty(n * world, dtype=dtype, device=device)
        fn = (lambda: dist.all_gather_into_tensor(dst, src, group=group)) if world > 1 else (lambda: dst.copy_(src))
        out.append((n * world * src.element_size(), _time(fn)))
        n <<= 2
    return out


def run_reduce_scatter(device, dtype, max_numel: int, group=None) -> List[Tuple[int, float]]:
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    out = []
    n = 1 << 12
    while n <= max_numel:
        src = torch.empty(n * world, dtype=dtype, device=device)
        dst = torch.empty(n, dtype=dtype, device=device)
        fn = (lambda: dist.reduce_scatter_tensor(dst, src, group=group)) if world > 1 else \
            (lambda: dst.copy_(src[:n]))
        out.append((n * world * src.element_size(), _time(fn)))
        n <<= 2
    return out


def create_predictor(table: List[Tuple[int, float]]) -> Callable[[int], float]:
    """Piecewise-linear ``bytes -> microseconds`` from a measured table (extrapolates with the last slope)."""
    table = sorted(table)
    xs = [b for b, _ in table]
    ys = [t for _, t in table]

    def predict(nbytes: int) -> float:
        if not xs:
            return 0.0
        if len(xs) == 1:
            return ys[0] * nbytes / max(xs[0], 1)
        i = min(max(bisect.bisect_left(xs, nbytes), 1), len(xs) - 1)
        x0, x1, y0, y1 = xs[i - 1], xs[i], ys[i - 1], ys[i]
        return max(0.0, y0 + (y1 - y0) * (nbytes - x0) / max(x1 - x0, 1))

    return predict


def estimate_us_per_byte(engine) -> Tuple[float, float]:
    """Cheap default when no sweep was run: NVLink-5 all-gather / reduce-scatter at ~350 GB/s effective per GPU."""
    world = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
    if world == 1:
        return 0.0, 0.0
    per_byte = 1e6 / 350e9
    return per_byte, per_byte
