"""Hand-runnable self-check of :class:`ContiguousMemoryAllocator` (reference: ``runtime/zero/test.py``, two ad-hoc
functions run by hand).  ``python -m shuffle_exchange_b200.runtime.zero.test`` runs a randomized allocate / release
workload and verifies after every operation that live tensors kept their contents (defragmentation moves them), never
overlap, and that the allocator's accounting matches; ``tests/test_misc_runtime.py`` runs the same checker."""
from __future__ import annotations

import random

import torch

from .contiguous_memory_allocator import ContiguousMemoryAllocator


def check_invariants(mem: ContiguousMemoryAllocator, live: dict) -> None:
    """``live``: tag -> tensor handed out by ``mem`` and filled with ``tag``."""
    spans = []
    for tag, t in live.items():
        assert bool((t == tag).all()), f"tensor {tag} lost its contents"
        off = t.data_ptr() - mem.buffer.data_ptr()
        assert off % t.element_size() == 0 and 0 <= off // t.element_size() <= mem.size - t.numel()
        spans.append((off // t.element_size(), t.numel()))
    spans.sort()
    for (a, n), (b, _) in zip(spans, spans[1:]):
        assert a + n <= b, "live tensors overlap"
    used = sum(n for _, n in spans)
    assert mem.total_free == mem.size - used, (mem.total_free, mem.size - used)
    assert mem.largest_contiguous <= mem.total_free


def random_workload(steps: int = 400, size: int = 4096, seed: int = 0, dtype=torch.float16, device="cpu") -> dict:
    rng = random.Random(seed)
    mem = ContiguousMemoryAllocator(size, dtype, device)
    live, tag, defrags = {}, 0, 0
    for _ in range(steps):
        want = rng.choice([16, 32, 64, 64, 128, 256, 512])
        if live and (rng.random() < 0.45 or want > mem.total_free):
            victim = rng.choice(list(live))
            mem.release_tensor(live.pop(victim))
        elif want <= mem.total_free:
            fragmented = mem.largest_contiguous < want
            tag += 1
            t = mem.allocate_tensor(want)
            t.fill_(tag % 1024)
            live[tag % 1024] = t
            defrags += int(fragmented)
        check_invariants(mem, live)
    return {"allocations": tag, "defragmentations": defrags, "max_allocated": mem.max_allocated}


def param_rebinding() -> None:
    """A parameter registered with ``assign_to_param`` follows its tensor through a compaction."""
    mem = ContiguousMemoryAllocator(256, torch.float32, "cpu")
    a = mem.allocate_tensor(64); b = mem.allocate_tensor(64); c = mem.allocate_tensor(64)
    p = torch.nn.Parameter(torch.empty(0))
    mem.assign_to_param(c, p, 48, (6, 8))
    c.fill_(7.0)
    mem.release_tensor(a)
    d = mem.allocate_tensor(128)            # 64 free in front + 64 at the tail: needs compaction
    d.fill_(1.0)
    assert p.shape == (6, 8) and bool((p.data == 7.0).all()) and p.data_ptr() == c.data_ptr()
    assert bool((b == 0).all())


if __name__ == "__main__":
    print(random_workload())
    param_rebinding()
    print("ok")
