"""Sub-allocator over one pre-allocated flat buffer, with defragmentation.

Reference: ``runtime/zero/contiguous_memory_allocator.py:16-287`` (used for contiguous activation-checkpoint and
parameter buffers).  Same contract: ``allocate_tensor(size)`` returns a view of the backing buffer; when the free
space is sufficient but fragmented, live tensors are compacted towards the front (their ``.data`` and any
parameters registered with ``assign_to_param`` are re-pointed) and the request is then served from the tail.
Implementation: an ordered ``{offset: size}`` table for live blocks, free space derived from the gaps
(first-fit-smallest), one ``copy_`` per moved block.
"""
from __future__ import annotations

import torch


class ContiguousMemoryAllocator:
    def __init__(self, size, dtype, device):
        self.buffer = torch.zeros(size, dtype=dtype, device=device)
        self.size = size
        self.total_free = size
        self.largest_contiguous = size
        self.max_allocated = 0
        self.count = 0
        self._live = {}            # offset -> size
        self._tensors = {}         # id -> (offset, tensor)
        self._params = {}          # id -> (param, numel, shape)

    # ------------------------------------------------------------------------------------ queries
    def _gaps(self):
        gaps, cursor = [], 0
        for off in sorted(self._live):
            if off > cursor:
                gaps.append((cursor, off - cursor))
            cursor = off + self._live[off]
        if cursor < self.size:
            gaps.append((cursor, self.size - cursor))
        return gaps

    def _refresh_stats(self):
        gaps = self._gaps()
        self.largest_contiguous = max((g[1] for g in gaps), default=0)
        self.max_allocated = max(self.max_allocated, self.size - self.total_free)

    def print_allocation(self, resolution=200):
        cells = ["."] * resolution
        for off, sz in self._live.items():
            for c in range(int(off * resolution / self.size), max(int((off + sz) * resolution / self.size), int(off * resolution / self.size) + 1)):
                if c < resolution:
                    cells[c] = "#"
        print("".join(cells))

    # ---------------------------------------------------------------------------------- allocate
    def allocate_tensor(self, size):
        assert size <= self.total_free, "Not enough memory in buffer. Allocation failed"
        if self.largest_contiguous < size:
            self._defragment_memory()
        gaps = [g for g in self._gaps() if g[1] >= size]
        off = min(gaps, key=lambda g: g[1])[0]         # smallest gap that fits
        self._live[off] = size
        self.total_free -= size
        t = self.buffer.narrow(0, off, size)
        self._tensors[id(t)] = (off, t)
        self.count += 1
        self._refresh_stats()
        return t

    def assign_to_param(self, tensor, param, numel, shape):
        tid = id(tensor)
        assert tid in self._tensors, "No such tensor allocated by the allocator."
        assert tid not in self._params, "This tensor has already been assigned to a param"
        self._params[tid] = (param, numel, shape)
        param.data = tensor.narrow(0, 0, numel).view(shape)

    def release_tensor(self, tensor):
        tid = id(tensor)
        off, _ = self._tensors.pop(tid)
        self._params.pop(tid, None)
        self.total_free += self._live.pop(off)
        self._refresh_stats()

    def release_tensor_with_id(self, tensor_id):
        off, t = self._tensors[tensor_id]
        self.release_tensor(t)

    # ---------------------------------------------------------------------------------- compaction
    def _defragment_memory(self):
        cursor = 0
        by_offset = sorted(self._tensors.items(), key=lambda kv: kv[1][0])
        new_live = {}
        for tid, (off, t) in by_offset:
            sz = self._live[off]
            if off != cursor:
                dst = self.buffer.narrow(0, cursor, sz)
                if cursor + sz <= off:
                    dst.copy_(t)
                else:                                   # overlapping move towards the front
                    dst.copy_(t.clone())
                t.data = dst
                self._tensors[tid] = (cursor, t)
                if tid in self._params:
                    param, numel, shape = self._params[tid]
                    param.data = t.narrow(0, 0, numel).view(shape)
            new_live[cursor] = sz
            cursor += sz
        self._live = new_live
        self._refresh_stats()
