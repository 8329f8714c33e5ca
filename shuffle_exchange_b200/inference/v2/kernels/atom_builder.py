"""Attention-atom schedule builder (reference ``kernels/ragged_ops/atom_builder``: C++ ``build_atoms``).  Native host
code: ``csrc/host/ragged_host.cpp``.  An atom is one q-block of one sequence together with the KV extent it attends to;
schedulers and tile-based attention kernels consume the ``[n_atoms, 6]`` int32 table
``(seq_index, q_start_idx, q_len, global_q_idx, kv_blocks, total_extent)``."""
from __future__ import annotations

import ctypes
from typing import Tuple

import torch

from ....native import host as _host

c_p, c_i = ctypes.c_void_p, ctypes.c_int
_host.register({"sxbh_build_atoms": [c_p, c_i, c_p, c_p, c_p, c_i, c_i, c_i],
                "sxbh_fill_ragged_batch": [c_p, c_p, c_p, c_p, c_p, c_i]})


class AtomBuilder:

    def __call__(self, atoms: torch.Tensor, seq_start: torch.Tensor, seq_tokens: torch.Tensor, seq_seen: torch.Tensor,
                 q_block_size: int, kv_block_size: int) -> Tuple[torch.Tensor, int]:
        """All tensors int32 on the CPU; ``atoms`` is ``[max_atoms, 6]``."""
        if atoms.device.type != "cpu" or atoms.dtype != torch.int32:
            raise RuntimeError("AtomBuilder must be called on int32 CPU tensors")
        lib = _host.load()
        n = lib.sxbh_build_atoms(atoms.data_ptr(), atoms.shape[0], seq_start.data_ptr(), seq_tokens.data_ptr(),
                                 seq_seen.data_ptr(), seq_tokens.numel(), int(q_block_size), int(kv_block_size))
        if n < 0:
            raise RuntimeError("atoms buffer too small for this batch")
        return atoms, n


def build_atoms_from_batch(batch, seqs, q_block_size: int, kv_block_size: int) -> torch.Tensor:
    """Convenience: atoms for the sequences of a finalized batch (``seqs``: descriptors in insertion order)."""
    n_tok = torch.tensor([s.in_flight_tokens for s in seqs], dtype=torch.int32)
    seen = torch.tensor([s.seen_tokens for s in seqs], dtype=torch.int32)
    start = (torch.cumsum(n_tok, 0) - n_tok).to(torch.int32)
    max_atoms = int(((n_tok + q_block_size - 1) // q_block_size).sum())
    atoms = torch.empty(max(max_atoms, 1), 6, dtype=torch.int32)
    atoms, n = AtomBuilder()(atoms, start, n_tok, seen, q_block_size, kv_block_size)
    return atoms[:n]
