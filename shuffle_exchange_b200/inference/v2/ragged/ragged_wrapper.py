"""Host-side assembly of one ragged batch and its upload (reference: ``inference/v2/ragged/ragged_wrapper.py:31-292``).

All metadata of a forward pass lives in ONE pinned int32 staging buffer laid out as
``[token ids | token->sequence | token position | last-token index per sequence | block tables | attention atoms]`` and
is uploaded with a single ``cudaMemcpyAsync`` in ``finalize`` (the reference uses three pinned "fast host buffers",
``ragged/csrc/fast_host_buffer.cu``).  The attention atoms (one per 128-query tile of a sequence, with the KV extent
it attends to) are cut by the native scheduler ``sxbh_build_atoms`` (``csrc/host/ragged_host.cpp``) at ``finalize``
and consumed by the tensor-core paged attention kernel (``sxb_attn_fwd_atoms``)."""
from __future__ import annotations

import torch

ATOM_Q_BLOCK = 128          # query rows of one attention atom = the kernel's q tile
ATOM_KV_BLOCK = 64


class RaggedBatchWrapper:
    def __init__(self, max_ragged_batch_size: int, max_ragged_sequence_count: int, max_blocks_per_seq: int,
                 device=None) -> None:
        self.max_tokens, self.max_seqs, self.max_blocks = max_ragged_batch_size, max_ragged_sequence_count, max_blocks_per_seq
        self.device = device if device is not None else ("cuda" if torch.cuda.is_available() else "cpu")
        self.max_atoms = self.max_tokens // ATOM_Q_BLOCK + self.max_seqs
        n = 3 * self.max_tokens + self.max_seqs + self.max_seqs * self.max_blocks + 6 * self.max_atoms
        self._seq_meta = torch.zeros(3, self.max_seqs, dtype=torch.int32)      # start / in-flight tokens / seen tokens
        pin = torch.cuda.is_available()
        self._host = torch.zeros(n, dtype=torch.int32, pin_memory=pin)
        self._dev = torch.zeros(n, dtype=torch.int32, device=self.device)
        self.clear()

    def clear(self) -> None:
        self._n_tokens = 0
        self._n_seqs = 0
        self._n_atoms = 0
        self._finalized = False

    def _views(self, buf):
        T, S, B = self.max_tokens, self.max_seqs, self.max_blocks
        e = 3 * T + S + S * B
        return (buf[0:T], buf[T:2 * T], buf[2 * T:3 * T], buf[3 * T:3 * T + S], buf[3 * T + S:e].view(S, B),
                buf[e:].view(self.max_atoms, 6))

    def insert_sequence(self, seq_descriptor, tokens: torch.Tensor, do_checks: bool = True) -> None:
        n = tokens.numel()
        if do_checks:
            if self._n_tokens + n > self.max_tokens:
                raise RuntimeError(f"Ragged batch is full due to tokens: {self._n_tokens} + {n} > {self.max_tokens}")
            if self._n_seqs + 1 > self.max_seqs:
                raise RuntimeError("Ragged batch is full due to sequences")
        ids, t2s, pos, last, tables, _ = self._views(self._host)
        a, b = self._n_tokens, self._n_tokens + n
        self._seq_meta[0, self._n_seqs], self._seq_meta[1, self._n_seqs] = a, n
        self._seq_meta[2, self._n_seqs] = seq_descriptor.seen_tokens
        ids[a:b] = tokens.to(torch.int32)
        t2s[a:b] = self._n_seqs
        pos[a:b] = torch.arange(seq_descriptor.seen_tokens, seq_descriptor.seen_tokens + n, dtype=torch.int32)
        last[self._n_seqs] = b - 1
        blocks = seq_descriptor.kv_blocks()
        tables[self._n_seqs, : blocks.numel()] = blocks
        self._n_tokens, self._n_seqs = b, self._n_seqs + 1

    def finalize(self, padding: bool = False) -> None:
        if self._n_seqs:
            from ..kernels.atom_builder import AtomBuilder
            m = self._seq_meta[:, : self._n_seqs].contiguous()
            atoms, self._n_atoms = AtomBuilder()(self._views(self._host)[5], m[0], m[1], m[2], ATOM_Q_BLOCK, ATOM_KV_BLOCK)
            # longest first: the kernel runs the short tail (decode tokens, chunk remainders) with the GQA group packed
            # into one tile, and needs those atoms contiguous
            live = atoms[: self._n_atoms]
            order = torch.argsort(live[:, 2], descending=True, stable=True)
            live.copy_(live[order])
            self._atom_q_lens = live[:, 2].clone()
        self._dev.copy_(self._host, non_blocking=True)
        self._finalized = True

    # device views ------------------------------------------------------------------------------
    def input_ids(self):
        return self._views(self._dev)[0][: self._n_tokens]

    def token_to_seq(self):
        return self._views(self._dev)[1][: self._n_tokens]

    def token_positions(self):
        return self._views(self._dev)[2][: self._n_tokens]

    def last_token_index(self):
        return self._views(self._dev)[3][: self._n_seqs]

    def block_tables(self):
        return self._views(self._dev)[4]

    def atoms(self):
        """int32 ``[n_atoms, 6]`` on the device: ``(seq, q_start_idx, q_len, global_q_idx, kv_blocks, total_extent)``."""
        return self._views(self._dev)[5][: self._n_atoms]

    def atoms_longer_than(self, q_len: int) -> int:
        """How many (leading) atoms have more than ``q_len`` queries; host-side, no synchronisation."""
        return int((self._atom_q_lens > q_len).sum()) if self._n_atoms else 0

    @property
    def current_tokens(self) -> int:
        return self._n_tokens

    @property
    def current_sequences(self) -> int:
        return self._n_seqs
