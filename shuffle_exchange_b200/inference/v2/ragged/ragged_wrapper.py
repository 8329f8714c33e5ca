"""Host-side assembly of one ragged batch and its upload (reference: ``inference/v2/ragged/ragged_wrapper.py:31-292``).

All metadata of a forward pass lives in ONE pinned int32 staging buffer laid out as
``[token ids | token->sequence | token position | last-token index per sequence | block tables]`` and is uploaded with
a single ``cudaMemcpyAsync`` in ``finalize`` (the reference uses three pinned "fast host buffers",
``ragged/csrc/fast_host_buffer.cu``)."""
from __future__ import annotations

import torch


class RaggedBatchWrapper:
    def __init__(self, max_ragged_batch_size: int, max_ragged_sequence_count: int, max_blocks_per_seq: int,
                 device=None) -> None:
        self.max_tokens, self.max_seqs, self.max_blocks = max_ragged_batch_size, max_ragged_sequence_count, max_blocks_per_seq
        self.device = device if device is not None else ("cuda" if torch.cuda.is_available() else "cpu")
        n = 3 * self.max_tokens + self.max_seqs + self.max_seqs * self.max_blocks
        pin = torch.cuda.is_available()
        self._host = torch.zeros(n, dtype=torch.int32, pin_memory=pin)
        self._dev = torch.zeros(n, dtype=torch.int32, device=self.device)
        self.clear()

    def clear(self) -> None:
        self._n_tokens = 0
        self._n_seqs = 0
        self._finalized = False

    def _views(self, buf):
        T, S, B = self.max_tokens, self.max_seqs, self.max_blocks
        return (buf[0:T], buf[T:2 * T], buf[2 * T:3 * T], buf[3 * T:3 * T + S], buf[3 * T + S:].view(S, B))

    def insert_sequence(self, seq_descriptor, tokens: torch.Tensor, do_checks: bool = True) -> None:
        n = tokens.numel()
        if do_checks:
            if self._n_tokens + n > self.max_tokens:
                raise RuntimeError(f"Ragged batch is full due to tokens: {self._n_tokens} + {n} > {self.max_tokens}")
            if self._n_seqs + 1 > self.max_seqs:
                raise RuntimeError("Ragged batch is full due to sequences")
        ids, t2s, pos, last, tables = self._views(self._host)
        a, b = self._n_tokens, self._n_tokens + n
        ids[a:b] = tokens.to(torch.int32)
        t2s[a:b] = self._n_seqs
        pos[a:b] = torch.arange(seq_descriptor.seen_tokens, seq_descriptor.seen_tokens + n, dtype=torch.int32)
        last[self._n_seqs] = b - 1
        blocks = seq_descriptor.kv_blocks()
        tables[self._n_seqs, : blocks.numel()] = blocks
        self._n_tokens, self._n_seqs = b, self._n_seqs + 1

    def finalize(self, padding: bool = False) -> None:
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

    @property
    def current_tokens(self) -> int:
        return self._n_tokens

    @property
    def current_sequences(self) -> int:
        return self._n_seqs
