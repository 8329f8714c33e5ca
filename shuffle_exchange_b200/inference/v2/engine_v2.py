"""``InferenceEngineV2``: continuous-batching inference over ragged batches with a paged KV cache.

Reference: ``inference/v2/engine_v2.py:30-268`` — ``put(batch_uids, batch_tokens)`` schedules one forward over an
arbitrary mix of prompt chunks and decode tokens and returns the next-token logits per sequence; ``query(uid,
max_request_tokens, max_request_blocks)`` tells a scheduler how much of a request fits; ``can_schedule`` checks a
whole candidate batch; ``flush(uid)`` returns a finished sequence's KV blocks.
"""
from __future__ import annotations

from typing import Iterable, List, Tuple

import torch

from .config_v2 import RaggedInferenceEngineConfig
from .ragged.ragged_manager import DSStateManager
from .ragged.ragged_wrapper import RaggedBatchWrapper


from .scheduling_utils import SchedulingError, SchedulingResult  # noqa: E402,F401


class InferenceEngineV2:
    def __init__(self, policy_or_model, engine_config: RaggedInferenceEngineConfig = None, device=None) -> None:
        """``policy_or_model``: a callable ``(state_manager) -> model`` or a training ``LlamaForCausalLM``."""
        self._config = engine_config or RaggedInferenceEngineConfig()
        self._base_mp_group = self._initialize_tp_group()
        if hasattr(policy_or_model, "set_mp_group"):
            policy_or_model.set_mp_group(self._base_mp_group)
        from .model_implementations.llama_v2 import RaggedLlamaModel
        if callable(policy_or_model) and not isinstance(policy_or_model, torch.nn.Module):
            builder = policy_or_model
        else:
            builder = lambda sm: RaggedLlamaModel(policy_or_model, sm)
        probe = builder(None)
        kv_cfg = probe.kv_cache_config(self._config.kv_block_size)
        self._state_manager = DSStateManager(self._config.state_manager, kv_cfg, device=device,
                                             mp_group=self._base_mp_group)
        self._model = builder(self._state_manager)
        sm = self._config.state_manager
        self._batch = RaggedBatchWrapper(sm.max_ragged_batch_size, sm.max_ragged_sequence_count,
                                         self._state_manager.max_blocks_per_seq, device=device)

    def _initialize_tp_group(self):
        """Tensor-parallel group over the first ``tp_size`` ranks (reference ``engine_v2.py:95-105``); every rank of
        the job must construct the engine so the collective group creation matches."""
        tp = self._config.tensor_parallel.tp_size
        if tp <= 1:
            return None
        import torch.distributed as dist
        from ...comm import comm
        if not dist.is_initialized():
            comm.init_distributed()
        if dist.get_world_size() < tp:
            raise ValueError(f"tensor_parallel.tp_size={tp} exceeds the world size {dist.get_world_size()}")
        return dist.new_group(ranks=list(range(tp)))

    @property
    def free_blocks(self) -> int:
        return self._state_manager.free_blocks

    @property
    def n_kv_cache_groups(self) -> int:
        return 1

    def model(self):
        return self._model

    def put(self, batch_uids: Iterable[int], batch_tokens: Iterable[torch.Tensor], do_checks: bool = True) -> torch.Tensor:
        batch_uids, batch_tokens = list(batch_uids), list(batch_tokens)
        if do_checks:
            res = self.can_schedule(batch_uids, [t.numel() for t in batch_tokens])
            if res != SchedulingResult.Success:
                raise SchedulingError(res)
        self._batch.clear()
        seqs = []
        for uid, tokens in zip(batch_uids, batch_tokens):
            seq = self._state_manager.get_or_create_sequence(uid)
            if seq.is_offloaded:
                self._state_manager.restore_sequence(uid)
            self._model.maybe_allocate_kv(seq, tokens.numel())
            seq.pre_forward(tokens.numel())
            self._batch.insert_sequence(seq, tokens.reshape(-1).cpu(), do_checks=do_checks)
            seqs.append(seq)
        self._batch.finalize()
        logits = self._model.forward(self._batch)
        for seq in seqs:
            seq.post_forward()
        return logits

    def query(self, uid: int, max_request_tokens: int, max_request_blocks: int) -> Tuple[int, int]:
        seq = self._state_manager.get_sequence(uid)
        if seq is None:
            if self._state_manager.n_tracked_sequences >= self._config.state_manager.max_tracked_sequences:
                return 0, 0
            from .ragged.sequence_descriptor import PlaceholderSequenceDescriptor
            seq = PlaceholderSequenceDescriptor()
        return self._model.get_kv_requirements(seq, max_request_tokens, max_request_blocks)

    def can_schedule(self, uids: Iterable[int], lengths: Iterable[int]) -> int:
        uids, lengths = list(uids), list(lengths)
        sm = self._config.state_manager
        new = sum(1 for u in uids if self._state_manager.get_sequence(u) is None)
        if self._state_manager.n_tracked_sequences + new > sm.max_tracked_sequences:
            return SchedulingResult.EngineSequenceLimitExceeded
        if len(uids) > sm.max_ragged_sequence_count:
            return SchedulingResult.BatchSequenceLimitExceeded
        if sum(lengths) > sm.max_ragged_batch_size:
            return SchedulingResult.BatchTokenLimitExceeded
        free = self._state_manager.free_blocks
        need_total = 0
        for uid, n in zip(uids, lengths):
            tokens, blocks = self.query(uid, n, free)
            if tokens != n:
                return SchedulingResult.KVCacheLimitExceeded
            seq = self._state_manager.get_sequence(uid)
            if (seq.seen_tokens if seq else 0) + n > sm.max_context:
                return SchedulingResult.SequenceTokenLimitExceeded
            need_total += blocks
        if need_total > free:
            return SchedulingResult.KVCacheLimitExceeded
        return SchedulingResult.Success

    def get_remaining_block_capacity(self, uid: int) -> int:
        seq = self._state_manager.get_sequence(uid)
        if seq is None:
            return 0
        return seq.cur_allocated_blocks * self._state_manager.kv_block_size - seq.seen_tokens

    def flush(self, uid: int) -> None:
        self._state_manager.flush_sequence(uid)

    def offload(self, uid: int) -> int:
        """Suspend an idle sequence: its KV blocks go to host memory and become free for other sequences; returns the
        number of blocks released.  ``put`` restores the sequence by itself when it is scheduled again."""
        return self._state_manager.offload_sequence(uid)

    def restore(self, uid: int) -> int:
        return self._state_manager.restore_sequence(uid)

    def serialize(self, save_path: str) -> None:
        """Write this rank's (sharded, fused, possibly quantised) parameters as one flat buffer plus its table of
        contents, and on TP rank 0 the model config (reference ``engine_v2.py:251-268``).  ``build_hf_engine`` /
        ``build_engine_from_ds_checkpoint`` reload the directory without touching the original checkpoint."""
        import os
        import pickle
        from .model_implementations.flat_model_helpers import (flatten_inference_model, make_metadata_filename,
                                                              make_model_config_filename, make_param_filename)
        m = self._model
        os.makedirs(save_path, exist_ok=True)
        if getattr(m, "flattened_params", None) is None:
            m.flattened_params, m.flattened_param_metadata = flatten_inference_model(m.weights, m.tp_rank, m.tp_size)
        torch.save(m.flattened_params.cpu(), make_param_filename(save_path, m.tp_rank, m.tp_size))
        with open(make_metadata_filename(save_path, m.tp_rank, m.tp_size), "w") as f:
            f.write(m.flattened_param_metadata.json())
        if m.tp_rank == 0:
            with open(make_model_config_filename(save_path), "wb") as f:
                pickle.dump(getattr(m, "model_config", None) or m.spec, f)
