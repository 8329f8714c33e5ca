"""Ragged inference engine: continuous batching (prompt chunks + decode steps, two interleaved sequences) must give
the same next-token logits as a dense causal forward of each full sequence."""
import torch

from shuffle_exchange_b200.inference.v2.config_v2 import RaggedInferenceEngineConfig
from shuffle_exchange_b200.inference.v2.engine_factory import build_engine_from_model
from shuffle_exchange_b200.inference.v2.engine_v2 import SchedulingResult
from shuffle_exchange_b200.models.llama import LlamaForCausalLM, llama_config


def _engine(dtype=torch.float32, device="cpu"):
    torch.manual_seed(0)
    cfg = llama_config("llama-tiny", activation_checkpointing=False)
    lm = LlamaForCausalLM(cfg, device=device, dtype=dtype).eval()
    ec = RaggedInferenceEngineConfig(kv_block_size=4)
    ec.state_manager.num_kv_blocks = 64
    ec.state_manager.max_context = 64
    ec.state_manager.max_ragged_batch_size = 64
    ec.state_manager.max_ragged_sequence_count = 8
    return lm, build_engine_from_model(lm, ec, device=device)


def test_put_matches_dense_forward_cpu():
    lm, eng = _engine()
    a = torch.randint(0, 1000, (11,))
    b = torch.randint(0, 1000, (7,))
    # prompt of A in two chunks, B joins in the second forward; then one decode token each
    eng.put([1], [a[:6]])
    out = eng.put([1, 2], [a[6:], b])
    with torch.no_grad():
        ref_a = lm(a[None])[0, -1]
        ref_b = lm(b[None])[0, -1]
    assert torch.allclose(out[0], ref_a, atol=2e-4, rtol=1e-3)
    assert torch.allclose(out[1], ref_b, atol=2e-4, rtol=1e-3)
    na, nb = out[0].argmax().view(1), out[1].argmax().view(1)
    out2 = eng.put([2, 1], [nb, na])
    with torch.no_grad():
        ref_a2 = lm(torch.cat([a, na])[None])[0, -1]
        ref_b2 = lm(torch.cat([b, nb])[None])[0, -1]
    assert torch.allclose(out2[1], ref_a2, atol=2e-4, rtol=1e-3)
    assert torch.allclose(out2[0], ref_b2, atol=2e-4, rtol=1e-3)


def test_scheduling_and_flush_cpu():
    _, eng = _engine()
    free0 = eng.free_blocks
    assert eng.can_schedule([5], [200]) != SchedulingResult.Success        # more tokens than a ragged batch holds
    tokens, blocks = eng.query(5, 10, free0)
    assert (tokens, blocks) == (10, 3)                                      # ceil(10 / 4)
    eng.put([5], [torch.randint(0, 1000, (10,))])
    assert eng.free_blocks == free0 - 3
    assert eng.get_remaining_block_capacity(5) == 2
    eng.flush(5)
    assert eng.free_blocks == free0


def test_offloaded_sequence_resumes_with_identical_logits_cpu():
    """Suspend a sequence to host memory, let another sequence take (and overwrite) its KV blocks, then continue the
    first one: the next-token logits must equal an uninterrupted run's."""
    lm, eng = _engine()
    a = torch.randint(0, 1000, (13,))
    b = torch.randint(0, 1000, (30,))
    free0 = eng.free_blocks
    out = eng.put([1], [a])
    nxt = out[0].argmax().view(1)
    released = eng.offload(1)
    assert released == 4 and eng.free_blocks == free0                      # ceil(13 / 4) blocks back on the free list
    assert eng.offload(1) == 0                                              # already on the host
    eng.put([2], [b])                                                       # reuses the released block ids
    tokens, blocks = eng.query(1, 1, eng.free_blocks)
    assert (tokens, blocks) == (1, 4)                                       # resuming needs its 4 blocks again (14 tokens)
    got = eng.put([1], [nxt])                                               # restores by itself
    with torch.no_grad():
        ref = lm(torch.cat([a, nxt])[None])[0, -1]
    assert torch.allclose(got[0], ref, atol=2e-4, rtol=1e-3)
    eng.offload(2)
    eng.flush(2)                                                            # flushing a suspended sequence frees nothing twice
    eng.flush(1)
    assert eng.free_blocks == free0
