"""FastGen v2: every supported architecture served through its policy must reproduce the HuggingFace logits
(prompt chunk, then a continuation chunk that reuses the paged KV cache).  CPU, fp32, random tiny models."""
import pytest
import torch

transformers = pytest.importorskip("transformers")
from transformers import AutoConfig, AutoModelForCausalLM  # noqa: E402

from shuffle_exchange_b200.inference.v2.config_v2 import RaggedInferenceEngineConfig  # noqa: E402
from shuffle_exchange_b200.inference.v2.engine_factory import build_engine_from_hf_model, build_hf_engine  # noqa: E402

_C = dict(vocab_size=128, hidden_size=64, num_hidden_layers=2, num_attention_heads=4, max_position_embeddings=128)
CASES = {
    "llama": dict(model_type="llama", intermediate_size=96, num_key_value_heads=2, **_C),
    "mistral": dict(model_type="mistral", intermediate_size=96, num_key_value_heads=2, sliding_window=None, **_C),
    "qwen2": dict(model_type="qwen2", intermediate_size=96, num_key_value_heads=2, **_C),
    "phi3": dict(model_type="phi3", intermediate_size=96, num_key_value_heads=4, pad_token_id=0, **_C),
    "mixtral": dict(model_type="mixtral", intermediate_size=96, num_key_value_heads=2, num_local_experts=4,
                    num_experts_per_tok=2, **_C),
    "qwen2_moe": dict(model_type="qwen2_moe", intermediate_size=96, num_key_value_heads=2, num_experts=4,
                      num_experts_per_tok=2, moe_intermediate_size=48, shared_expert_intermediate_size=80,
                      decoder_sparse_step=1, mlp_only_layers=[], **_C),
    "opt": dict(model_type="opt", vocab_size=128, hidden_size=64, num_hidden_layers=2, num_attention_heads=4,
                ffn_dim=96, max_position_embeddings=128, word_embed_proj_dim=64),
    "phi": dict(model_type="phi", intermediate_size=96, partial_rotary_factor=0.5, **_C),
    "falcon": dict(model_type="falcon", vocab_size=128, hidden_size=64, num_hidden_layers=2, num_attention_heads=4,
                   multi_query=True, parallel_attn=True, new_decoder_architecture=False, alibi=False, bias=False),
    "falcon_new": dict(model_type="falcon", vocab_size=128, hidden_size=64, num_hidden_layers=2,
                       num_attention_heads=4, num_kv_heads=2, parallel_attn=True, new_decoder_architecture=True,
                       alibi=False, bias=False),
}


def _model(name):
    torch.manual_seed(0)
    m = AutoModelForCausalLM.from_config(AutoConfig.for_model(**CASES[name])).float().eval()
    g = torch.Generator().manual_seed(11)
    with torch.no_grad():          # HF zero-initialises biases: randomise them so a misplaced bias shows up
        for n, p in m.named_parameters():
            if n.endswith("bias"):
                p.copy_(torch.randn(p.shape, generator=g) * 0.2)
            elif p.dim() == 1 and n.endswith("weight"):          # norm scales are initialised to exactly 1
                p.copy_(1 + torch.randn(p.shape, generator=g) * 0.1)
    return m


@pytest.mark.parametrize("name", sorted(CASES))
def test_policy_matches_huggingface_logits(name):
    m = _model(name)
    ids = torch.randint(0, 128, (1, 9), generator=torch.Generator().manual_seed(1))
    with torch.no_grad():
        ref = m(ids).logits[0]
    eng = build_engine_from_hf_model(m, RaggedInferenceEngineConfig())
    out = eng.put([1], [ids[0, :6]])
    assert (out[0] - ref[5]).abs().max() < 1e-4
    out = eng.put([1], [ids[0, 6:9]])              # continuation reads the cached keys/values
    assert (out[0] - ref[8]).abs().max() < 1e-4
    eng.flush(1)


def test_two_sequences_in_one_ragged_batch_and_hf_directory(tmp_path):
    m = _model("llama")
    a = torch.randint(0, 128, (7,), generator=torch.Generator().manual_seed(2))
    b = torch.randint(0, 128, (4,), generator=torch.Generator().manual_seed(3))
    with torch.no_grad():
        ra, rb = m(a[None]).logits[0, -1], m(b[None]).logits[0, -1]
    m.save_pretrained(str(tmp_path), safe_serialization=True)
    eng = build_hf_engine(str(tmp_path), RaggedInferenceEngineConfig())
    out = eng.put([10, 11], [a, b]).float()
    # a checkpoint directory is served in its stored dtype (fp32 here)
    assert (out[0] - ra).abs().max() < 1e-3 and (out[1] - rb).abs().max() < 1e-3


def test_unsupported_model_type_is_rejected():
    m = AutoModelForCausalLM.from_config(AutoConfig.for_model(model_type="gpt2", n_layer=1, n_embd=32, n_head=2,
                                                              vocab_size=64))
    with pytest.raises(ValueError):
        build_engine_from_hf_model(m)
