"""Hybrid engine: train <-> generate switching with LoRA folding, under ZeRO-1 and ZeRO-3 (CPU / gloo)."""
import pytest
import torch

from conftest import run_dist


class _LoRALinear(torch.nn.Module):
    """DeepSpeed-Chat style LoRA linear (``lora_right_weight [in, r]``, ``lora_left_weight [r, out]``)."""

    def __init__(self, i, o, r=4):
        super().__init__()
        self.weight = torch.nn.Parameter(torch.randn(o, i) * 0.1, requires_grad=False)
        self.lora_right_weight = torch.nn.Parameter(torch.randn(i, r) * 0.1)
        self.lora_left_weight = torch.nn.Parameter(torch.randn(r, o) * 0.1)
        self.lora_scaling = 0.5
        self.fuse_calls = 0

    def forward(self, x):
        y = torch.nn.functional.linear(x, self.weight)
        if not getattr(self, "_folded", False):
            y = y + (x @ self.lora_right_weight @ self.lora_left_weight) * self.lora_scaling
        return y


class _Toy(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.emb = torch.nn.Embedding(32, 16)
        self.l1 = _LoRALinear(16, 16)
        self.head = torch.nn.Linear(16, 32)
        self.seen_weight = None

    def forward(self, ids, labels=None):
        h = self.head(torch.tanh(self.l1(self.emb(ids))))
        if labels is None:
            return h
        return torch.nn.functional.cross_entropy(h.view(-1, 32).float(), labels.view(-1))

    def generate(self, input_ids, max_new_tokens=2, **kw):
        # records the (gathered, LoRA-folded) weight it ran with; greedy decode
        self.seen_weight = self.l1.weight.detach().clone()
        self.l1._folded = True
        try:
            ids = input_ids
            for _ in range(max_new_tokens):
                nxt = self.forward(ids)[:, -1].argmax(-1, keepdim=True)
                ids = torch.cat([ids, nxt], 1)
            return ids
        finally:
            self.l1._folded = False


def _worker(rank, world, stage):
    import shuffle_exchange_b200 as ds
    from shuffle_exchange_b200.runtime.hybrid_engine import DeepSpeedHybridEngine
    torch.manual_seed(0)
    model = _Toy()
    cfg = {"train_micro_batch_size_per_gpu": 2, "steps_per_print": 10 ** 9, "bf16": {"enabled": False},
           "optimizer": {"type": "AdamW", "params": {"lr": 1e-2}},
           "zero_optimization": {"stage": stage}, "hybrid_engine": {"enabled": True, "pin_parameters": True,
                                                                   "release_inference_cache": True}}
    params = [p for p in model.parameters() if p.requires_grad]
    engine, _, _, _ = ds.initialize(model=model, config=cfg, model_parameters=params)
    assert isinstance(engine, DeepSpeedHybridEngine)
    w0 = model.l1.weight.detach().clone() if stage < 3 else None
    ids = torch.randint(0, 32, (2, 5), generator=torch.Generator().manual_seed(rank))
    engine.eval()
    out = engine.generate(ids, max_new_tokens=3)
    assert out.shape == (2, 8)
    assert not engine.is_lora_fused
    if stage < 3:
        # generate ran on W + s * (R L)^T and the base weight was restored exactly afterwards
        delta = 0.5 * (model.l1.lora_right_weight @ model.l1.lora_left_weight).t()
        assert torch.allclose(model.seen_weight, w0 + delta, atol=1e-6)
        assert torch.allclose(model.l1.weight, w0, atol=1e-6)
        # folded generation == unfolded forward
        ref = ids
        for _ in range(3):
            ref = torch.cat([ref, model(ref)[:, -1].argmax(-1, keepdim=True)], 1)
        assert torch.equal(out, ref)
    else:
        assert model.seen_weight.numel() == 16 * 16          # whole parameter was visible during generate
    engine.train()
    loss = engine(ids, labels=ids)
    engine.backward(loss)
    engine.step()
    engine.eval()                                            # second eval() prints the latency breakdown
    out2 = engine.generate(input_ids=ids, max_new_tokens=1)
    assert out2.shape == (2, 6)
    lb = engine.latency_breakdown()
    assert lb["iters"] == 1.0 and lb["generate"] > 0


@pytest.mark.parametrize("stage", [1, 3])
def test_hybrid_engine_generate_train_cycle(stage):
    run_dist(_worker, world=2, args=(stage,))


def _hybrid_containers(rank, world):
    transformers = pytest.importorskip("transformers")
    from transformers import AutoConfig, AutoModelForCausalLM
    import shuffle_exchange_b200 as ds
    torch.manual_seed(0)
    model = AutoModelForCausalLM.from_config(AutoConfig.for_model(
        model_type="gpt2", vocab_size=64, n_embd=32, n_layer=2, n_head=4, n_positions=32)).float()
    cfg = {"train_micro_batch_size_per_gpu": 2, "steps_per_print": 1 << 30,
           "optimizer": {"type": "AdamW", "params": {"lr": 1e-2}}, "zero_optimization": {"stage": 1},
           "hybrid_engine": {"enabled": True}}
    engine, *_ = ds.initialize(model=model, config=cfg, model_parameters=model.parameters())
    assert engine.create_inference_module() == 2
    ids = torch.randint(0, 64, (2, 6), generator=torch.Generator().manual_seed(3))
    engine.eval()
    with torch.no_grad():
        ref = engine.module(ids, use_cache=False).logits.clone()
        engine.use_inference_layers(True)
        fused = engine.module(ids, use_cache=False).logits.clone()
        engine.use_inference_layers(False)
    assert (fused - ref).abs().max() < 2e-3
    # one training step changes the weights; a refresh carries them into the fused layers
    engine.train()
    out = engine(ids, labels=ids)
    engine.backward(out.loss)
    engine.step()
    engine.eval()
    engine.refresh_inference_module()
    with torch.no_grad():
        ref2 = engine.module(ids, use_cache=False).logits.clone()
        engine.use_inference_layers(True)
        fused2 = engine.module(ids, use_cache=False).logits.clone()
        engine.use_inference_layers(False)
    assert (ref2 - ref).abs().max() > 1e-4 and (fused2 - ref2).abs().max() < 2e-3
    lin = torch.nn.Sequential(torch.nn.Linear(4, 3))
    x = torch.randn(2, 4)
    y = lin(x)
    engine.convert_to_linear_transposed(lin)
    assert type(lin[0]).__name__ == "TLinear" and torch.allclose(lin(x), y, atol=1e-6)


def test_hybrid_engine_inference_containers_follow_training():
    run_dist(_hybrid_containers, 2)
