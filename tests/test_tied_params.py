"""A parameter used more than once in the forward (tied embedding / LM head, a shared projection) delivers one
gradient contribution per use through the main-grad protocol.  The sharded optimizers must launch the group / unit
reduction only after the LAST contribution (reference behaviour: autograd sums all uses before the per-parameter hook
fires, runtime/zero/stage_1_and_2.py:1090-1154, stage3.py:1175-1259)."""
import copy

import pytest
import torch
import torch.nn as nn

from conftest import run_dist


class TiedNet(nn.Module):
    """``shared`` is used twice through the protocol-following linear op; ``mid`` once; ``blocks`` form units."""

    def __init__(self, h=32):
        super().__init__()
        self.shared = nn.Parameter(torch.randn(h, h) * 0.2)
        self.blocks = nn.ModuleList([nn.Linear(h, h, bias=False) for _ in range(2)])

    def forward(self, x, protocol=True):
        from shuffle_exchange_b200.ops.functional import _Linear
        lin = (lambda a, w: _Linear.apply(a, w, None)) if protocol else (lambda a, w: torch.nn.functional.linear(a, w))
        h = lin(x, self.shared)
        for b in self.blocks:
            h = torch.tanh(lin(h, b.weight))
        h = lin(h, self.shared)            # second use of the tied weight, first to be differentiated
        return (h.float() ** 2).mean()


def _worker(rank, world, stage):
    import shuffle_exchange_b200 as ds
    from shuffle_exchange_b200.utils.tensor_fragment import safe_get_full_grad
    torch.manual_seed(0)
    model = TiedNet().to(torch.bfloat16)
    oracle = copy.deepcopy(model).float()
    cfg = {"train_micro_batch_size_per_gpu": 4, "steps_per_print": 10 ** 9, "bf16": {"enabled": True},
           "optimizer": {"type": "AdamW", "params": {"lr": 1e-3}},
           "zero_optimization": {"stage": stage, "fused_step": False}}
    engine, _, _, _ = ds.initialize(model=model, config=cfg, model_parameters=model.parameters(), slice_count=world)
    g = torch.Generator().manual_seed(3)
    xs = [torch.randn(4, 32, generator=g) for _ in range(world)]
    for it in range(2):                     # second iteration: the use counts come from a fresh window
        loss = engine(xs[rank].to(torch.bfloat16))
        engine.backward(loss)
        got = {n: safe_get_full_grad(p).float().clone() for n, p in engine.module.named_parameters()}
        oracle.zero_grad()
        for x in xs:
            (oracle(x.to(torch.bfloat16).float(), protocol=False) / world).backward()
        for n, p in oracle.named_parameters():
            ref = p.grad
            err = (got[n] - ref).abs().max().item()
            assert err <= 0.05 * ref.abs().max().item() + 1e-4, (stage, it, n, err, ref.abs().max().item())
        engine.step()
        for p, q in zip(oracle.parameters(), engine.module.parameters()):
            from shuffle_exchange_b200.utils.tensor_fragment import safe_get_full_fp32_param
            p.data.copy_(safe_get_full_fp32_param(q).to(torch.bfloat16).float())


@pytest.mark.parametrize("stage", [1, 2, 3])
@pytest.mark.parametrize("world", [1, 2])
def test_tied_parameter_gradients(stage, world):
    run_dist(_worker, world, (stage,))
