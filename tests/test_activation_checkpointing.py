"""``deepspeed.checkpointing``: recomputation must not change outputs or gradients in any of its memory modes —
plain, inputs partitioned across the model-parallel group, partitions parked in host memory, partitions packed into
one contiguous arena — and must replay the random state (dropout) of the first forward."""
import pytest
import torch

from conftest import run_dist


class _Mpu:
    def __init__(self, group, rank, world):
        self._g, self._r, self._w = group, rank, world

    def get_model_parallel_rank(self):
        return self._r

    def get_model_parallel_world_size(self):
        return self._w

    def get_model_parallel_group(self):
        return self._g


def _block():
    torch.manual_seed(3)
    return torch.nn.Sequential(torch.nn.Linear(16, 32), torch.nn.GELU(), torch.nn.Dropout(0.25), torch.nn.Linear(32, 16))


def _run(fn, x, seed):
    torch.manual_seed(seed)
    y = fn(x)
    y.square().sum().backward()
    return y.detach().clone(), x.grad.detach().clone()


def _worker(rank, world, mode):
    import torch.distributed as td
    from shuffle_exchange_b200.runtime.activation_checkpointing import checkpointing as ck
    ck.reset()
    net = _block()
    net.train()
    x0 = torch.randn(4, 6, 16, generator=torch.Generator().manual_seed(5))        # replicated over the MP group
    xa = x0.clone().requires_grad_()
    want_y, want_g = _run(net, xa, seed=11)
    want_w = net[0].weight.grad.clone()
    net.zero_grad()
    kw = {"plain": {}, "partition": {"partition_activations": True},
          "cpu": {"partition_activations": True, "checkpoint_in_cpu": True},
          "contiguous": {"partition_activations": True, "contiguous_checkpointing": True, "num_checkpoints": 2}}[mode]
    ck.configure(_Mpu(td.group.WORLD, rank, world), **kw)
    assert ck.is_configured()
    for call in (ck.checkpoint, ck.non_reentrant_checkpoint):
        net.zero_grad()
        xb = x0.clone().requires_grad_()
        got_y, got_g = _run(lambda t: call(net, t), xb, seed=11)
        assert torch.equal(got_y, want_y), (mode, call.__name__)                   # same dropout mask in the first pass
        assert torch.allclose(got_g, want_g, atol=1e-6), (mode, call.__name__, float((got_g - want_g).abs().max()))
        assert torch.allclose(net[0].weight.grad, want_w, atol=1e-6)               # ... and in the recomputation
    if mode != "plain":
        # what a rank keeps of a replicated input is its 1/world slice
        assert ck.get_partition_size(x0) == x0.numel() // world
        assert ck.get_partition_start(x0) == rank * (x0.numel() // world)
    ck.reset()                      # frees the contiguous arenas; the configuration stays (as in the reference)
    assert ck.is_configured()


@pytest.mark.parametrize("mode", ["plain", "partition", "cpu", "contiguous"])
def test_checkpointed_block_matches_plain_backward(mode):
    run_dist(_worker, 2, (mode,))
