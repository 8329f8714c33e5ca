"""1-bit communication: packbits round trip, compressed all-reduce accuracy and 1-bit Adam convergence (gloo, 2 ranks)."""
import pytest
import torch

from conftest import run_dist


def test_packbits_roundtrip_cpu():
    from shuffle_exchange_b200.runtime.comm.compressed import packbits, unpackbits
    x = torch.randn(1003)
    p = packbits(x)
    assert p.numel() == (1003 + 7) // 8
    y = unpackbits(p, 1003, torch.float32, 2.0)
    assert torch.equal(y, torch.where(x >= 0, torch.tensor(2.0), torch.tensor(-2.0)))


def _compressed_allreduce_worker(rank, world):
    import torch.distributed as td
    from shuffle_exchange_b200.runtime.comm.compressed import CompressedBackend
    torch.manual_seed(7 + rank)
    be = CompressedBackend()
    n = 16 * world * 8
    werr, serr = torch.zeros(n), torch.zeros(n // world)
    # error feedback: the running SUM of compressed results tracks the running sum of exact averages
    exact_sum, comp_sum = torch.zeros(n), torch.zeros(n)
    for it in range(60):
        x = torch.randn(n) + 0.5
        exact = x.clone()
        td.all_reduce(exact)
        exact /= world
        out = be.compressed_allreduce(x.clone(), werr, serr, 0)
        exact_sum += exact
        comp_sum += out
    rel = (comp_sum - exact_sum).norm() / exact_sum.norm()
    assert rel < 0.25, rel


def test_compressed_allreduce_error_feedback():
    run_dist(_compressed_allreduce_worker, 2)


def _onebit_adam_worker(rank, world):
    from shuffle_exchange_b200.runtime.fp16.onebit import OnebitAdam
    import torch.distributed as td
    torch.manual_seed(0)
    w = torch.nn.Parameter(torch.zeros(64))
    target = torch.where(torch.arange(64) % 2 == 0, 1.0, -1.0)   # equal gradient scale per element
    opt = OnebitAdam([w], lr=0.02, freeze_step=20, comm_backend_name="compressed")
    g = torch.Generator().manual_seed(100 + rank)
    for it in range(300):
        noise = torch.randn(64, generator=g) * 0.1
        loss = ((w - target + noise) ** 2).sum()
        opt.zero_grad()
        loss.backward()
        if not opt.adam_freeze_key:        # warm-up: the "engine" averages gradients
            td.all_reduce(w.grad)
            w.grad /= world
        opt.step()
    assert opt.adam_freeze_key
    err = (w.detach() - target).abs().max().item()
    assert err < 0.2, err
    ws = [torch.zeros_like(w.data) for _ in range(world)]
    td.all_gather(ws, w.data)
    assert torch.allclose(ws[0], ws[1], atol=1e-6)     # replicas stay identical under compressed communication


def test_onebit_adam_converges():
    run_dist(_onebit_adam_worker, 2)


def _other_onebit_worker(rank, world, which):
    """1-bit LAMB and 0/1 Adam on the same noisy quadratic: pass their warm-up / variance-freeze point, converge, and
    keep the replicas identical under compressed communication."""
    import torch.distributed as td
    from shuffle_exchange_b200.runtime.fp16.onebit import OnebitLamb, ZeroOneAdam
    torch.manual_seed(0)
    ws_ = [torch.nn.Parameter(torch.zeros(64)), torch.nn.Parameter(torch.zeros(8, 16))]
    target = [torch.where(torch.arange(64) % 2 == 0, 1.0, -1.0), torch.full((8, 16), 0.5)]
    if which == "lamb":
        opt = OnebitLamb(ws_, lr=0.05, freeze_step=20, comm_backend_name="compressed", max_coeff=2.0, min_coeff=0.1)
        frozen = lambda: opt.lamb_freeze_key            # noqa: E731
    else:
        opt = ZeroOneAdam(ws_, lr=0.02, var_freeze_step=20, var_update_scaler=4, local_step_scaler=40,
                          local_step_clipper=4, comm_backend_name="compressed")
        frozen = lambda: opt.freeze_key                 # noqa: E731
    g = torch.Generator().manual_seed(100 + rank)
    first = None
    for it in range(300):
        loss = sum(((w - t + torch.randn(w.shape, generator=g) * 0.1) ** 2).sum() for w, t in zip(ws_, target))
        first = first if first is not None else float(loss)
        opt.zero_grad()
        loss.backward()
        if which == "lamb" and not frozen():           # warm-up of 1-bit LAMB: the "engine" averages gradients
            for w in ws_:
                td.all_reduce(w.grad)
                w.grad /= world
        opt.step()
    assert frozen()
    errs = [float((w.detach() - t).abs().max()) for w, t in zip(ws_, target)]
    assert max(errs) < 0.3, errs
    for w in ws_:
        parts = [torch.zeros_like(w.data) for _ in range(world)]
        td.all_gather(parts, w.data)
        assert torch.allclose(parts[0], parts[1], atol=1e-5)
    sd = opt.state_dict()
    assert len(sd["state"]) == 2 and "exp_avg" in sd["state"][0]


@pytest.mark.parametrize("which", ["lamb", "zoadam"])
def test_onebit_lamb_and_zero_one_adam_converge(which):
    run_dist(_other_onebit_worker, 2, (which,))
