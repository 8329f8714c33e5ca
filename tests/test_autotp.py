"""AutoTP on 2 gloo ranks: a sharded HF-style block must reproduce the unsharded forward and (training mode) give
gradients that match the corresponding slices of the full model's gradients."""
import torch
import torch.nn as nn

from conftest import run_dist


class Attn(nn.Module):
    def __init__(self, h, heads):
        super().__init__()
        self.num_heads, self.head_dim, self.hidden_size = heads, h // heads, h
        self.q_proj, self.k_proj, self.v_proj, self.o_proj = (nn.Linear(h, h, bias=False) for _ in range(4))

    def forward(self, x):
        B, S, _ = x.shape
        q = self.q_proj(x).view(B, S, self.num_heads, self.head_dim).transpose(1, 2)
        k = self.k_proj(x).view(B, S, self.num_heads, self.head_dim).transpose(1, 2)
        v = self.v_proj(x).view(B, S, self.num_heads, self.head_dim).transpose(1, 2)
        o = torch.nn.functional.scaled_dot_product_attention(q, k, v, is_causal=True)
        return self.o_proj(o.transpose(1, 2).reshape(B, S, self.num_heads * self.head_dim))


class MLP(nn.Module):
    def __init__(self, h):
        super().__init__()
        self.gate_proj, self.up_proj, self.down_proj = nn.Linear(h, 4 * h, bias=False), nn.Linear(h, 4 * h, bias=False), nn.Linear(4 * h, h, bias=True)

    def forward(self, x):
        return self.down_proj(torch.nn.functional.silu(self.gate_proj(x)) * self.up_proj(x))


class Block(nn.Module):
    def __init__(self, h, heads):
        super().__init__()
        self.self_attn, self.mlp = Attn(h, heads), MLP(h)
        self.ln1, self.ln2 = nn.LayerNorm(h), nn.LayerNorm(h)

    def forward(self, x):
        x = x + self.self_attn(self.ln1(x))
        return x + self.mlp(self.ln2(x))


class Tiny(nn.Module):
    def __init__(self, h=32, heads=4, layers=2):
        super().__init__()
        self.layers = nn.ModuleList([Block(h, heads) for _ in range(layers)])

    def forward(self, x):
        for l in self.layers:
            x = l(x)
        return x


def test_tp_parser_finds_row_parallel_linears():
    from shuffle_exchange_b200.module_inject import AutoTP
    (cls, gem), = AutoTP.tp_parser(Tiny())
    assert cls is Block and set(gem) == {"o_proj", "down_proj"}


def _worker(rank, world, overlap=False):
    import copy
    import types
    from shuffle_exchange_b200.module_inject import tp_model_init
    from shuffle_exchange_b200.module_inject import layers as L
    from shuffle_exchange_b200.module_inject.layers import LinearAllreduce, LinearLayer
    # ``tensor_parallel.tp_overlap_comm``: column-parallel layers take the asynchronous-all-reduce backward
    L.configure_tensor_parallel_runtime(types.SimpleNamespace(tp_grain_size=1, tp_overlap_comm=overlap))
    assert L.TensorParallel_Layer.tp_overlap_comm is overlap
    taken = []
    if overlap:
        orig = L.AsyncColumnParallel.forward
        L.AsyncColumnParallel.forward = staticmethod(lambda ctx, *a: (taken.append(1), orig(ctx, *a))[1])
    torch.manual_seed(0)
    full = Tiny()
    x = torch.randn(2, 6, 32)
    ref = full(x)
    ref.sum().backward()
    tp = tp_model_init(copy.deepcopy(full).requires_grad_(True), tp_size=2, dtype=None)
    for p in tp.parameters():
        p.grad = None
    assert isinstance(tp.layers[0].self_attn.q_proj, LinearLayer) and isinstance(tp.layers[0].mlp.down_proj, LinearAllreduce)
    assert tp.layers[0].self_attn.num_heads == 2
    out = tp(x)
    assert torch.allclose(out, ref, atol=1e-5), (out - ref).abs().max()
    out.sum().backward()
    g_full = full.layers[1].mlp.gate_proj.weight.grad
    g_tp = tp.layers[1].mlp.gate_proj.weight.grad
    assert torch.allclose(g_tp, g_full.chunk(2, 0)[rank], atol=1e-5)
    g_full = full.layers[0].self_attn.o_proj.weight.grad
    g_tp = tp.layers[0].self_attn.o_proj.weight.grad
    assert torch.allclose(g_tp, g_full.chunk(2, 1)[rank], atol=1e-5)
    if overlap:
        assert len(taken) >= 4, "tp_overlap_comm did not route the column-parallel layers through AsyncColumnParallel"
        L.configure_tensor_parallel_runtime(types.SimpleNamespace(tp_grain_size=1, tp_overlap_comm=False))


def test_autotp_matches_unsharded():
    run_dist(_worker, 2)


def test_autotp_with_overlapped_input_gradient_all_reduce():
    run_dist(_worker, 2, (True,))
