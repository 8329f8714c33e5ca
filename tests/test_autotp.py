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


def _meta_skeleton_worker(rank, world, root):
    """The large-model flow of the reference: skeleton on the meta device + checkpoint folder + tensor parallelism.
    Every TP layer is filled from the checkpoint and sliced by its own rule, replicated tensors and tied / recomputed
    buffers are materialised, and the logits equal the unsharded model's."""
    import os
    import pytest as _pt
    transformers = _pt.importorskip("transformers")
    import shuffle_exchange_b200 as sxb
    cfg = transformers.AutoConfig.for_model(model_type="llama", vocab_size=96, hidden_size=32, intermediate_size=48,
                                            num_hidden_layers=2, num_attention_heads=4, num_key_value_heads=2,
                                            max_position_embeddings=64)
    torch.manual_seed(0)
    full = transformers.AutoModelForCausalLM.from_config(cfg).float().eval()
    ids = torch.randint(0, 96, (2, 9), generator=torch.Generator().manual_seed(1))
    with torch.no_grad():
        ref = full(ids, use_cache=False).logits.clone()
    if rank == 0:
        torch.save(full.state_dict(), os.path.join(root, "pytorch_model.bin"))
    torch.distributed.barrier()
    with sxb.OnDevice(dtype=torch.float32, device="meta"):
        skeleton = transformers.AutoModelForCausalLM.from_config(cfg).eval()
    assert all(p.is_meta for p in skeleton.parameters())
    eng = sxb.init_inference(skeleton, dtype=torch.float32, tensor_parallel={"tp_size": world}, checkpoint=root)
    assert not any(p.is_meta for p in eng.module.parameters()) and not any(b.is_meta for b in eng.module.buffers())
    assert tuple(eng.module.model.layers[0].mlp.down_proj.weight.shape) == (32, 48 // world)      # row-parallel slice
    assert tuple(eng.module.model.layers[0].self_attn.q_proj.weight.shape) == (32 // world, 32)   # column-parallel
    with torch.no_grad():
        out = eng.module(ids, use_cache=False).logits
    assert torch.allclose(out, ref, atol=1e-5), float((out - ref).abs().max())
    # a checkpoint that does not cover the model is an error, not a crash inside .to(device)
    if rank == 0:
        sd = {k: v for k, v in full.state_dict().items() if "layers.1.mlp" not in k}
        os.makedirs(os.path.join(root, "partial"), exist_ok=True)
        torch.save(sd, os.path.join(root, "partial", "pytorch_model.bin"))
    torch.distributed.barrier()
    with sxb.OnDevice(dtype=torch.float32, device="meta"):
        skeleton = transformers.AutoModelForCausalLM.from_config(cfg).eval()
    try:
        sxb.init_inference(skeleton, dtype=torch.float32, tensor_parallel={"tp_size": world},
                           checkpoint=os.path.join(root, "partial"))
        raised = False
    except RuntimeError as e:
        raised = "no data" in str(e)
    assert raised


def test_meta_skeleton_plus_checkpoint_with_tensor_parallelism(tmp_path):
    run_dist(_meta_skeleton_worker, 2, (str(tmp_path),))
