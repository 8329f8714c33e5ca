"""state_dict_factory (Megatron TP merge/split), MoQ quantizer schedule, weight quantizer, autocast / compiler helpers."""
import os

import pytest
import torch


def _megatron_sd(h=16):
    torch.manual_seed(0)
    pre = "transformer.layers.0."
    return {pre + "attention.query_key_value.weight": torch.randn(3 * h, h),
            pre + "attention.query_key_value.bias": torch.randn(3 * h),
            pre + "attention.dense.weight": torch.randn(h, h), pre + "attention.dense.bias": torch.randn(h),
            pre + "mlp.dense_h_to_4h.weight": torch.randn(4 * h, h), pre + "mlp.dense_h_to_4h.bias": torch.randn(4 * h),
            pre + "mlp.dense_4h_to_h.weight": torch.randn(h, 4 * h), pre + "mlp.dense_4h_to_h.bias": torch.randn(h),
            "word_embeddings.weight": torch.randn(32, h)}


@pytest.mark.parametrize("version", [0, 2.0])
def test_sd_loader_split_then_merge_roundtrip(tmp_path, version):
    from shuffle_exchange_b200.runtime.state_dict_factory import SDLoaderFactory
    full = _megatron_sd()
    p = str(tmp_path / "full.pt")
    torch.save({"module": full, "checkpoint_version": version}, p)
    loader = SDLoaderFactory.get_sd_loader([p], None, "Megatron", version=version)
    files = []
    for r in range(2):
        path, sd, (scales, merge_count) = loader.load(2, r)
        assert path == p and scales is None and merge_count == 1
        assert sd["module"]["transformer.layers.0.attention.dense.weight"].shape == (16, 8)
        assert sd["module"]["transformer.layers.0.attention.query_key_value.weight"].shape == (24, 16)
        assert sd["module"]["transformer.layers.0.attention.dense.bias"].shape == (16,)      # replicated
        f = str(tmp_path / f"mp_rank_0{r}.pt")
        torch.save({"module": sd["module"], "checkpoint_version": version, "mp_world_size": 2}, f)
        files.append(f)
    merged_loader = SDLoaderFactory.get_sd_loader_json({"type": "Megatron", "checkpoints": files,
                                                        "version": version}, None)
    _, sd, (_, merge_count) = merged_loader.load(1, 0)
    assert merge_count == 2
    for k, v in full.items():
        assert torch.equal(sd["module"][k], v), k
    # same degree: own file, optionally quantised
    _, sd0, _ = merged_loader.load(2, 0)
    assert torch.equal(sd0["module"]["word_embeddings.weight"], full["word_embeddings.weight"][:16])
    _, sdq, (scales, _) = merged_loader.load(1, 0, quantize=True, quantize_groups=4)
    assert sdq["module"]["transformer.layers.0.mlp.dense_4h_to_h.weight"].dtype == torch.int8
    assert scales.shape[0] == 1 and scales.shape[1] == 4
    with pytest.raises(AssertionError):
        SDLoaderFactory.get_sd_loader(files, None, "Unknown")
    assert SDLoaderFactory.get_sd_loader_json({"type": "bloom", "checkpoints": [], "version": 1}, None)["type"] == "bloom"


def test_weight_quantization_error_bound_and_model_quantize():
    from shuffle_exchange_b200.runtime.weight_quantizer import WeightQuantization
    torch.manual_seed(0)
    wq = WeightQuantization(mlp_extra_grouping=True, mp_size=1)
    w = torch.randn(64, 16)
    q, s = wq.quantize_data(w, 8, 4)
    assert q.dtype == torch.int8 and s.shape == (4, 1)
    deq = (q.float().reshape(4, -1) / s).reshape(w.shape)
    assert (deq - w).abs().max() <= (1.0 / s.min()).item() + 1e-6      # the positive extreme clamps 128 -> 127
    assert wq.is_mlp(torch.empty(64, 16)) and wq.is_qkv(torch.empty(48, 16)) and not wq.is_qkv(torch.empty(16, 16))

    class BertLayer(torch.nn.Module):            # matched by name through the built-in policy list
        def __init__(self):
            super().__init__()
            self.qkv = torch.nn.Linear(16, 48)
            self.out = torch.nn.Linear(16, 16)
            self.up = torch.nn.Linear(16, 64)
            self.down = torch.nn.Linear(64, 16)

    model = torch.nn.Sequential(BertLayer(), BertLayer())
    _, scales = WeightQuantization().model_quantize(model, None, 8, 2)
    assert scales.shape[0] == 2 and scales.shape[1] == 4
    assert float(model[0].qkv.weight.abs().max()) <= 128 and float(model[0].qkv.weight.frac().abs().max()) == 0.0


def test_moq_schedule_reduces_precision_to_target():
    from shuffle_exchange_b200.runtime.quantize import Quantizer
    torch.manual_seed(0)
    q = Quantizer(q_groups=4, q_type="symmetric", q_rounding="nearest")
    w = torch.nn.Parameter(torch.randn(8, 16))
    w.start_bits, w.target_bits, w.q_period = 8, 4, 2
    seen = set()
    for _ in range(40):
        q.quantize([[w]], overflow=False, eigenvalue_enabled=False)
        seen.add(w.start_bits)
    assert w.start_bits == 4 and seen == {8, 7, 6, 5, 4}
    assert w.data.reshape(4, -1)[0].unique().numel() <= 16
    q.quantize([[w]], overflow=True, eigenvalue_enabled=False)      # skipped on overflow
    assert q.qsteps == 40
    t = Quantizer(q_groups=1, q_type="symmetric", q_rounding="nearest")
    x = torch.randn(4, 8)
    assert t.quantize_binary(x).abs().unique().numel() == 1
    assert t.quantize_tenary(x).unique().numel() <= 3
    mixed = Quantizer(q_groups=1, q_mixed_fp16=True, q_change_ratio=0.5, q_type="asymmetric", q_rounding="nearest")
    mixed.update_fp16_ratio()
    mixed.update_fp16_ratio()
    mixed.update_fp16_ratio()
    assert mixed.quantize_real_ratio == 0.0


def test_autocast_and_compiler_helpers():
    from shuffle_exchange_b200.runtime import compiler, torch_autocast as ta

    class Eng:
        module = torch.nn.Sequential(torch.nn.Linear(4, 4), torch.nn.LayerNorm(4))
        communication_data_type = torch.float32
        def fp16_enabled(self): return False
        def bfloat16_enabled(self): return False
        def torch_autocast_enabled(self): return True

    eng = Eng()
    ta.init_autocast_params(eng, torch.bfloat16, None)
    assert ta.is_autocast_initialized()
    lin, ln = eng.module[0], eng.module[1]
    assert ta.get_autocast_dtype(lin.weight) == torch.bfloat16 and not ta.has_autocast_dtype(ln.weight)
    assert ta.get_all_autocast_dtypes(eng.module.parameters()) == {torch.bfloat16, torch.float32}
    assert ta.sort_dtypes([torch.float32, torch.bfloat16]) == [torch.bfloat16, torch.float32]
    with pytest.raises(ValueError):
        ta.init_autocast_params(eng, torch.bfloat16, ["not.a.module.Class"])
    with torch.autocast("cpu", dtype=torch.bfloat16):
        ta.validate_nested_autocast(eng)            # allowed (warns once)
        eng.torch_autocast_enabled = lambda: False
        with pytest.raises(AssertionError):
            ta.validate_nested_autocast(eng)

    @compiler.enable
    def f(x):
        return x + 1

    @compiler.enable("99.0.0")
    def g(x):
        return x + 2

    assert f(1) == 2 and g(1) == 3 and compiler.is_compile_supported() and not compiler.is_compiling()
    assert compiler.disable(f)(2) == 3


def test_inference_post_init_weight_quantization():
    from torch import nn
    from shuffle_exchange_b200.inference.quantization import (DeQuantizer, QuantizedEmbedding, QuantizedLinear,
                                                              Quantizer, _init_group_wise_weight_quantization)
    torch.manual_seed(0)

    class M(nn.Module):
        def __init__(self):
            super().__init__()
            self.embed = nn.Embedding(64, 32)
            self.fc1 = nn.Linear(32, 64)
            self.fc2 = nn.Linear(64, 32)
            self.head = nn.Linear(32, 64, bias=False)

        def forward(self, ids):
            return self.head(self.fc2(torch.relu(self.fc1(self.embed(ids)))))

    m = M().to(torch.bfloat16)
    ids = torch.randint(0, 64, (2, 5))
    ref = m(ids).float()
    q8 = {"num_bits": 8, "group_size": 16, "group_dim": 1, "symmetric": False}
    q4 = {"num_bits": 4, "group_size": 16, "group_dim": 1, "symmetric": False}
    m = _init_group_wise_weight_quantization(m, {"weight_quantization": {"post_init_quant": {"fc": q8, "embed": q4}}})
    assert isinstance(m.fc1, QuantizedLinear) and isinstance(m.embed, QuantizedEmbedding)
    assert type(m.head) is nn.Linear                                  # no key matched
    assert m.fc1.weight.dtype == torch.uint8 and m.embed.weight.numel() < 64 * 32
    assert (m(ids).float() - ref).abs().max() < 0.1
    w = torch.randn(4, 16)
    qz = Quantizer({"num_bits": 4, "group_size": 8, "group_dim": 1, "symmetric": False})
    codes, scale, mn = qz.quantize(w)
    assert codes.shape == (4, 8) and codes.dtype == torch.uint8
    back = DeQuantizer(qz.config, torch.float32).dequantize(codes, scale, mn)
    step = (w.reshape(4, 2, 8).amax(-1) - w.reshape(4, 2, 8).amin(-1)) / 15
    assert (back - w).abs().max() <= step.max() / 2 + 1e-5
    with pytest.raises(AssertionError):
        Quantizer({"num_bits": 3, "group_size": 8, "group_dim": 1, "symmetric": False})
    with pytest.raises(AssertionError):
        _init_group_wise_weight_quantization(M().half(), {"weight_quantization": {"post_init_quant": {"fc": q8, "c1": q8}}})


def test_contiguous_allocator_randomized_workload_and_param_rebinding():
    from shuffle_exchange_b200.runtime.zero import test as alloc_check
    stats = alloc_check.random_workload(steps=600, size=2048, seed=3)
    assert stats["allocations"] > 100 and stats["defragmentations"] > 0 and stats["max_allocated"] <= 2048
    alloc_check.param_rebinding()


def test_fused_qkv_attention_entry_point_matches_split_path_on_cpu():
    """``flash_attention_qkv`` (one fused tensor in, fused gradient out) is the same function as slicing q, k, v and
    calling ``flash_attention`` — forward and gradient — for GQA and MHA layouts."""
    import torch
    from shuffle_exchange_b200.ops import functional as F
    for (B, S, hq, hk, d) in ((2, 9, 4, 2, 8), (1, 5, 3, 3, 4)):
        torch.manual_seed(0)
        qkv = torch.randn(B, S, (hq + 2 * hk) * d, requires_grad=True)
        out = F.flash_attention_qkv(qkv, hq, hk, d, causal=True)
        g = torch.randn_like(out)
        out.backward(g)
        ref_in = qkv.detach().clone().requires_grad_()
        q, k, v = F._split_qkv(ref_in, hq, hk, d)
        ref = F.flash_attention(q, k, v, causal=True)
        ref.backward(g)
        assert out.shape == (B, S, hq, d)
        assert torch.allclose(out, ref, atol=1e-6) and torch.allclose(qkv.grad, ref_in.grad, atol=1e-6)


def _engine_helpers(rank, world):
    import torch
    import shuffle_exchange_b200 as sxb
    from shuffle_exchange_b200.runtime.sparse_tensor import SparseTensor
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Embedding(16, 4), torch.nn.Linear(4, 4))
    engine, *_ = sxb.initialize(model=model, model_parameters=model.parameters(), config={
        "train_micro_batch_size_per_gpu": 2, "steps_per_print": 1 << 30,
        "optimizer": {"type": "Adam", "params": {"lr": 1e-3}},
        "autotuning": {"enabled": True, "start_profile_step": 4, "model_info": {"profile": True}},
        "compression_training": {"weight_quantization": {"shared_parameters": {"enabled": True, "quantize_groups": 4},
                                                         "different_groups": {"g": {"params": {"target_bits": 4}}}}}})
    assert engine.autotuning_start_profile_step() == 4 and engine.autotuning_end_profile_step() == 5
    assert engine.autotuning_profile_model_info() and engine.autotuning_metric_path().endswith("autotuning_metric.json")
    q = engine.quantize_training()
    assert len(q) == 13 and q[1] is True and q[2] == 4 and q[6] == 4
    assert engine.is_map_style_dataset([1, 2]) and not engine.is_iterable_style_dataset([1, 2])
    assert engine.get_pld_theta() is None and not engine.checkpoint_writer_enabled() and engine.autotp_size() in (0, 1, None)
    # rows touched differ per rank: rank r has rows {r, 5}
    dense = torch.zeros(8, 3)
    dense[rank] = 1.0 + rank
    dense[5] = 10.0
    sp = SparseTensor(dense)
    group = torch.distributed.group.WORLD
    out = engine.sparse_allreduce(sp, group)
    ref = dense.clone()
    torch.distributed.all_reduce(ref)
    assert torch.allclose(out.to_dense(), ref / world)
    gathered = engine.sparse_all_gather(torch.arange(rank + 1, dtype=torch.float32), group)
    assert [t.numel() for t in gathered] == [1, 2]
    seen = []
    engine.register_compile_pass("noop_for_test", lambda plan, ctx: seen.append(1) or plan)
    from shuffle_exchange_b200.compile.backend import opt_passes
    assert "noop_for_test" in opt_passes and engine.get_compile_time() == 0.0


def test_engine_helper_methods_two_ranks():
    from conftest import run_dist
    run_dist(_engine_helpers, world=2)


def test_autotuner_micro_batch_size_search_with_a_fake_runner():
    from types import SimpleNamespace
    from shuffle_exchange_b200.autotuning.autotuner import Autotuner
    from shuffle_exchange_b200.autotuning import constants as C
    t = Autotuner.__new__(Autotuner)
    t.user_config = {"gradient_accumulation_steps": "auto", "zero_optimization": {"stage": 1}}
    t.autotuning_config = SimpleNamespace(arg_mappings={"gradient_accumulation_steps": "--gas"}, mp_size=1)
    t.args = SimpleNamespace(user_args=["--lr", "1", "--gas", "2"])
    t.exp_num_gpus, t.exp_num_nodes, t.records = 2, 1, {}
    assert t.get_val_from_user_args("gradient_accumulation_steps") == "2" and t.get_gas_from_user_config() == 2
    calls = []

    def fake_run(ds_config, name):                     # runs up to mbs 11; throughput saturates around mbs 8
        mbs = ds_config[C.TRAIN_MICRO_BATCH_SIZE_PER_GPU]
        calls.append(mbs)
        assert ds_config[C.TRAIN_BATCH_SIZE] == mbs * ds_config[C.GRADIENT_ACCUMULATION_STEPS] * 2
        val = None if mbs > 11 else 100.0 * min(mbs, 8) + mbs
        return {"name": name, "ds_config": ds_config}, val

    t.run_ds_config = fake_run
    assert t.get_min_max_micro_batch_size(1, 1, 32) == (1, 11)            # estimate too high: bisection
    assert max(calls) == 32 and len(calls) <= 8
    t.records.clear(); calls.clear()
    assert t.get_min_max_micro_batch_size(1, 1, 2) == (1, 8)              # estimate too low: doubling until 16 fails
    assert t.get_min_max_micro_batch_size(1, 64, 64) == (-1, -1)
    t.records.clear()
    assert t.run_tuning_micro_batch_sizes([2, 4, 8, 10, 16], None, 1, stage=1) == 10
    name = f"{C.TUNING_MICRO_BATCH_SIZE_PREFIX}1"
    assert t.get_plateau_mbs(name) == 8
    assert t.run_tuning_micro_batch_sizes([4, 8], 6, 1, stage=1) == 4     # batch-size cap


def test_runtime_utils_reference_helpers(tmp_path):
    import torch
    from shuffle_exchange_b200.runtime import utils as U
    nested = {"a": [torch.ones(2), (torch.zeros(3, dtype=torch.float16), 5)], "b": "x"}
    assert U.get_tensor_bytes(nested) == 2 * 4 + 3 * 2
    c = U.copy_to_device(nested, "cpu", torch.is_tensor)
    assert U.compare_tensors_in_structures(nested, c) and c["a"][0] is not None
    c["a"][0][0] = 7
    assert U.compare_tensors_in_structures(nested, c) == bool(nested["a"][0][0] == 7)
    assert not U.compare_tensors_in_structures([1], (1,))
    t = torch.tensor([1.0, float("inf"), float("nan")])
    assert U.mask_nan_or_inf_with_val_inplace(t, val=-1.0).tolist() == [1.0, -1.0, -1.0]
    lin = torch.nn.Linear(4, 3)
    tl = U.TLinear(lin)
    x = torch.randn(2, 4)
    assert tl.weight.shape == (4, 3) and torch.allclose(tl(x), lin(x), atol=1e-6)
    p = torch.nn.Parameter(torch.tensor([3.0, 0.0, 4.0]))
    p.grad = torch.tensor([3.0, 0.0, 4.0])
    assert abs(U.get_flattened_grad_norm([p]) - 5.0) < 1e-6 and U.get_grad_zeros([p]) == 1.0
    assert abs(U.get_flattened_grad_norm([p], grad_norm_mask=[torch.tensor([[0, 1]])]) - 4.0) < 1e-6
    assert abs(U.get_weight_norm([p]) - 5.0) < 1e-6
    assert U.get_norm_with_moe_layers(3.0, None, {"ep": [torch.tensor([4.0])]}) == 5.0
    assert U.get_norm_with_moe_layers(float("inf"), None, {}) == -1
    (tmp_path / "tag").mkdir(); (tmp_path / "tag" / "f.bin").write_bytes(b"x" * 2048)
    assert abs(U.get_checkpoint_folder_size(str(tmp_path), "tag") - 2048 / 1024 ** 3) < 1e-12
    assert U.get_checkpoint_folder_size(str(tmp_path), "tag", local_rank=1) == 0.0
    assert U.DummyOptim([p]).param_groups[0]["params"] == [p] and U.get_ma_status() == 0
    with U.noop_context(lambda: 3) as v:
        assert v == 3
    assert U.get_inactive_params([p]) == []


def _compress_tp(rank, world):
    import torch
    import torch.distributed as td
    from shuffle_exchange_b200.compression.basic_layer import ColumnParallelLinear_Compress, RowParallelLinear_Compress

    class MPU:
        def get_model_parallel_world_size(self): return world
        def get_model_parallel_rank(self): return rank
        def get_model_parallel_group(self): return td.group.WORLD

    torch.manual_seed(0)
    full1, full2 = torch.nn.Linear(8, 12), torch.nn.Linear(12, 8)
    col = ColumnParallelLinear_Compress(MPU(), 8, 12, gather_output=False)
    row = RowParallelLinear_Compress(MPU(), 12, 8, input_is_parallel=True)
    n = 12 // world
    with torch.no_grad():
        col.weight.copy_(full1.weight[rank * n:(rank + 1) * n]); col.bias.copy_(full1.bias[rank * n:(rank + 1) * n])
        row.weight.copy_(full2.weight[:, rank * n:(rank + 1) * n]); row.bias.copy_(full2.bias)
    x = torch.randn(3, 8, requires_grad=True)
    xr = x.detach().clone().requires_grad_()
    h, _ = col(x)
    y, _ = row(torch.relu(h))
    ref = full2(torch.relu(full1(xr)))
    assert torch.allclose(y, ref, atol=1e-5)
    y.sum().backward(); ref.sum().backward()
    assert torch.allclose(x.grad, xr.grad, atol=1e-5)                  # input grad all-reduced over the group
    assert torch.allclose(col.weight.grad, full1.weight.grad[rank * n:(rank + 1) * n], atol=1e-5)
    col.enable_weight_quantization(8, 8, 1, True, "symmetric", 1)       # compression methods still apply
    assert col(x.detach())[0].shape == (3, n)


def test_compression_tensor_parallel_layers_match_dense():
    from conftest import run_dist
    run_dist(_compress_tp, world=2)
