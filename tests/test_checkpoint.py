"""Checkpoint layout, resume, DP re-partitioning and offline fp32 consolidation (gloo)."""
import os

import pytest
import torch

from conftest import run_dist


def _cfg(stage):
    return {"train_micro_batch_size_per_gpu": 2, "steps_per_print": 10 ** 9, "bf16": {"enabled": True},
            "optimizer": {"type": "AdamW", "params": {"lr": 5e-3}},
            "scheduler": {"type": "WarmupLR", "params": {"warmup_num_steps": 10, "warmup_max_lr": 5e-3}},
            "zero_optimization": {"stage": stage}}


def _make(stage, world):
    import shuffle_exchange_b200 as ds
    from shuffle_exchange_b200.models import GPT2LMHeadModel, gpt2_config
    torch.manual_seed(0)
    model = GPT2LMHeadModel(gpt2_config("gpt2-tiny"))
    return ds.initialize(model=model, config=_cfg(stage), model_parameters=model.parameters(), slice_count=world)[0]


def _save_load_worker(rank, world, stage, tmp):
    from shuffle_exchange_b200.utils.tensor_fragment import safe_get_full_fp32_param
    e1 = _make(stage, world)
    ids = torch.randint(0, 512, (2, 16), generator=torch.Generator().manual_seed(rank))
    for _ in range(3):
        e1.backward(e1(ids, labels=ids))
        e1.step()
    e1.save_checkpoint(tmp, client_state={"note": "hi"})
    tag = "global_step3"
    files = set(os.listdir(os.path.join(tmp, tag)))
    assert open(os.path.join(tmp, "latest")).read() == tag
    if stage == 3:
        assert {f"zero_pp_rank_{r}_mp_rank_00_model_states.pt" for r in range(world)} <= files
    else:
        assert "mp_rank_00_model_states.pt" in files
    assert {f"bf16_zero_pp_rank_{r}_mp_rank_00_optim_states.pt" for r in range(world)} <= files
    assert "zero_to_fp32.py" in files
    next_loss = float(e1(ids, labels=ids))
    e2 = _make(stage, world)
    path, client = e2.load_checkpoint(tmp)
    assert path is not None and client["note"] == "hi"
    assert e2.global_steps == 3 and e2.lr_scheduler.last_batch_iteration == e1.lr_scheduler.last_batch_iteration
    assert abs(float(e2(ids, labels=ids)) - next_loss) < 1e-4
    for (n, p1), (_, p2) in zip(e1.module.named_parameters(), e2.module.named_parameters()):
        assert torch.equal(safe_get_full_fp32_param(p1), safe_get_full_fp32_param(p2)), n
    # optimizer state survived: one more identical step gives identical weights
    for e in (e1, e2):
        e.backward(e(ids, labels=ids))
        e.step()
    p1, p2 = next(e1.module.parameters()), next(e2.module.parameters())
    assert torch.allclose(safe_get_full_fp32_param(p1), safe_get_full_fp32_param(p2), atol=1e-7)
    if rank == 0:
        from shuffle_exchange_b200.utils.zero_to_fp32 import get_fp32_state_dict_from_zero_checkpoint
        sd = get_fp32_state_dict_from_zero_checkpoint(tmp)
        e3 = _make(stage, world) if False else None
        assert set(sd.keys()) >= {n for n, _ in e1.module.named_parameters()}
        # sharded + indexed output (HuggingFace layout), single-file output, safetensors
        import json
        from shuffle_exchange_b200.utils.zero_to_fp32 import _parse_size, convert_zero_checkpoint_to_fp32_state_dict
        assert _parse_size("5GB") == 5 * 1000 ** 3 and _parse_size("2MiB") == 2 * 1024 ** 2 and _parse_size(7) == 7
        biggest = max(v.numel() * v.element_size() for v in sd.values())
        out = os.path.join(tmp, "fp32_sharded")
        convert_zero_checkpoint_to_fp32_state_dict(tmp, out, max_shard_size=biggest)
        index = json.load(open(os.path.join(out, "pytorch_model.bin.index.json")))
        assert set(index["weight_map"]) == set(sd) and len(set(index["weight_map"].values())) > 1
        merged = {}
        for f in set(index["weight_map"].values()):
            merged.update(torch.load(os.path.join(out, f), weights_only=True))
        assert all(torch.equal(merged[k], sd[k]) for k in sd)
        one = os.path.join(tmp, "fp32_one")
        convert_zero_checkpoint_to_fp32_state_dict(tmp, one)
        assert os.path.exists(os.path.join(one, "pytorch_model.bin"))
        single = os.path.join(tmp, "all.pt")
        convert_zero_checkpoint_to_fp32_state_dict(tmp, single)
        assert set(torch.load(single, weights_only=True)) == set(sd)
        try:
            import safetensors.torch as st
            convert_zero_checkpoint_to_fp32_state_dict(tmp, os.path.join(tmp, "fp32_st"), safe_serialization=True)
            loaded = st.load_file(os.path.join(tmp, "fp32_st", "model.safetensors"))
            assert all(torch.equal(loaded[k], sd[k]) for k in sd)
        except ImportError:
            pass


@pytest.mark.parametrize("stage", [1, 2, 3])
def test_save_load_roundtrip(stage, tmp_path):
    run_dist(_save_load_worker, 2, (stage, str(tmp_path)))


def _consolidate_worker(rank, world, stage, tmp):
    from shuffle_exchange_b200.utils.tensor_fragment import safe_get_full_fp32_param
    from shuffle_exchange_b200.utils.zero_to_fp32 import get_fp32_state_dict_from_zero_checkpoint
    e = _make(stage, world)
    ids = torch.randint(0, 512, (2, 16))
    e.backward(e(ids, labels=ids))
    e.step()
    e.save_checkpoint(tmp, tag="t")
    full = {n: safe_get_full_fp32_param(p).clone() for n, p in e.module.named_parameters()}
    torch.distributed.barrier()
    if rank == 0:
        sd = get_fp32_state_dict_from_zero_checkpoint(tmp, tag="t")
        for n, v in full.items():
            assert torch.equal(sd[n], v), n


@pytest.mark.parametrize("stage", [2, 3])
def test_zero_to_fp32_consolidation(stage, tmp_path):
    run_dist(_consolidate_worker, 2, (stage, str(tmp_path)))


def _save_worker(rank, world, tmp):
    e = _make(2, world)
    ids = torch.randint(0, 512, (2, 16))
    for _ in range(2):
        e.backward(e(ids, labels=ids))
        e.step()
    e.save_checkpoint(tmp, tag="dp2")


def _load_other_dp_worker(rank, world, tmp):
    from shuffle_exchange_b200.utils.tensor_fragment import safe_get_full_fp32_param, safe_get_full_optimizer_state
    from shuffle_exchange_b200.utils.zero_to_fp32 import get_fp32_state_dict_from_zero_checkpoint
    e = _make(2, world)
    path, _ = e.load_checkpoint(tmp, tag="dp2")
    assert path is not None
    sd = get_fp32_state_dict_from_zero_checkpoint(tmp, tag="dp2")
    for n, p in e.module.named_parameters():
        assert torch.equal(safe_get_full_fp32_param(p).cpu(), sd[n]), n
        assert safe_get_full_optimizer_state(p, "exp_avg").abs().sum() > 0


def test_elastic_repartition_dp2_to_dp1(tmp_path):
    run_dist(_save_worker, 2, (str(tmp_path),))
    run_dist(_load_other_dp_worker, 1, (str(tmp_path),))


def _universal_worker(rank, world, tmp):
    """Save under ZeRO-2 dp=2, convert to universal, reload under ZeRO-1 in a fresh engine: masters and Adam moments of
    every parameter must come back bit-identical."""
    import os
    import torch
    import shuffle_exchange_b200 as sxb
    from shuffle_exchange_b200.checkpoint.ds_to_universal import convert
    torch.manual_seed(0)

    def make(stage, universal=False):
        torch.manual_seed(0)
        m = torch.nn.Sequential(torch.nn.Linear(16, 32), torch.nn.Tanh(), torch.nn.Linear(32, 8))
        cfg = {"train_micro_batch_size_per_gpu": 2, "bf16": {"enabled": True},
               "optimizer": {"type": "Adam", "params": {"lr": 1e-2}}, "zero_optimization": {"stage": stage},
               "checkpoint": {"load_universal": universal}}
        e, _, _, _ = sxb.initialize(model=m, config=cfg, model_parameters=m.parameters())
        return e

    e = make(2)
    g = torch.Generator().manual_seed(5 + rank)
    for _ in range(3):
        x = torch.randn(2, 16, generator=g).bfloat16()
        loss = e(x).float().pow(2).mean()
        e.backward(loss)
        e.step()
    e.save_checkpoint(tmp, tag="t3")
    import torch.distributed as td
    td.barrier()
    if rank == 0:
        convert(os.path.join(tmp, "t3"), os.path.join(tmp, "t3_universal"))
    td.barrier()
    ref = {}
    for fg in e.optimizer.groups:
        st = e.optimizer._optimizer_state_for(fg)
        ref[tuple(fg.names)] = (fg.hp_param.data.clone(), st["exp_avg"].clone(), st["exp_avg_sq"].clone())
    e2 = make(1, universal=True)
    path, _ = e2.load_checkpoint(tmp, tag="t3_universal")
    assert path is not None
    for fg in e2.optimizer.groups:
        hp, m1, m2 = ref[tuple(fg.names)]
        st = e2.optimizer._optimizer_state_for(fg)
        valid = min(fg.part, max(0, sum(fg.numels) + fg.offsets[0] - fg.rank * fg.part))
        assert torch.equal(fg.hp_param.data[:valid], hp[:valid])
        assert torch.equal(st["exp_avg"][:valid], m1[:valid]) and torch.equal(st["exp_avg_sq"][:valid], m2[:valid])
    assert e2.global_steps == 3


def test_universal_checkpoint_roundtrip(tmp_path):
    from conftest import run_dist
    run_dist(_universal_worker, 2, (str(tmp_path),))


def test_2d_and_3d_layout_reshape_and_zero_checkpoint_merge(tmp_path):
    import torch
    from shuffle_exchange_b200.checkpoint import (ZeROCheckpoint, get_model_3d_descriptor, get_mpu_ranks,
                                                  model_3d_desc, reshape_meg_2d_parallel)
    m = reshape_meg_2d_parallel(old_pp_degree=2, old_tp_degree=4, new_pp_degree=1, new_tp_degree=2)
    assert m.get_data(0, 0) == [0, 1, 4, 5] and m.get_data(0, 1) == [2, 3, 6, 7]
    tp, pp, dp = get_mpu_ranks(tp_size=2, pp_size=2, dp_size=2)
    assert tp == [[0, 1], [2, 3], [4, 5], [6, 7]] and pp == [[0, 4], [1, 5], [2, 6], [3, 7]]
    assert dp == [[0, 2], [1, 3], [4, 6], [5, 7]]
    src = model_3d_desc(pp_degree=1, tp_degree=2, dp_degree=4)
    ok, errs = src.can_reshape(model_3d_desc(1, 4, 1))
    assert not ok and "TP" in errs[0]
    maps = src.reshape(model_3d_desc(1, 1, 2))
    assert len(maps) == 2 and maps[0].get_data(0, 0) == [0, 1, 2, 3] and maps[1].get_data(0, 0) == [4, 5, 6, 7]
    # a folder with tp=1, pp=1, dp=4 ZeRO files: merge to dp=2 concatenates the flat partitions without their padding
    for r in range(4):
        torch.save({"module": {}}, tmp_path / "mp_rank_00_model_states.pt")
        sd = {"optimizer_state_dict": {"base_optimizer_state": {"state": {0: {"exp_avg": torch.full((6,), float(r))}}},
                                       "group_paddings": [2], "partition_count": [4]}}
        torch.save(sd, tmp_path / f"zero_pp_rank_{r}_mp_rank_00_optim_states.pt")
    d = get_model_3d_descriptor(str(tmp_path))
    assert (d.pp_degree, d.tp_degree, d.dp_degree) == (1, 1, 4)
    z = ZeROCheckpoint(str(tmp_path))
    assert z.get_src_dp_degree() == 4 and len(z.get_files_for_rank(0, 0, 3)) == 1
    z.reshape(model_3d_desc(1, 1, 2))
    assert [f.split("zero_pp_rank_")[1][0] for f in z.get_files_for_rank(0, 0, 1)] == ["2", "3"]
    st = z.get_state_for_rank(0, 0, 1)
    ea = st["optimizer_state_dict"]["base_optimizer_state"]["state"][0]["exp_avg"]
    assert ea.tolist() == [2.0] * 4 + [3.0] * 4
    assert st["optimizer_state_dict"]["partition_count"] == [2] and st["optimizer_state_dict"]["group_paddings"] == [0]


def test_deepspeed_checkpoint_inspector_maps_and_merges(tmp_path, capsys):
    """A Megatron-DeepSpeed style folder (tp=2, pp=2, dp=2; embedding + 4 blocks + final norm) read under its own
    layout and under tp=1 / pp=1: file tables, TP merges along the right dimension, mp_rank and ZeRO file selection."""
    import torch
    from shuffle_exchange_b200.checkpoint import DeepSpeedCheckpoint
    H = 4
    for layer in range(6):                                  # 00 embedding, 01-04 blocks, 05 final norm
        for tp in range(2):
            if layer == 0:
                sd = {"word_embeddings.weight": torch.full((3, H), float(tp))}
            elif layer == 5:
                sd = {"weight": torch.full((H,), 7.0)}
            else:
                sd = {"input_layernorm.weight": torch.full((H,), float(layer)),
                      "self_attention.dense.weight": torch.full((H, 2), float(10 * layer + tp)),
                      "mlp.dense_h_to_4h.weight": torch.full((2, H), float(10 * layer + tp))}
            torch.save(sd, tmp_path / f"layer_{layer:02d}-model_{tp:02d}-model_states.pt")
    for mp in range(4):
        torch.save({"iteration": 11, "args": {"hidden": H}, "checkpoint_info": {"mp": mp},
                    "module": None}, tmp_path / f"mp_rank_{mp:02d}_model_states.pt")
        for dp in range(2):
            torch.save({"optimizer_state_dict": {"base_optimizer_state": {"state": {}}, "group_paddings": [0],
                                                 "partition_count": [2]}},
                       tmp_path / f"zero_pp_rank_{dp}_mp_rank_{mp:02d}_optim_states.pt")

    c = DeepSpeedCheckpoint(str(tmp_path))
    assert (c.original_tp_degree, c.original_pp_degree, c.original_dp_degree) == (2, 2, 2)
    assert not (c.is_change_tp_degree() or c.is_change_pp_degree() or c.is_change_dp_degree())
    assert c.get_iteration() == 11 and c.get_args() == {"hidden": H} and c.get_checkpoint_info() == {"mp": 0}
    assert c.get_embedding_layer_id() == "00" and c.get_final_norm_layer_id() == "05"
    assert c.get_pp_transformer_map(0) == ["01", "02"] and c.get_pp_transformer_map(1) == ["03", "04"]
    assert [len(v) for v in c.tp_to_embedding_map.values()] == [1, 1]
    blocks = c.get_transformer_state(tp_index=1, pp_index=1)
    assert len(blocks) == 2 and float(blocks[0]["self_attention.dense.weight"][0, 0]) == 31.0
    assert [os.path.basename(f) for f in c.get_2d_parallel_files(tp_index=1, pp_index=1)] == ["mp_rank_03_model_states.pt"]
    assert len(c.get_zero_files(pp_index=1, tp_index=0, dp_index=1)) == 1 and len(c.get_zero_files()) == 8

    m = DeepSpeedCheckpoint(str(tmp_path), tp_degree=1, pp_degree=1, dp_degree=1)
    assert m.is_change_tp_degree() and m.is_change_pp_degree() and m.is_change_dp_degree() and m.world_size == 1
    assert len(m.get_embedding_files(0)) == 2 and m.get_embedding_state(0)["word_embeddings.weight"].shape == (6, H)
    blocks = m.get_transformer_state(0, 0)
    assert len(blocks) == 4
    b = blocks[2]                                                        # layer 03, both TP slices merged
    assert b["input_layernorm.weight"].shape == (H,)                     # sequential: taken once
    assert b["self_attention.dense.weight"].shape == (H, 4)             # row-parallel weight: concatenated on dim 1
    assert b["self_attention.dense.weight"][0].tolist() == [30.0, 30.0, 31.0, 31.0]
    assert b["mlp.dense_h_to_4h.weight"].shape == (4, H)                # column-parallel weight: dim 0
    assert m.get_final_norm_state(0)["weight"].tolist() == [7.0] * H
    assert len(m.get_2d_parallel_files(0, 0)) == 4 and len(m.get_zero_files(0, 0, 0)) == 8
    assert m.get_zero_checkpoint_state(0, 0, 0)["optimizer_state_dict"]["partition_count"] == [1]
    m.show_2d_mapping(); m.show_tp_embedding_map(); m.show_pp_transformer_map(); m.show_transformer_file_map()
    m.show_tp_final_norm_map(); m.validate_files()
    out = capsys.readouterr().out
    assert "[0, 0] = [0, 1, 2, 3]" in out and "pp_to_transformer_layers" in out
    import pytest
    with pytest.raises(AssertionError):
        DeepSpeedCheckpoint(str(tmp_path), tp_degree=4)                  # growing a degree is not a reshape


def _decoupled_worker(rank, world, root):
    """Decoupled checkpointing: save_checkpoint returns while files are still being written; the ``latest`` marker moves
    only at the next optimizer step (after the files are durable), or at destroy(); a slow writer must not let it move
    early; the result loads back."""
    import threading
    import time
    import shuffle_exchange_b200 as ds
    from shuffle_exchange_b200.models import GPT2LMHeadModel, gpt2_config
    from shuffle_exchange_b200.runtime.checkpoint_engine import fast_checkpoint_engine as fce
    from shuffle_exchange_b200.runtime.checkpoint_engine.decoupled_checkpoint_engine import DecoupledCheckpointEngine

    gate = threading.Event()
    slow_save = fce.FastCheckpointEngine.save

    def held_back(self, state, path):
        gate.wait(timeout=60)                           # the background writer is stuck until the test lets it go
        return slow_save(self, state, path)

    cfg = {"train_micro_batch_size_per_gpu": 2, "steps_per_print": 10 ** 9, "bf16": {"enabled": True},
           "optimizer": {"type": "AdamW", "params": {"lr": 1e-3}}, "zero_optimization": {"stage": 2},
           "checkpoint": {"writer": {"type": "python", "decoupled": True}}}
    torch.manual_seed(0)
    model = GPT2LMHeadModel(gpt2_config("gpt2-tiny"))
    engine, _, _, _ = ds.initialize(model=model, config=cfg, model_parameters=model.parameters())
    assert isinstance(engine.checkpoint_engine, DecoupledCheckpointEngine) and engine.checkpoint_engine.is_decoupled()
    ids = torch.randint(0, 512, (2, 16))

    def step():
        engine.backward(engine(ids, labels=ids))
        engine.step()

    step()
    fce.FastCheckpointEngine.save = held_back
    try:
        t0 = time.time()
        engine.save_checkpoint(root, tag="a")
        assert time.time() - t0 < 30                    # did not wait for the (blocked) writer
        info = engine.checkpoint_engine.get_commit_info()
        assert info is not None and (info.tag, info.save_dir, info.save_latest) == ("a", root, True)
        assert not os.path.exists(os.path.join(root, "latest"))
        threading.Timer(1.0, gate.set).start()
        step()                                          # commits: blocks until the writer has finished
    finally:
        fce.FastCheckpointEngine.save = slow_save
    assert gate.is_set() and engine.checkpoint_engine.get_commit_info() is None
    td = torch.distributed
    td.barrier()
    assert open(os.path.join(root, "latest")).read() == "a"
    files = os.listdir(os.path.join(root, "a"))
    assert "mp_rank_00_model_states.pt" in files and sum("optim_states" in f for f in files) == world
    # a save right before destroy still gets its marker
    engine.save_checkpoint(root, tag="b")
    engine.destroy()
    td.barrier()
    assert open(os.path.join(root, "latest")).read() == "b"
    # and loads
    torch.manual_seed(0)
    e2, _, _, _ = ds.initialize(model=GPT2LMHeadModel(gpt2_config("gpt2-tiny")), config=cfg)
    path, _ = e2.load_checkpoint(root)
    assert path is not None and e2.global_steps == 2
    e2.destroy()


def test_decoupled_checkpoint_engine_commits_only_when_durable(tmp_path):
    from conftest import run_dist
    run_dist(_decoupled_worker, 2, (str(tmp_path),))
