"""End-to-end on a B200: the flagship model through the public API."""
import os

import pytest
import torch


@pytest.mark.gpu
@pytest.mark.timeout(300)
def test_smoke_entry_point():
    import __graft_entry__ as g
    g.smoke()


@pytest.mark.gpu
@pytest.mark.timeout(300)
@pytest.mark.parametrize("stage", [1, 2, 3])
def test_llama_trains_under_each_zero_stage(stage):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(29580 + stage), RANK="0", WORLD_SIZE="1", LOCAL_RANK="0")
    import shuffle_exchange_b200 as sxb
    from shuffle_exchange_b200.models.llama import LlamaForCausalLM, llama_config
    from shuffle_exchange_b200.native import lib
    torch.manual_seed(0)
    cfg = llama_config("llama-tiny", hidden_size=512, num_attention_heads=4, num_key_value_heads=2)
    model = LlamaForCausalLM(cfg, device="cuda", dtype=torch.bfloat16)
    engine, _, _, _ = sxb.initialize(model=model, model_parameters=model.parameters(), config={
        "train_micro_batch_size_per_gpu": 2, "gradient_accumulation_steps": 2, "bf16": {"enabled": True},
        "steps_per_print": 1 << 30, "gradient_clipping": 1.0,
        "optimizer": {"type": "AdamW", "params": {"lr": 1e-3}}, "zero_optimization": {"stage": stage}})
    ids = torch.randint(0, cfg.vocab_size, (2, 256), device="cuda")
    losses = []
    lib.reset_launch_count()
    for _ in range(8):
        loss = engine(ids, labels=ids)
        engine.backward(loss)
        engine.step()
        losses.append(float(loss))
    assert lib.launch_count > 100
    assert losses[-1] < losses[0] - 0.5, losses
    assert engine.get_global_grad_norm() is not None


@pytest.mark.gpu
@pytest.mark.timeout(300)
def test_checkpoint_roundtrip_gpu(tmp_path):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT="29590", RANK="0", WORLD_SIZE="1", LOCAL_RANK="0")
    import shuffle_exchange_b200 as sxb
    from shuffle_exchange_b200.models.llama import LlamaForCausalLM, llama_config

    def make():
        torch.manual_seed(0)
        cfg = llama_config("llama-tiny", hidden_size=512, num_attention_heads=4, num_key_value_heads=2)
        m = LlamaForCausalLM(cfg, device="cuda", dtype=torch.bfloat16)
        e, _, _, _ = sxb.initialize(model=m, model_parameters=m.parameters(), config={
            "train_micro_batch_size_per_gpu": 2, "bf16": {"enabled": True}, "steps_per_print": 1 << 30,
            "optimizer": {"type": "AdamW", "params": {"lr": 1e-3}}, "zero_optimization": {"stage": 3}})
        return e, cfg
    e1, cfg = make()
    ids = torch.randint(0, cfg.vocab_size, (2, 128), device="cuda")
    for _ in range(3):
        e1.backward(e1(ids, labels=ids))
        e1.step()
    e1.save_checkpoint(str(tmp_path), tag="t3", client_state={"hello": 7})
    files = sorted(os.listdir(tmp_path / "t3"))
    assert "zero_pp_rank_0_mp_rank_00_model_states.pt" in files
    assert "bf16_zero_pp_rank_0_mp_rank_00_optim_states.pt" in files and (tmp_path / "latest").read_text() == "t3"
    l_next = float(e1(ids, labels=ids))
    e2, _ = make()
    path, client = e2.load_checkpoint(str(tmp_path))
    assert path is not None and client["hello"] == 7 and e2.global_steps == 3
    assert abs(float(e2(ids, labels=ids)) - l_next) < 1e-3


@pytest.mark.gpu
@pytest.mark.timeout(300)
@pytest.mark.parametrize("stage", [1, 2, 3])
def test_tied_embedding_gradients_match_fp32_oracle(stage):
    """Llama with ``tie_word_embeddings``: the shared table receives one wgrad (LM head) and one scatter-add
    (embedding) through the main-grad protocol; the reduced gradient must contain both."""
    import copy
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(29600 + stage), RANK="0", WORLD_SIZE="1", LOCAL_RANK="0")
    import shuffle_exchange_b200 as sxb
    from shuffle_exchange_b200.models.llama import LlamaForCausalLM, llama_config
    from shuffle_exchange_b200.utils.tensor_fragment import safe_get_full_grad
    torch.manual_seed(0)
    cfg = llama_config("llama-tiny", hidden_size=256, num_attention_heads=4, num_key_value_heads=2,
                       tie_word_embeddings=True)
    model = LlamaForCausalLM(cfg, device="cuda", dtype=torch.bfloat16)
    oracle = copy.deepcopy(model).float().cpu()          # CPU branch of every op = plain PyTorch fp32
    engine, _, _, _ = sxb.initialize(model=model, model_parameters=model.parameters(), config={
        "train_micro_batch_size_per_gpu": 2, "bf16": {"enabled": True}, "steps_per_print": 1 << 30,
        "optimizer": {"type": "AdamW", "params": {"lr": 1e-3}},
        "zero_optimization": {"stage": stage, "fused_step": False}})
    ids = torch.randint(0, cfg.vocab_size, (2, 128), device="cuda")
    loss = engine(ids, labels=ids)
    engine.backward(loss)
    got = {n: safe_get_full_grad(p).float().cpu() for n, p in engine.module.named_parameters()}
    oracle(ids.cpu(), labels=ids.cpu()).backward()
    for n, p in oracle.named_parameters():
        ref = p.grad
        err = (got[n] - ref).abs()
        tol = 0.06 * ref.abs().max() + 0.06 * ref.abs()
        assert bool((err <= tol).all()), (stage, n, err.max().item(), ref.abs().max().item())


@pytest.mark.gpu
@pytest.mark.timeout(300)
@pytest.mark.parametrize("stage", [0, 2])
def test_llama_trains_in_fp16_with_dynamic_loss_scale(stage):
    """The reference's fp16 mode (dynamic loss scaling, overflow skips) on a GPU: the model's GEMMs and attention take
    the library path for fp16 (the tcgen05 kernels are bf16), everything else stays on the native kernels."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(29570 + stage), RANK="0", WORLD_SIZE="1", LOCAL_RANK="0")
    import warnings
    import shuffle_exchange_b200 as sxb
    from shuffle_exchange_b200.models.llama import LlamaForCausalLM, llama_config
    torch.manual_seed(0)
    cfg = llama_config("llama-tiny", hidden_size=256, num_attention_heads=4, num_key_value_heads=2)
    model = LlamaForCausalLM(cfg, device="cuda", dtype=torch.float16)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        engine, _, _, _ = sxb.initialize(model=model, model_parameters=model.parameters(), config={
            "train_micro_batch_size_per_gpu": 2, "steps_per_print": 1 << 30,
            "fp16": {"enabled": True, "initial_scale_power": 12, "loss_scale_window": 4, "hysteresis": 1},
            "optimizer": {"type": "AdamW", "params": {"lr": 1e-3}}, "zero_optimization": {"stage": stage}})
        ids = torch.randint(0, cfg.vocab_size, (2, 128), device="cuda")
        losses = []
        for _ in range(10):
            loss = engine(ids, labels=ids)
            engine.backward(loss)
            engine.step()
            losses.append(float(loss))
    assert all(l == l for l in losses), losses
    assert losses[-1] < losses[0] - 0.3, losses
    assert engine.optimizer.loss_scaler.dynamic and engine.optimizer.loss_scaler.cur_scale >= 1
