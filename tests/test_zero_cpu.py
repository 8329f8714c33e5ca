"""ZeRO-1/2/3 on CPU/gloo: sharded training must match an unsharded fp32-master AdamW run."""
import copy

import pytest
import torch

from conftest import run_dist


def _cfg(stage, gas=1, clip=0.0, extra=None):
    c = {"train_micro_batch_size_per_gpu": 2, "gradient_accumulation_steps": gas, "steps_per_print": 10 ** 9,
         "bf16": {"enabled": True}, "gradient_clipping": clip,
         "optimizer": {"type": "AdamW", "params": {"lr": 5e-3, "betas": [0.9, 0.95], "weight_decay": 0.05}},
         "zero_optimization": {"stage": stage}}
    if extra:
        c["zero_optimization"].update(extra)
    return c


def _reference_run(model, batches, gas, clip, steps):
    """Plain PyTorch: bf16 forward weights, fp32 master, AdamW, mean gradient over all micro-batches."""
    master = [p.detach().float().clone().requires_grad_() for p in model.parameters()]
    opt = torch.optim.AdamW(master, lr=5e-3, betas=(0.9, 0.95), weight_decay=0.05)
    it = iter(batches)
    for _ in range(steps):
        for p, m in zip(model.parameters(), master):
            p.data = m.detach().to(torch.bfloat16)
        group = [next(it) for _ in range(gas)]
        n = sum(len(g) for g in group)
        acc = [torch.zeros_like(m) for m in master]
        for micro in group:
            for ids in micro:                      # one entry per data-parallel rank
                model.zero_grad()
                (model(ids, labels=ids) / n).backward()
                for a, p in zip(acc, model.parameters()):
                    a += p.grad.float()
        for a, m in zip(acc, master):
            m.grad = a
        if clip > 0:
            torch.nn.utils.clip_grad_norm_(master, clip)
        opt.step()
    return [m.detach() for m in master]


def _worker(rank, world, stage, gas, clip):
    import shuffle_exchange_b200 as ds
    from shuffle_exchange_b200.models import GPT2LMHeadModel, gpt2_config
    from shuffle_exchange_b200.utils.tensor_fragment import safe_get_full_fp32_param
    torch.manual_seed(0)
    model = GPT2LMHeadModel(gpt2_config("gpt2-tiny"))
    ref_model = copy.deepcopy(model).to(torch.bfloat16)
    steps = 3
    g = torch.Generator().manual_seed(5)
    batches = [[torch.randint(0, 512, (2, 16), generator=g) for _ in range(world)] for _ in range(steps * gas)]
    engine, opt, _, _ = ds.initialize(model=model, config=_cfg(stage, gas, clip), model_parameters=model.parameters(),
                                      slice_count=world)
    for micro in batches:
        loss = engine(micro[rank], labels=micro[rank])
        engine.backward(loss)
        engine.step()
    init = [p.detach().float().clone() for p in ref_model.parameters()]
    want = _reference_run(ref_model, batches, gas, clip, steps)
    # Adam normalises every element's update to ~lr, so elements whose gradient is ~0 are decided by
    # rounding noise; compare the update *direction* and bound the worst element by 2*lr*steps.
    num = den_a = den_b = 0.0
    for (n, p), w, w0 in zip(engine.module.named_parameters(), want, init):
        got = safe_get_full_fp32_param(p).cpu()
        da, db = (got - w0).double().flatten(), (w - w0).double().flatten()
        num += float(da @ db); den_a += float(da @ da); den_b += float(db @ db)
        assert (got - w).abs().max().item() <= 2 * 5e-3 * steps + 1e-6, (stage, n)
    cos = num / (den_a ** 0.5 * den_b ** 0.5)
    assert cos > 0.99, (stage, gas, clip, cos)


@pytest.mark.parametrize("stage,gas,clip", [(1, 2, 1.0), (2, 2, 0.0), (3, 1, 0.0), (3, 2, 0.5)])
def test_zero_matches_unsharded(stage, gas, clip):
    run_dist(_worker, 2, (stage, gas, clip))


def _fragment_worker(rank, world):
    import shuffle_exchange_b200 as ds
    from shuffle_exchange_b200.models import GPT2LMHeadModel, gpt2_config
    from shuffle_exchange_b200.utils import tensor_fragment as tf
    torch.manual_seed(0)
    model = GPT2LMHeadModel(gpt2_config("gpt2-tiny"))
    want = {n: p.detach().to(torch.bfloat16).float().clone() for n, p in model.named_parameters()}
    engine, opt, _, _ = ds.initialize(model=model, config=_cfg(3), model_parameters=model.parameters())
    for n, p in engine.module.named_parameters():
        assert torch.equal(tf.safe_get_full_fp32_param(p).cpu(), want[n]), n
    p = next(engine.module.parameters())
    tf.safe_set_full_fp32_param(p, torch.full_like(want["wte"], 0.25))
    assert torch.all(tf.safe_get_full_fp32_param(p) == 0.25)
    ids = torch.randint(0, 512, (2, 16))
    engine.backward(engine(ids, labels=ids))
    engine.step()
    assert tf.safe_get_full_optimizer_state(p, "exp_avg").shape == want["wte"].shape
    # the per-parameter mapping object (reference ``param._hp_mapping``): pieces held by the ranks tile the parameter
    hm = p._hp_mapping
    assert isinstance(hm, tf.tensor_fragment)
    a_lp, a_hp = hm.lp_fragment_address, hm.get_hp_fragment_address()
    assert a_lp.numel == a_hp.numel == hm.get_hp_fragment().numel() == hm.lp_fragment.numel()
    held = torch.tensor([a_lp.numel], dtype=torch.int64)
    torch.distributed.all_reduce(held)
    assert int(held) == p.ds_numel if hasattr(p, "ds_numel") else True
    assert {"exp_avg", "exp_avg_sq"} <= set(hm.get_optim_state_keys())
    assert hm.get_hp_fragment("exp_avg").numel() == a_hp.numel
    full = tf.safe_get_full_fp32_param(p).reshape(-1)
    assert torch.equal(hm.get_hp_fragment().cpu(), full[a_lp.start: a_lp.start + a_lp.numel].cpu())
    with ds.zero.GatheredParameters(list(engine.module.parameters()), modifier_rank=0):
        if rank == 0:
            engine.module.ln_f_w.data.fill_(3.0)
    assert torch.all(tf.safe_get_full_fp32_param(engine.module.ln_f_w) == 3.0)


def test_fragment_api_and_gathered_parameters():
    run_dist(_fragment_worker, 2)


def _zero_init_worker(rank, world):
    import shuffle_exchange_b200 as ds
    from shuffle_exchange_b200.models import GPT2LMHeadModel, gpt2_config
    with ds.zero.Init(dtype=torch.bfloat16):
        model = GPT2LMHeadModel(gpt2_config("gpt2-tiny"))
    assert all(p.numel() == 0 and p.ds_tensor is not None for p in model.parameters())
    engine, _, _, _ = ds.initialize(model=model, config=_cfg(3), model_parameters=model.parameters())
    ids = torch.randint(0, 512, (2, 16))
    l0 = None
    for _ in range(4):
        loss = engine(ids, labels=ids)
        engine.backward(loss)
        engine.step()
        l0 = l0 or loss.item()
    assert loss.item() < l0


def test_zero_init_context():
    run_dist(_zero_init_worker, 2)


def _optimizer_surface(rank, world, stage):
    import shuffle_exchange_b200 as sxb
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(8, 16), torch.nn.ReLU(), torch.nn.Linear(16, 8))
    engine, opt, *_ = sxb.initialize(model=model, model_parameters=model.parameters(), config={
        "train_micro_batch_size_per_gpu": 2, "steps_per_print": 1 << 30, "gradient_clipping": 0.5,
        "optimizer": {"type": "Adam", "params": {"lr": 1e-2}}, "zero_optimization": {"stage": stage}})
    x = torch.randn(2, 8, generator=torch.Generator().manual_seed(rank))
    loss = engine(x).pow(2).mean()
    engine.backward(loss)
    p = next(engine.module.parameters())
    g_full = opt.get_fp32_grad_for_param(p)
    assert g_full is None or g_full.shape == (16, 8)
    assert not opt.has_overflow() and opt.scaled_global_norm() > 0
    f = opt.unscale_and_clip_grads(total_norm=2.0)
    assert abs(f - 0.5 / (2.0 + 1e-6)) < 1e-6                      # loss scale 1, clip 0.5 at norm 2
    engine.step()
    w = opt.get_full_hp_param(p)
    assert w.shape == (16, 8) and opt.get_full_hp_param(p, "exp_avg").shape == (16, 8)
    opt.set_full_hp_param(torch.full((16, 8), 0.25), p)
    assert torch.allclose(opt.get_full_hp_param(p), torch.full((16, 8), 0.25))
    assert opt.get_param_id(p) == 0 and len(opt.get_bit16_param_group(0)) >= 1
    lean = opt.get_lean_optimizer_state()
    assert len(lean) == len(opt.groups) and "exp_avg" in lean[0]
    assert len(opt.get_groups_without_padding()) == len(opt.groups)
    assert opt.get_hp_param_device(p).type == "cpu"
    opt.check_overflow(); assert opt.overflow is False
    opt.checkpoint_event_prologue(); opt.checkpoint_event_epilogue(); opt.reset_swap_buffers()


@pytest.mark.parametrize("stage", [2, 3])
def test_optimizer_object_surface(stage):
    run_dist(_optimizer_surface, world=2, args=(stage,))


def _bf16_optimizer_methods(rank, world):
    import shuffle_exchange_b200 as sxb
    from shuffle_exchange_b200.runtime.bf16_optimizer import BF16_Optimizer
    torch.manual_seed(0)
    model = torch.nn.Linear(8, 8).to(torch.bfloat16)
    engine, opt, *_ = sxb.initialize(model=model, model_parameters=model.parameters(), config={
        "train_micro_batch_size_per_gpu": 2, "steps_per_print": 1 << 30, "bf16": {"enabled": True},
        "optimizer": {"type": "Adam", "params": {"lr": 1e-2}}})
    assert isinstance(opt, BF16_Optimizer), type(opt)
    loss = engine(torch.randn(2, 8, dtype=torch.bfloat16)).float().pow(2).mean()
    engine.backward(loss)
    gs = opt.get_grads_for_norm()
    assert len(gs) >= 1 and all(torch.is_tensor(t) for t in gs)
    non_expert, expert = opt.get_grads_for_norm(for_clipping=True)
    assert expert == {} and len(non_expert) == len(gs)
    engine.step()
    p = next(model.parameters())
    p.grad = torch.ones_like(p)
    opt.accumulate_hp_grads_and_remove_lp(p, 0, 0)
    assert p.grad is None and float(opt.groups[0].grad_view(0).sum()) == p.numel()
    opt.clear_hp_grads()
    assert float(opt.groups[0].flat_grad.abs().sum()) == 0.0 and opt.create_grad_acc_hooks() is None


def test_bf16_optimizer_reference_named_methods():
    run_dist(_bf16_optimizer_methods, world=2)


def _offload_states_worker(rank, world, stage):
    """engine.offload_states / reload_states between steps must not change training: every state tensor is replaced by
    a moved copy (on this CPU box the "move" is forced to reallocate, which is what exposes stale references)."""
    import shuffle_exchange_b200 as ds
    from shuffle_exchange_b200.models import GPT2LMHeadModel, gpt2_config
    from shuffle_exchange_b200.runtime.zero import offload_states as osm
    from shuffle_exchange_b200.runtime.zero.offload_states import OffloadStateTypeEnum as K

    moved = []

    def relocating_to(t, device, pin_memory, non_blocking):
        if t is None:
            return t
        moved.append(t.data_ptr())
        return t.clone()

    def run(offload):
        torch.manual_seed(0)
        model = GPT2LMHeadModel(gpt2_config("gpt2-tiny"))
        engine, _, _, _ = ds.initialize(model=model, config=_cfg(stage), model_parameters=model.parameters())
        g = torch.Generator().manual_seed(7 + rank)
        losses = []
        for step in range(4):
            ids = torch.randint(0, 512, (2, 16), generator=g)
            loss = engine(ids, labels=ids)
            engine.backward(loss)
            engine.step()
            losses.append(float(loss))
            if offload and step in (0, 2):
                include = None if step == 0 else [K.optim_states, K.hp_params]
                engine.offload_states(include=include, device="cpu", pin_memory=False)
                assert osm.get_state_devices(engine, K.hp_params) == {torch.device("cpu")}
                assert osm.get_state_devices(engine, K.optim_states) == {torch.device("cpu")}
                engine.reload_states()
        return losses

    base = run(False)
    orig = osm._to
    osm._to = relocating_to
    try:
        got = run(True)
    finally:
        osm._to = orig
    assert moved, "nothing was offloaded"
    assert got == base, (got, base)


@pytest.mark.parametrize("stage", [2, 3])
def test_offload_and_reload_states_keep_training_identical(stage):
    run_dist(_offload_states_worker, 2, (stage,))


def _zero_init_module_granularity_worker(rank, world):
    """``Init.granularity = "module"``: a Hugging Face model (whose post_init runs ``self.apply(self._init_weights)``) is
    built without any rank ever holding it whole — every sub-module is sharded when its constructor returns — the HF
    initialiser still takes effect (through the gathered ``apply``), every rank ends up with rank 0's values, and the
    model trains."""
    import pytest as _pt
    transformers = _pt.importorskip("transformers")
    import shuffle_exchange_b200 as ds
    from shuffle_exchange_b200.runtime.zero import partition_parameters as pp
    cfg = transformers.AutoConfig.for_model(model_type="llama", vocab_size=128, hidden_size=32, intermediate_size=64,
                                            num_hidden_layers=2, num_attention_heads=4, num_key_value_heads=2,
                                            max_position_embeddings=64, initializer_range=0.5)
    held = []                      # elements still whole under the module being sharded, at each sharding event
    orig_convert = pp.Init._convert_module

    def tracking(self, module, recurse=True):
        held.append(sum(p.numel() for p in module.parameters(recurse=True)))
        orig_convert(self, module, recurse)

    pp.Init._convert_module = tracking
    pp.Init.granularity = "module"
    try:
        torch.manual_seed(100 + rank)               # ranks would initialise differently ...
        with ds.zero.Init(dtype=torch.float32):
            model = transformers.AutoModelForCausalLM.from_config(cfg)
    finally:
        pp.Init._convert_module = orig_convert
        pp.Init.granularity = "model"
    assert max(held) <= 128 * 32, max(held)         # never more than one sub-module (the embedding) whole at a time
    params = list(model.parameters())
    assert all(p.numel() == 0 and p.ds_tensor is not None for p in params)
    total = sum(p.ds_numel for p in params)
    w = model.model.layers[0].mlp.down_proj.weight
    with ds.zero.GatheredParameters([w]):
        full = w.data.clone()
    assert abs(float(full.std()) - 0.5) < 0.1, float(full.std())      # HF's normal_(std=initializer_range) was applied
    ref = full.clone()
    torch.distributed.broadcast(ref, 0)
    assert torch.equal(ref, full)                                      # ... but all hold rank 0's initial values
    assert torch.nn.Module.apply.__name__ == "apply"                  # the patch is lifted on exit
    engine, _, _, _ = ds.initialize(model=model, config=_cfg(3), model_parameters=model.parameters())
    ids = torch.randint(0, 128, (2, 16))
    first = None
    for _ in range(4):
        loss = engine(input_ids=ids, labels=ids).loss
        engine.backward(loss)
        engine.step()
        first = first if first is not None else float(loss)
    assert float(loss) < first
    assert total == sum(p.ds_numel for p in model.parameters())


def test_zero_init_shards_per_module_with_hf_post_init():
    run_dist(_zero_init_module_granularity_worker, 2)
