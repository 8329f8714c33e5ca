"""Reference arm of bench.py: the UNMODIFIED reference package (staged by baseline/make_ref.py under
baseline/_ref) driving a stock Hugging Face ``LlamaForCausalLM`` (random init, sdpa attention,
gradient checkpointing) through the reference's own public API::

    engine, *_ = deepspeed.initialize(model=model, model_parameters=..., config=zero3_bf16_config)
    loss = engine(input_ids=ids, labels=ids).loss ; engine.backward(loss) ; engine.step()

Same metric, model config, batch, sequence length, optimizer hyper-parameters and timing protocol as
the product arm.  Nothing from ``shuffle_exchange_b200`` is imported here.  The optimizer is
``torch.optim.AdamW`` via the reference's ``torch_adam: true`` switch because the reference's fused-Adam
CUDA sources are not in its tree.
"""
from __future__ import annotations

import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "_ref")

LLAMA3_8B = dict(vocab_size=128256, hidden_size=4096, intermediate_size=14336, num_hidden_layers=32,
                 num_attention_heads=32, num_key_value_heads=8, max_position_embeddings=8192, rope_theta=500000.0,
                 rms_norm_eps=1e-5, tie_word_embeddings=False, hidden_act="silu", attention_bias=False, mlp_bias=False)


def run(args, ClockSampler, ds_config_fn):
    if not str(args.model).startswith("llama3-8b") or getattr(args, "sp", 1) > 1:
        # the arm drives Llama-3-8B under ZeRO-3 data parallelism only; a run of another model or of sequence
        # parallelism labelled with that name would be a wrong number, so say so instead
        raise NotImplementedError(f"reference arm covers llama3-8b ZeRO-3 DP only (asked: model={args.model}, "
                                  f"sp={getattr(args, 'sp', 1)})")
    from baseline import make_ref
    if not make_ref.is_staged():
        make_ref.main()          # raises with the staging instructions when /root/reference is not mounted
    sys.path.insert(0, REF)
    import torch
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29534")
    os.environ.setdefault("RANK", "0")
    os.environ.setdefault("WORLD_SIZE", "1")
    os.environ.setdefault("LOCAL_RANK", "0")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    import deepspeed  # the reference
    assert os.path.realpath(deepspeed.__file__).startswith(os.path.realpath(REF)), deepspeed.__file__
    from transformers import LlamaConfig, LlamaForCausalLM

    kw = dict(LLAMA3_8B)
    if args.layers:
        kw["num_hidden_layers"] = args.layers
    kw["max_position_embeddings"] = max(args.seq, 8192)
    cfg = LlamaConfig(**kw)
    cfg._attn_implementation = "sdpa"
    torch.manual_seed(1234)
    old = torch.get_default_dtype()
    torch.set_default_dtype(torch.bfloat16)
    with torch.device(dev):
        model = LlamaForCausalLM(cfg)
    torch.set_default_dtype(old)
    model.gradient_checkpointing_enable(gradient_checkpointing_kwargs={"use_reentrant": False})
    model.config.use_cache = False
    model.train()

    conf = ds_config_fn(args, world)
    conf["optimizer"]["params"]["torch_adam"] = True
    conf["zero_optimization"].pop("fused_step", None)
    conf["zero_optimization"].update({"reduce_bucket_size": int(2e8), "stage3_prefetch_bucket_size": int(5e7),
                                      "stage3_max_live_parameters": int(1e9), "stage3_max_reuse_distance": int(1e9)})
    engine, _, _, _ = deepspeed.initialize(model=model, model_parameters=model.parameters(), config=conf)
    B, S = args.mbs, args.seq
    # Same policy as our arm: recompute only when HBM demands it.  HF's switch is all-or-nothing; a
    # non-checkpointed HF Llama layer keeps ~2.3 GB at B*S = 8192 tokens (fp32 norm upcasts, q/k/v + rotated
    # copies, gate/up/act/product), scaled linearly with the token count.
    ckpt_mode = "per-layer (HF gradient_checkpointing)"
    if getattr(args, "ckpt_layers", "auto") == "auto":
        total = torch.cuda.get_device_properties(dev).total_memory
        # torch.optim.AdamW allocates exp_avg / exp_avg_sq (8 bytes per owned parameter) lazily in the FIRST step —
        # they are not in memory_allocated() yet (omitting them let a 1-GPU run drop recomputation and then run out
        # of memory in its second step)
        n_params = sum(getattr(p, "ds_numel", p.numel()) for p in model.parameters())
        lazy_adam = 8 * n_params // world
        need = cfg.num_hidden_layers * int(2.3e9 * (B * S) / 8192) + 36 * 2 ** 30 + lazy_adam
        fits = torch.tensor([1 if total - torch.cuda.memory_allocated(dev) >= need else 0], device=dev)
        if world > 1:
            torch.distributed.all_reduce(fits, op=torch.distributed.ReduceOp.MIN)
        if int(fits.item()):
            model.gradient_checkpointing_disable()
            ckpt_mode = "none (fits in HBM)"
    elif args.ckpt_layers not in ("all",) and int(args.ckpt_layers) == 0:
        model.gradient_checkpointing_disable()
        ckpt_mode = "none (--ckpt-layers 0)"
    g = torch.Generator().manual_seed(1000 + rank)
    n_batches = 4
    host = [torch.randint(0, cfg.vocab_size, (B, S), generator=g, dtype=torch.int64).pin_memory() for _ in range(n_batches)]
    devb = [h.to(dev, non_blocking=True) for h in host]

    def step(ids):
        loss = engine(input_ids=ids, labels=ids).loss
        engine.backward(loss)
        engine.step()
        return loss

    import torch.distributed as td

    def barrier_sync():
        if td.is_initialized() and td.get_world_size() > 1:
            td.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if td.is_initialized() and td.get_world_size() > 1:
            td.all_reduce(t, op=td.ReduceOp.MAX)
        return float(t.item())

    for i in range(args.warmup):
        step(devb[i % n_batches])
    barrier_sync()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier_sync()
    e0.record()
    last = None
    for i in range(args.steps):
        last = step(devb[i % n_batches])
    e1.record()
    barrier_sync()
    clocks = sampler.stop() if rank == 0 else None
    ms = max_over_ranks(e0.elapsed_time(e1))
    tokens_per_step = B * S * world
    value = tokens_per_step * args.steps / (ms / 1e3)
    e2e = None
    if not args.no_e2e:
        float(step(host[0].to(dev, non_blocking=True)).item())
        barrier_sync()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        for i in range(args.steps):
            float(step(host[i % n_batches].to(dev, non_blocking=True)).item())
        t1.record()
        barrier_sync()
        ems = max_over_ranks(t0.elapsed_time(t1))
        e2e = {"value": tokens_per_step * args.steps / (ems / 1e3), "unit": "tokens/s",
               "h2d_bytes_per_step": B * S * 8, "d2h_bytes_per_step": 4, "ms_per_step": ems / args.steps}
    if rank == 0:
        print(json.dumps({
            "metric": "tokens/sec/box Llama-3-8B ZeRO-3", "value": value, "unit": "tokens/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "impl": "reference",
            "config": {"model": args.model + ("" if not args.layers else f"-DEBUG-{args.layers}-layers"),
                       "global_batch": B * world, "seq_len": S, "parallelism": f"zero3-dp{world}",
                       "activation_checkpointing": ckpt_mode,
                       "optimizer": "torch.optim.AdamW (torch_adam)", "attention": "torch sdpa",
                       "l2_policy": "inputs larger than L2; no explicit flush"},
            "clocks": clocks, "e2e": e2e, "gpu_launches": 0, "loss": float(last.item()),
            "peak_mem_gb": torch.cuda.max_memory_allocated() / 2 ** 30}))
    if world > 1:
        td.barrier()
