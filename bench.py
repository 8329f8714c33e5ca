#!/usr/bin/env python
"""Headline benchmark: Llama-3-8B, ZeRO-3, bf16, synthetic tokens — tokens/s over the whole job.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

N > 1 is launched by the driver as ``python -m torch.distributed.run --nproc-per-node N ... bench.py
--gpus N`` (one rank per GPU, NCCL).  Prints ONE JSON line on rank 0 (contract in the task brief).

Timed region = K full training steps (forward + backward with activation recomputation + ZeRO-3
gradient reduction + AdamW update on fp32 master shards), bracketed by barrier + synchronize, CUDA
events on the launching stream, MAX over ranks.  Working set (16 GB of bf16 weights alone) is far
larger than the 126 MB L2, so no explicit L2 flush is needed between iterations (stated in config).
``e2e`` repeats the measurement through the public API with the per-step host->device copy of the
token batch from pinned memory and the device->host read of the loss inside the timed region.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--model", default="llama3-8b")
    ap.add_argument("--seq", type=int, default=8192)
    ap.add_argument("--mbs", type=int, default=1)
    ap.add_argument("--layers", type=int, default=0, help="debug only: override layer count (result is then INVALID)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--ckpt-layers", default="auto",
                    help="decoder layers that recompute activations in backward: 'auto' = as few as HBM allows, "
                         "'all', or a count")
    ap.add_argument("--no-fused", action="store_true", help="NCCL baseline path for the ZeRO-3 collectives")
    ap.add_argument("--sp", type=int, default=1,
                    help="Ulysses sequence-parallel degree: --seq is the FULL sequence, every rank holds seq/sp tokens "
                         "(config 4: --sp 8 --seq 131072); ZeRO-3 still shards over all ranks")
    ap.add_argument("--moe-nccl", action="store_true", help="Mixtral: NCCL all_to_all_single exchange (the baseline)")
    ap.add_argument("--moe-fused", action="store_true",
                    help="Mixtral: symmetric-memory dispatch/combine + grouped expert GEMM (moe/fused_ep.py) — the "
                         "default since the 8-GPU run (57.7k vs 55.6k tok/s); kept as an explicit flag")
    ap.add_argument("--fp8-experts", action="store_true", help="Mixtral: e4m3 expert GEMMs in forward")
    ap.add_argument("--ep-capacity-factor", type=float, default=2.0,
                    help="Mixtral fused exchange: receive-buffer rows per rank as a multiple of the balanced share (a "
                         "random-init router is unbalanced: 1.25 overflowed at 8 GPUs; overflow raises, nothing is dropped)")
    ap.add_argument("--no-extras", action="store_true",
                    help="skip the secondary measurements (exposed communication, fused-path rooflines, the line at "
                         "the reference's recompute setting)")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi sampler running during the timed region (profiling recipe's clocks line)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.rows = []
        self.proc = None
        self.gpu_index = gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.gpu_index)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                pass
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def dist_env(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world == 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29533")
        os.environ.setdefault("RANK", "0")
        os.environ.setdefault("WORLD_SIZE", "1")
        os.environ.setdefault("LOCAL_RANK", "0")
    return rank, world, local


def max_over_ranks(x: float, device):
    import torch
    import torch.distributed as td
    t = torch.tensor([x], dtype=torch.float64, device=device)
    if td.is_initialized() and td.get_world_size() > 1:
        td.all_reduce(t, op=td.ReduceOp.MAX)
    return float(t.item())


def barrier_sync():
    import torch
    import torch.distributed as td
    if td.is_initialized() and td.get_world_size() > 1:
        td.barrier()
    torch.cuda.synchronize()


def ds_config(args, world):
    return {
        "train_micro_batch_size_per_gpu": args.mbs,
        "gradient_accumulation_steps": 1,
        "steps_per_print": 1 << 30,
        "bf16": {"enabled": True},
        "optimizer": {"type": "AdamW", "params": {"lr": 1e-5, "betas": [0.9, 0.95], "eps": 1e-8, "weight_decay": 0.1}},
        "zero_optimization": {"stage": 3, "overlap_comm": True, "contiguous_gradients": True,
                              "stage3_param_persistence_threshold": 100000},
        "gradient_clipping": 0.0,
    } if not getattr(args, "no_fused", False) else _with_nccl_collectives(args)


def _with_nccl_collectives(args):
    import copy
    a = copy.copy(args)
    a.no_fused = False
    c = ds_config(a, 0)
    c["zero_optimization"]["fused_collectives"] = False
    return c


def pick_checkpoint_layers(args, cfg, B, S, dev, world, extra_reserve=0):
    """Spend spare HBM on NOT recomputing: every layer left un-checkpointed keeps
    ``cfg.activation_bytes_per_layer`` bytes alive until its backward.  The reserve covers the transient peak of
    a fully checkpointed step (logits + loss grad, two gathered units, lazily allocated gradient arenas, one
    layer being recomputed) measured at 13.4 GiB for this model, plus allocator slack."""
    import torch
    L = cfg.num_hidden_layers
    if args.ckpt_layers == "all":
        return L
    if args.ckpt_layers != "auto":
        return max(0, min(L, int(args.ckpt_layers)))
    total = torch.cuda.get_device_properties(dev).total_memory
    resident = torch.cuda.memory_allocated(dev)
    reserve = 24 * 2 ** 30 + int(extra_reserve)
    per_layer = cfg.activation_bytes_per_layer(B, S)
    free_layers = int(max(0, total - resident - reserve) // per_layer)
    t = torch.tensor([free_layers], device=dev, dtype=torch.int64)
    if world > 1:
        import torch.distributed as td
        td.all_reduce(t, op=td.ReduceOp.MIN)
    return max(0, L - int(t.item()))


def run_ours(args):
    import torch
    rank, world, local = dist_env(args)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    import shuffle_exchange_b200 as sxb
    from shuffle_exchange_b200.utils.logging import logger as _sxb_logger
    for _h in _sxb_logger.handlers:          # stdout carries exactly one JSON line; progress logs go to stderr
        if hasattr(_h, "setStream"):
            _h.setStream(sys.stderr)
    from shuffle_exchange_b200.models.llama import LlamaForCausalLM, llama_config
    from shuffle_exchange_b200.native import lib as nlib

    is_moe = args.model.startswith("mixtral")
    if is_moe:
        # Second half of the headline metric: Mixtral with expert parallelism over all GPUs of the box
        # (experts sharded 1/N per rank, ZeRO-2 for the dense part: the reference asserts MoE + ZeRO-3 unsupported,
        # runtime/engine.py:1760).  The full 8x7B model needs >= 8 GPUs of 180 GB for training state.
        from shuffle_exchange_b200.models.mixtral import MixtralForCausalLM, mixtral_config
        cfg = mixtral_config(args.model, max_position_embeddings=max(args.seq, 8192), ep_size=world,
                             fused_exchange=(False if args.moe_nccl else None), fp8_experts=args.fp8_experts,
                             ep_capacity_factor=args.ep_capacity_factor)
        if args.layers:
            cfg.num_hidden_layers = args.layers
        per_rank = cfg.num_parameters() - cfg.num_hidden_layers * (cfg.num_local_experts - cfg.num_local_experts // world) * 3 * cfg.hidden_size * cfg.intermediate_size
        need = per_rank * 4 + (per_rank - 1.7e9) * 12 + 1.7e9 * 12 / world
        total = torch.cuda.get_device_properties(dev).total_memory
        if need > 0.92 * total and not args.layers:
            if rank == 0:
                print(json.dumps({"impl": "ours", "metric": "tokens/sec/box Mixtral-8x7B EP", "n_gpus": world,
                                  "unavailable": f"training state of {args.model} with EP={world} needs ~{need / 2**30:.0f} GiB "
                                                 f"per GPU (> {total / 2**30:.0f} GiB); run with --gpus 8"}))
            return
        torch.manual_seed(1234)
        model = MixtralForCausalLM(cfg, device=dev, dtype=torch.bfloat16)
        conf = ds_config(args, world)
        conf["zero_optimization"] = {"stage": 2, "overlap_comm": True, "contiguous_gradients": True,
                                     "reduce_bucket_size": int(5e8)}
    else:
        cfg = llama_config(args.model, max_position_embeddings=max(args.seq, 8192))
        if args.layers:
            cfg.num_hidden_layers = args.layers
        if args.sp > 1:
            assert args.sp == world and args.seq % args.sp == 0 and args.mbs == 1, "--sp must equal the number of ranks"
            sxb.init_distributed(dist_backend="nccl")
            from shuffle_exchange_b200.sequence.fused_ulysses import UlyssesPeers
            cfg.ulysses = UlyssesPeers(torch.distributed.group.WORLD, args.seq // args.sp, cfg.num_attention_heads,
                                       cfg.num_key_value_heads, cfg.head_dim)
        torch.manual_seed(1234)
        model = LlamaForCausalLM(cfg, device=dev, dtype=torch.bfloat16)
        conf = ds_config(args, world)
    engine, _, _, _ = sxb.initialize(model=model, config=conf, model_parameters=model.parameters())
    engine.train()
    sp = args.sp
    B, S = args.mbs, args.seq // sp                 # tokens this rank processes per step
    # ZeRO-2 keeps a full-size low-precision gradient buffer that is allocated lazily in the first backward
    extra = sum(p.numel() for p in model.parameters()) * 2 if is_moe else 0
    if is_moe and not args.moe_nccl:
        # symmetric receive / return buffers, the per-layer copies the fused movers keep for backward (combined rows,
        # capacity-padded expert inputs) and allocator slack for their capacity-sized blocks: the first 8-GPU run with
        # a 2 GiB reserve ran out of memory at the start of backward with 171 GiB allocated
        extra += 16 * 2 ** 30
    cfg.checkpoint_layers = pick_checkpoint_layers(args, cfg, B, S, dev, world, extra)
    n_batches = 4
    if sp > 1:
        # one sequence for the whole group: every rank draws the same tokens and keeps its slice; labels are shifted
        # BEFORE slicing (the next-token target of a rank's last token lives on the next rank)
        g = torch.Generator().manual_seed(1000)
        full = [torch.randint(0, cfg.vocab_size, (1, args.seq), generator=g, dtype=torch.int64) for _ in range(n_batches)]
        host, host_lab = [], []
        for f in full:
            lab = torch.full_like(f, -100)
            lab[:, :-1] = f[:, 1:]
            host.append(f[:, rank * S:(rank + 1) * S].contiguous().pin_memory())
            host_lab.append(lab[:, rank * S:(rank + 1) * S].contiguous().pin_memory())
        pos = torch.arange(rank * S, (rank + 1) * S, device=dev).view(1, S)
        kw = dict(position_ids=pos, shift_labels=False)
    else:
        g = torch.Generator().manual_seed(1000 + rank)
        host = [torch.randint(0, cfg.vocab_size, (B, S), generator=g, dtype=torch.int64).pin_memory() for _ in range(n_batches)]
        host_lab = host
        kw = {}
    devb = [h.to(dev, non_blocking=True) for h in host]
    devl = [h.to(dev, non_blocking=True) for h in host_lab] if sp > 1 else devb

    def step_device(i):
        loss = engine(devb[i % n_batches], labels=devl[i % n_batches], **kw)
        engine.backward(loss)
        engine.step()
        return loss

    def step_e2e(i):
        ids = host[i % n_batches].to(dev, non_blocking=True)       # H2D from pinned memory, every step
        lab = host_lab[i % n_batches].to(dev, non_blocking=True) if sp > 1 else ids
        loss = engine(ids, labels=lab, **kw)
        engine.backward(loss)
        engine.step()
        return float(loss.item())                                    # D2H read of the step's result

    if is_moe and args.ckpt_layers == "auto":
        # ZeRO-2 allocates optimizer state and the gradient buffers lazily (first backward / first step), so memory
        # resident before the first step says little.  One untimed step with every layer recomputed, then size the
        # recompute count from what is resident AFTER it plus that step's measured transient peak.
        cfg.checkpoint_layers = cfg.num_hidden_layers
        torch.cuda.reset_peak_memory_stats(dev)
        step_device(0)
        torch.cuda.synchronize()
        resident = torch.cuda.memory_allocated(dev)
        transient = torch.cuda.max_memory_allocated(dev) - resident
        total = torch.cuda.get_device_properties(dev).total_memory
        slack = 10 * 2 ** 30                                  # allocator fragmentation, NCCL / symmetric workspaces
        free_layers = int(max(0, total - resident - transient - slack) // cfg.activation_bytes_per_layer(B, S))
        t = torch.tensor([free_layers], device=dev, dtype=torch.int64)
        if world > 1:
            import torch.distributed as td
            td.all_reduce(t, op=td.ReduceOp.MIN)
        cfg.checkpoint_layers = max(0, cfg.num_hidden_layers - int(t.item()))
        torch.cuda.reset_peak_memory_stats(dev)
    for i in range(args.warmup):
        step_device(i)
    barrier_sync()
    hang = nlib.hang_info("gemm") or nlib.hang_info("attn")
    if hang:
        raise RuntimeError(f"kernel watchdog fired during warm-up: {hang}")
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    nlib.reset_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier_sync()
    e0.record()
    last = None
    for i in range(args.steps):
        last = step_device(i)
    e1.record()
    barrier_sync()
    launches = nlib.reset_launch_count()
    clocks = sampler.stop() if rank == 0 else None
    ms = max_over_ranks(e0.elapsed_time(e1), dev)
    loss_val = float(last.item())
    tokens_per_step = B * S * world
    value = tokens_per_step * args.steps / (ms / 1e3)

    e2e = None
    if not args.no_e2e:
        step_e2e(0)
        barrier_sync()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        for i in range(args.steps):
            step_e2e(i)
        t1.record()
        barrier_sync()
        ems = max_over_ranks(t0.elapsed_time(t1), dev)
        e2e = {"value": tokens_per_step * args.steps / (ems / 1e3), "unit": "tokens/s",
               "h2d_bytes_per_step": B * S * 8, "d2h_bytes_per_step": 4, "ms_per_step": ems / args.steps}
    extras = {}
    if is_moe:
        for m in model.modules():
            if hasattr(m, "check_overflow"):
                m.check_overflow()          # raises if routing ever exceeded the receive-buffer capacity
                break
    if not args.no_extras and not is_moe and sp == 1:
        extras = secondary_measurements(args, engine, cfg, step_device, barrier_sync, max_over_ranks, dev, world, ms / args.steps,
                                        B, S)
    hang = nlib.hang_info("gemm") or nlib.hang_info("attn")
    if rank == 0:
        flops = cfg.flops_per_token(args.seq) * tokens_per_step / world
        out = {
            "metric": ("tokens/sec/box Mixtral-8x7B EP" if is_moe else
                       f"tokens/sec/box Llama-3-8B Ulysses SP={sp}" if sp > 1 else "tokens/sec/box Llama-3-8B ZeRO-3"),
            "value": value, "unit": "tokens/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "impl": "ours",
            "config": {"model": args.model + ("" if not args.layers else f"-DEBUG-{args.layers}-layers"),
                       "global_batch": B * world // sp, "seq_len": args.seq,
                       "parallelism": (f"ep{world}-zero2-dp{world}" if is_moe else
                                       f"ulysses-sp{sp}-zero3-dp{world}" if sp > 1 else f"zero3-dp{world}"),
                       "collectives": (("nccl all-to-all of routed rows" if args.moe_nccl else
                                        "fused symmetric-memory dispatch/combine + grouped expert GEMM"
                                        + (" (fp8 forward)" if args.fp8_experts else "")) + " + ZeRO-2 gradient reduce"
                                       if is_moe else
                                       "nccl" if args.no_fused or world == 1 else "fused symmetric-memory kernels"
                                       + (" + Ulysses exchanges in the GEMM / attention epilogues" if sp > 1 else "")),
                       "activation_checkpointing": f"{cfg.checkpoint_layers} of {cfg.num_hidden_layers} layers recompute "
                                                   f"(--ckpt-layers {args.ckpt_layers})",
                       "optimizer": "AdamW fp32 master",
                       "l2_policy": "inputs (16 GB weights + activations) larger than L2; no explicit flush",
                       "params": cfg.num_parameters()},
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches,
            "gpu_launches_per_step": launches / max(1, args.steps),
            "model_tflops_per_gpu": flops / (ms / args.steps / 1e3) / 1e12, "loss": loss_val, **extras,
            "peak_mem_gb": torch.cuda.max_memory_allocated() / 2 ** 30, "watchdog": hang,
        }
        print(json.dumps(out))
    if world > 1:
        import torch.distributed as td
        td.barrier()
        td.destroy_process_group()


def secondary_measurements(args, engine, cfg, step_device, barrier_sync, max_over_ranks, dev, world, ms_per_step, B, S):
    """Numbers that explain the headline, taken AFTER it (they never touch the timed region above):

    * ``fused_paths``: per fused compute+collective path, device time per step (CUDA events around every launch, max
      over ranks), the bytes it moves over NVLink and the fraction of the NVLink roofline (bytes / measured peer-copy
      bandwidth, B200_PROFILING.md: 770 GB/s per direction) that this is;
    * ``exposed_comm_ms``: step time minus the step time with every collective served from this rank's own memory
      (no NVLink traffic, no cross-rank waits — same kernels, same bytes through HBM), max over ranks; the model's
      weights are garbage afterwards, so this runs last;
    * ``at_reference_recompute``: the same metric with every layer recomputing activations — the policy the
      reference arm (all-or-nothing Hugging Face gradient checkpointing) falls back to when HBM is short."""
    import torch
    out = {}
    opt = engine.optimizer
    fused = getattr(opt, "fused", None)
    n = max(2, min(3, args.steps))

    def timed(k):
        barrier_sync()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for i in range(k):
            step_device(i)
        b.record()
        barrier_sync()
        return max_over_ranks(a.elapsed_time(b), dev) / k

    # like-for-like recompute policy
    if cfg.checkpoint_layers != cfg.num_hidden_layers:
        keep = cfg.checkpoint_layers
        cfg.checkpoint_layers = cfg.num_hidden_layers
        step_device(0)
        t = timed(n)
        cfg.checkpoint_layers = keep
        out["at_reference_recompute"] = {"activation_checkpointing": f"{cfg.num_hidden_layers} of {cfg.num_hidden_layers} layers",
                                         "ms_per_step": t, "value": B * S * world / (t / 1e3), "unit": "tokens/s"}
        step_device(0)
    if fused is None:
        out["exposed_comm_ms"] = 0.0
        out["exposed_comm_note"] = "single rank / NCCL path: no fused collective ran"
        return out
    # per-path device time and NVLink roofline
    LINK = 770e9
    fused.timing = []
    step_device(0)
    per = fused.timing_summary()
    fused.timing = None
    paths = {}
    for what, (cnt, t_ms, nbytes) in per.items():
        t_max = max_over_ranks(t_ms, dev)
        paths[what] = {"launches_per_step": cnt, "device_ms_per_step": t_max, "nvlink_bytes_per_step": nbytes,
                       "roofline_ms": nbytes / LINK * 1e3,
                       "roofline_fraction": (nbytes / LINK * 1e3) / t_max if t_max > 0 else None}
    out["fused_paths"] = paths
    out["fused_paths_note"] = ("device time of every launch of the path summed over a step (launches overlap compute "
                               "on side streams, so this is occupancy of the path, not exposed time); roofline = bytes "
                               "received over NVLink / 770 GB/s measured peer-copy bandwidth per direction")
    # exposed communication
    t_normal = timed(n)
    fused.local_mode = True
    step_device(0)
    t_local = timed(n)
    out["exposed_comm_ms"] = max(0.0, t_normal - t_local)
    out["exposed_comm_fraction"] = max(0.0, t_normal - t_local) / t_normal
    out["exposed_comm_note"] = (f"step {t_normal:.1f} ms vs {t_local:.1f} ms with all peer traffic served from local "
                                "memory (max over ranks, measured after the timed region)")
    return out


def run_reference(args):
    from baseline.reference_arm import run as ref_run
    ref_run(args, ClockSampler, ds_config)


def main():
    args = parse()
    if args.moe_fused:
        args.moe_nccl = False
    if args.impl == "reference":
        try:
            run_reference(args)
        except Exception as e:  # noqa
            if int(os.environ.get("RANK", "0")) == 0:
                print(json.dumps({"impl": "reference", "unavailable": f"{type(e).__name__}: {str(e)[:300]}"}))
        return
    run_ours(args)


if __name__ == "__main__":
    main()
