"""Kernel-time breakdown of one training step with torch.profiler (CUPTI), for configurations that ncu cannot cover
(multi-rank).  Usage (torchrun): python tools/step_profile.py --model mixtral-8x7b --layers 2 --steps 2

Prints, for rank 0: wall time per step, summed GPU kernel time, and the top kernels.  Numbers taken under the
profiler are diagnostic only (never bench values)."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="mixtral-8x7b")
    ap.add_argument("--layers", type=int, default=2)
    ap.add_argument("--seq", type=int, default=8192)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--out", default="gpurun_out/step_profile.json")
    ap.add_argument("--gaps", action="store_true", help="report idle gaps of the busiest stream and what ran meanwhile")
    ap.add_argument("--trace-kernel", default="", help="comma-separated kernel-name substrings: report which CPU ops launch them")
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    import shuffle_exchange_b200 as sxb
    is_moe = a.model.startswith("mixtral")
    if is_moe:
        from shuffle_exchange_b200.models.mixtral import MixtralForCausalLM, mixtral_config
        cfg = mixtral_config(a.model, max_position_embeddings=max(a.seq, 8192), ep_size=world, num_hidden_layers=a.layers,
                             activation_checkpointing=False)
        model = MixtralForCausalLM(cfg, device=dev, dtype=torch.bfloat16)
        zero = {"stage": 2, "overlap_comm": True, "contiguous_gradients": True, "reduce_bucket_size": int(5e8)}
    else:
        from shuffle_exchange_b200.models.llama import LlamaForCausalLM, llama_config
        cfg = llama_config(a.model, max_position_embeddings=max(a.seq, 8192), num_hidden_layers=a.layers,
                           activation_checkpointing=False)
        model = LlamaForCausalLM(cfg, device=dev, dtype=torch.bfloat16)
        zero = {"stage": 3, "overlap_comm": True, "contiguous_gradients": True}
    conf = {"train_micro_batch_size_per_gpu": 1, "gradient_accumulation_steps": 1, "steps_per_print": 1 << 30,
            "bf16": {"enabled": True}, "optimizer": {"type": "AdamW", "params": {"lr": 1e-5}},
            "zero_optimization": zero}
    engine, _, _, _ = sxb.initialize(model=model, config=conf, model_parameters=model.parameters())
    ids = torch.randint(0, cfg.vocab_size, (1, a.seq), device=dev)

    def step():
        loss = engine(ids, labels=ids)
        engine.backward(loss)
        engine.step()

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    from torch.profiler import ProfilerActivity, profile
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        e0.record()
        for _ in range(a.steps):
            step()
        e1.record()
        torch.cuda.synchronize()
    if rank == 0:
        wall = e0.elapsed_time(e1) / a.steps
        ev = [e for e in prof.key_averages() if e.device_type is not None and str(e.device_type).endswith("CUDA")]
        rows = sorted(((e.key, e.count / a.steps, e.device_time_total / 1000.0 / a.steps) for e in ev), key=lambda r: -r[2])
        tot = sum(r[2] for r in rows)
        out = {"model": a.model, "layers": a.layers, "world": world, "wall_ms_per_step": wall, "gpu_kernel_ms_per_step": tot,
               "top": [{"kernel": k[:90], "calls": c, "ms": round(ms, 3)} for k, c, ms in rows[:30]]}
        cpu = sorted(((e.key, e.count / a.steps, e.self_cpu_time_total / 1000.0 / a.steps) for e in prof.key_averages()),
                     key=lambda r: -r[2])[:12]
        out["cpu_top"] = [{"op": k[:70], "calls": c, "ms": round(ms, 3)} for k, c, ms in cpu]
        if a.trace_kernel:
            # which autograd / aten op chain launches the named kernels (finds stray fills and copies)
            from collections import Counter
            subs = [t for t in a.trace_kernel.split(",") if t]
            chains = {t: Counter() for t in subs}
            for e in prof.events():
                ks = getattr(e, "kernels", None) or []
                for k in ks:
                    for t in subs:
                        if t in k.name:
                            chain, cur = [], e
                            while cur is not None and len(chain) < 6:
                                chain.append(cur.name[:48])
                                cur = cur.cpu_parent
                            chains[t][" <- ".join(chain)] += 1
            out["kernel_origins"] = {t: [{"chain": c, "launches_per_step": n / a.steps} for c, n in ct.most_common(12)]
                                     for t, ct in chains.items()}
        if a.gaps:
            # timeline of the busiest (compute) stream: where it idles, for how long, what it waited for
            kev = [e for e in prof.profiler.kineto_results.events() if str(e.device_type()).endswith("CUDA")]
            by_stream = {}
            for e in kev:
                by_stream.setdefault(e.device_resource_id(), []).append((e.start_ns(), e.start_ns() + e.duration_ns(), e.name()))
            busy = {sid: sum(b - a_ for a_, b, _ in ev) for sid, ev in by_stream.items()}
            main = max(busy, key=busy.get)
            ev = sorted(by_stream[main])
            others = sorted((x for sid, v in by_stream.items() if sid != main for x in v))
            gaps = []
            for (s0, e0_, n0), (s1, e1_, n1) in zip(ev, ev[1:]):
                if s1 - e0_ > 50_000:
                    during = {}
                    for (so, eo, no) in others:
                        ov = min(eo, s1) - max(so, e0_)
                        if ov > 0:
                            during[no[:40]] = during.get(no[:40], 0) + ov
                    gaps.append({"gap_ms": (s1 - e0_) / 1e6, "after": n0[:50], "before": n1[:50],
                                 "meanwhile": {k: round(v / 1e6, 3) for k, v in sorted(during.items(), key=lambda t: -t[1])[:3]}})
            gaps.sort(key=lambda g: -g["gap_ms"])
            out["streams_busy_ms_per_step"] = {str(k): round(v / 1e6 / a.steps, 2) for k, v in busy.items()}
            out["main_stream_idle_ms_per_step"] = round(sum(g["gap_ms"] for g in gaps) / a.steps, 2)
            out["n_gaps_per_step"] = len(gaps) / a.steps
            out["top_gaps"] = gaps[:25]
            agg = {}
            for g in gaps:
                key = (g["after"][:36], g["before"][:36])
                agg[key] = agg.get(key, 0) + g["gap_ms"]
            out["gaps_by_neighbours"] = [{"after": k[0], "before": k[1], "ms_per_step": round(v / a.steps, 2)}
                                         for k, v in sorted(agg.items(), key=lambda t: -t[1])[:15]]
        os.makedirs(os.path.dirname(a.out), exist_ok=True)
        json.dump(out, open(a.out, "w"), indent=1)
        print(json.dumps(out)[:6000])
    if world > 1:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
