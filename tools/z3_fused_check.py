"""2+ GPU check: ZeRO-3 with the fused symmetric-memory collectives must train like the NCCL path — and both are
arbitrated by an oracle whose gradients are reduced and accumulated in fp32 (stage 0, bf16 optimizer with fp32
gradient accumulation), so a drift of ONE of the two paths shows up as a larger distance to the oracle.

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29558 tools/z3_fused_check.py
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
STEPS = int(os.environ.get("Z3_CHECK_STEPS", "50"))


def run(mode):
    import shuffle_exchange_b200 as sxb
    from shuffle_exchange_b200.models.llama import LlamaForCausalLM, llama_config
    torch.manual_seed(0)
    cfg = llama_config("llama-tiny", hidden_size=512, num_attention_heads=4, num_key_value_heads=2, num_hidden_layers=8)
    model = LlamaForCausalLM(cfg, device="cuda", dtype=torch.bfloat16)
    conf = {"train_micro_batch_size_per_gpu": 2, "bf16": {"enabled": True}, "steps_per_print": 1 << 30,
            "optimizer": {"type": "AdamW", "params": {"lr": 3e-4, "weight_decay": 0.1}}}
    if mode == "oracle":
        conf["zero_optimization"] = {"stage": 0}
        conf["data_types"] = {"grad_accum_dtype": "fp32"}
        conf["communication_data_type"] = "fp32"
    else:
        conf["zero_optimization"] = {"stage": 3, "fused_collectives": mode == "fused",
                                     "stage3_param_persistence_threshold": 0}
    engine, _, _, _ = sxb.initialize(model=model, model_parameters=model.parameters(), config=conf)
    if mode != "oracle":
        assert (engine.optimizer.fused is not None) == (mode == "fused")
    g = torch.Generator().manual_seed(7 + int(os.environ["RANK"]))
    losses = []
    for _ in range(STEPS):
        ids = torch.randint(0, cfg.vocab_size, (2, 256), generator=g).cuda()
        loss = engine(ids, labels=ids)
        engine.backward(loss)
        engine.step()
        losses.append(float(loss))
    torch.cuda.synchronize()
    engine.destroy()
    return losses


def main():
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    a = run("fused")
    b = run("nccl")
    try:
        o = run("oracle")
    except Exception as e:  # the arbiter is optional
        o = None
        oracle_err = f"{type(e).__name__}: {e}"
    from shuffle_exchange_b200.native import lib as n
    import shuffle_exchange_b200.comm.symm  # noqa: F401  (registers sxb_symm_hang_info)
    rel = lambda x, y: max(abs(p - q) / abs(q) for p, q in zip(x, y))
    out = {"steps": STEPS, "fused_last": a[-5:], "nccl_last": b[-5:], "max_rel_diff_fused_vs_nccl": rel(a, b),
           "watchdog": n.hang_info("symm")}
    if o is not None:
        out.update({"oracle_last": o[-5:], "max_rel_diff_fused_vs_oracle": rel(a, o), "max_rel_diff_nccl_vs_oracle": rel(b, o)})
        out["ok"] = out["max_rel_diff_fused_vs_oracle"] < max(2e-2, 2 * out["max_rel_diff_nccl_vs_oracle"])
    else:
        out["oracle_error"] = oracle_err
        out["ok"] = out["max_rel_diff_fused_vs_nccl"] < 2e-2
    if int(os.environ["RANK"]) == 0:
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        json.dump(out, open(os.path.join(ROOT, "gpurun_out", "z3_fused_check.json"), "w"), indent=1)
        print(json.dumps(out))


if __name__ == "__main__":
    main()
