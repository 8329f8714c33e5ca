"""2+ GPU check: ZeRO-3 with the fused symmetric-memory collectives must train like the NCCL path.
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29558 tools/z3_fused_check.py"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def run(fused):
    import shuffle_exchange_b200 as sxb
    from shuffle_exchange_b200.models.llama import LlamaForCausalLM, llama_config
    torch.manual_seed(0)
    cfg = llama_config("llama-tiny", hidden_size=512, num_attention_heads=4, num_key_value_heads=2, num_hidden_layers=4)
    model = LlamaForCausalLM(cfg, device="cuda", dtype=torch.bfloat16)
    engine, _, _, _ = sxb.initialize(model=model, model_parameters=model.parameters(), config={
        "train_micro_batch_size_per_gpu": 2, "bf16": {"enabled": True}, "steps_per_print": 1 << 30,
        "optimizer": {"type": "AdamW", "params": {"lr": 1e-3, "weight_decay": 0.1}},
        "zero_optimization": {"stage": 3, "fused_collectives": fused, "stage3_param_persistence_threshold": 0}})
    assert (engine.optimizer.fused is not None) == fused
    g = torch.Generator().manual_seed(7 + int(os.environ["RANK"]))
    losses = []
    for _ in range(6):
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
    a = run(True)
    b = run(False)
    from shuffle_exchange_b200.native import lib as n
    import shuffle_exchange_b200.comm.symm  # noqa: F401  (registers sxb_symm_hang_info)
    err = max(abs(x - y) / abs(y) for x, y in zip(a, b))
    if int(os.environ["RANK"]) == 0:
        print(json.dumps({"fused": a, "nccl": b, "max_rel_diff": err, "ok": err < 2e-2,
                          "watchdog": n.hang_info("symm")}))


if __name__ == "__main__":
    main()
