"""Prefill / mixed-batch timing of the ragged attention: atom-scheduled tcgen05 kernel vs the per-token SIMT kernel.
Writes gpurun_out/paged_attn_bench.json (CUDA events, 3 warm-ups, 10 timed calls)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from shuffle_exchange_b200.inference.v2.kernels import ragged_ops as R
from shuffle_exchange_b200.inference.v2.ragged.ragged_wrapper import RaggedBatchWrapper


class _Seq:
    def __init__(self, seen, blocks):
        self.seen_tokens, self._b = seen, torch.tensor(blocks, dtype=torch.int32)

    def kv_blocks(self):
        return self._b


def timed(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n


def case(name, plan, Hq=32, Hk=8, D=128, bs=64):
    total_blocks = sum((s + n + bs - 1) // bs for s, n in plan) + 1
    cache = torch.randn(total_blocks, bs, 2, Hk, D, device="cuda", dtype=torch.bfloat16)
    w = RaggedBatchWrapper(sum(n for _, n in plan), len(plan), max((s + n + bs - 1) // bs for s, n in plan), device="cuda")
    nxt = 0
    for seen, n in plan:
        need = (seen + n + bs - 1) // bs
        w.insert_sequence(_Seq(seen, list(range(nxt, nxt + need))), torch.zeros(n, dtype=torch.int64))
        nxt += need
    w.finalize()
    q = torch.randn(w.current_tokens, Hq, D, device="cuda", dtype=torch.bfloat16)
    t_atoms = timed(lambda: R.ragged_attention(q, cache, w))
    t_simt = timed(lambda: R.paged_attention(q, cache, w.token_to_seq(), w.token_positions(), w.block_tables()),
                   n=3 if w.current_tokens > 2048 else 10)
    flops = sum(4.0 * Hq * D * (n * seen + n * (n + 1) / 2) for seen, n in plan)
    return {"case": name, "tokens": w.current_tokens, "atoms": int(w.atoms().shape[0]), "atoms_ms": t_atoms,
            "simt_ms": t_simt, "speedup": t_simt / t_atoms, "atoms_tflops": flops / t_atoms / 1e9}


def main():
    out = [case("prefill 4096", [(0, 4096)]),
           case("prefill 4 x 1024", [(0, 1024)] * 4),
           case("chunked prefill 512 @ 8192 seen", [(8192, 512)]),
           case("decode 64 seqs @ 2048", [(2048, 1)] * 64),
           case("mixed: prefill 2048 + 32 decodes @ 1024", [(0, 2048)] + [(1024, 1)] * 32)]
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(out, open("gpurun_out/paged_attn_bench.json", "w"), indent=1)
    for r in out:
        print(json.dumps(r))


if __name__ == "__main__":
    main()
