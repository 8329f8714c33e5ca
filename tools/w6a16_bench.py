"""W6A16 linear at decode sizes: fused weight-only GEMV vs dequantise-to-scratch + tcgen05 GEMM (CUDA events)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from shuffle_exchange_b200.ops import functional as F
from shuffle_exchange_b200.ops.fp_quantizer import fp_dequantize, fp_gemv, fp_quantize


def timed(fn, n=20):
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


def main():
    out = []
    for (N, K) in [(14336 * 2, 4096), (4096, 14336), (6144, 4096)]:
        w = (torch.randn(N, K, device="cuda") * 0.05).to(torch.bfloat16)
        wq = fp_quantize(w, 512, 6)
        for rows in (1, 4, 8):
            x = torch.randn(rows, K, device="cuda", dtype=torch.bfloat16)
            pad = torch.zeros(max(8, rows), K, device="cuda", dtype=torch.bfloat16)
            pad[:rows] = x
            t_fused = timed(lambda: fp_gemv(x, wq, N, K, 6, 512))
            t_deq = timed(lambda: F.linear(pad, fp_dequantize(wq, 512, 6).view(N, K)))
            packed = N * K * 0.75 + N * K // 512 * 4
            out.append({"N": N, "K": K, "rows": rows, "fused_ms": t_fused, "dequant_gemm_ms": t_deq,
                        "speedup": t_deq / t_fused, "fused_weight_TBps": packed / t_fused / 1e9})
            print(json.dumps(out[-1]))
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(out, open("gpurun_out/w6a16_bench.json", "w"), indent=1)


if __name__ == "__main__":
    main()
