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


def int_layers():
    """INT4 / INT8 layers of inference/quantization at decode sizes: GEMV on the codes vs the layer's dequantise +
    F.linear path (the reference's only path)."""
    import torch.nn.functional as tF
    from shuffle_exchange_b200.inference.quantization import layers as ql
    out = []
    for (N, K) in [(14336 * 2, 4096), (4096, 14336)]:
        lin = torch.nn.Linear(K, N, bias=False, device="cuda", dtype=torch.bfloat16)
        for bits in (4, 8):
            ql.quantized_weight_registry.clear()
            cfg = {"num_bits": bits, "group_size": 64, "group_dim": 1, "symmetric": False}
            q = ql.QuantizedLinear(cfg, lin)
            for rows in (1, 4):
                x = torch.randn(rows, K, device="cuda", dtype=torch.bfloat16)
                t_fused = timed(lambda: q(x))
                t_deq = timed(lambda: tF.linear(x, q.dequantized_weight()))
                packed = N * K * bits / 8 + 2 * (N * K // 64) * 2
                out.append({"N": N, "K": K, "bits": bits, "rows": rows, "gemv_ms": t_fused, "dequant_linear_ms": t_deq,
                            "speedup": t_deq / t_fused, "gemv_weight_TBps": packed / t_fused / 1e9})
                print(json.dumps(out[-1]))
            del q
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(out, open("gpurun_out/int_gemv_bench.json", "w"), indent=1)


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "int":
        return int_layers()
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
