"""Per-shape GEMM table on the Llama-3-8B / Mixtral shapes: the base CTA-pair tcgen05 kernel, the extended kernel
(ungated, gated with all flags set, fp8 forward), and cuBLAS (torch.matmul) on the same box.

    python tools/gemm_table.py            # writes gpurun_out/gemm_table.{json,md}

Timing: CUDA events around each launch, 3 warm-ups, 10 timed iterations, L2 flushed (a 256 MB write) before every timed
launch; reported = median.  TFLOP/s = 2MNK / time."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

SHAPES = [  # (layout, M, N, K, what)
    (0, 8192, 6144, 4096, "qkv fwd"), (0, 8192, 4096, 4096, "o fwd"), (0, 8192, 28672, 4096, "gate|up fwd"),
    (0, 8192, 4096, 14336, "down fwd"), (1, 8192, 4096, 28672, "gate|up dgrad"), (1, 8192, 14336, 4096, "down dgrad"),
    (2, 28672, 4096, 8192, "gate|up wgrad"), (2, 4096, 14336, 8192, "down wgrad"), (0, 8192, 128256, 4096, "lm head"),
    (0, 8192, 8192, 8192, "square"),
]


def timed(fn, flush, iters=10, warm=3):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        flush.fill_(1.0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    from shuffle_exchange_b200.ops import functional as F
    from shuffle_exchange_b200.ops.gemm_ext import Gate, fp8_gemm_nt, fp8_rowwise_quant, gemm_gated
    dev = torch.device("cuda")
    flush = torch.empty(64 << 20, device=dev)          # 256 MB > 126 MB L2
    rows = []
    for layout, M, N, K, what in SHAPES:
        a = torch.randn((K, M) if layout == 2 else (M, K), device=dev).to(torch.bfloat16)
        b = torch.randn((N, K) if layout == 0 else (K, N), device=dev).to(torch.bfloat16)
        out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        fl = 2.0 * M * N * K
        r = {"layout": ["NT", "NN", "TN"][layout], "M": M, "N": N, "K": K, "what": what}
        r["base_ms"] = timed(lambda: F.gemm(a, b, layout, out=out), flush)
        r["ext_ms"] = timed(lambda: gemm_gated(a, b, layout, None, out=out), flush)
        if layout != 2:
            chunk = 1 << 20
            flags = torch.full(((b.numel() * 2 + chunk - 1) // chunk,), 3, dtype=torch.int32, device=dev)
            gate = Gate(flags, 3, chunk, b.data_ptr())
            r["ext_gated_ms"] = timed(lambda: gemm_gated(a, b, layout, gate, out=out), flush)
        ref = (lambda: torch.matmul(a, b.t(), out=out)) if layout == 0 else \
              ((lambda: torch.matmul(a, b, out=out)) if layout == 1 else (lambda: torch.matmul(a.t(), b, out=out)))
        r["cublas_ms"] = timed(ref, flush)
        if layout == 0:
            qa, sa = fp8_rowwise_quant(a)
            qb, sb = fp8_rowwise_quant(b)
            r["fp8_ms"] = timed(lambda: fp8_gemm_nt(qa, sa, qb, sb, out=out), flush)
        for k in list(r):
            if k.endswith("_ms"):
                r[k.replace("_ms", "_tflops")] = fl / r[k] / 1e9
        rows.append(r)
        del a, b, out
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(rows, open(os.path.join(ROOT, "gpurun_out", "gemm_table.json"), "w"), indent=1)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = peaks.get("bf16_tflops", 1590.0)
    with open(os.path.join(ROOT, "gpurun_out", "gemm_table.md"), "w") as f:
        f.write(f"# GEMM table (TFLOP/s; bf16 peak of measured = {peak:.0f}, L2 flushed between launches)\n\n")
        f.write("| op | layout | M | N | K | base pair | ext | ext gated | cuBLAS | base/cuBLAS | base of measured | fp8 ext |\n")
        f.write("|---|---|---|---|---|---|---|---|---|---|---|---|\n")
        for r in rows:
            f.write(f"| {r['what']} | {r['layout']} | {r['M']} | {r['N']} | {r['K']} | {r['base_tflops']:.0f} | "
                    f"{r['ext_tflops']:.0f} | {r.get('ext_gated_tflops', float('nan')):.0f} | {r['cublas_tflops']:.0f} | "
                    f"{r['base_tflops'] / r['cublas_tflops']:.2f} | {r['base_tflops'] / peak:.2f} | "
                    f"{r.get('fp8_tflops', float('nan')):.0f} |\n")
    print(open(os.path.join(ROOT, "gpurun_out", "gemm_table.md")).read())


if __name__ == "__main__":
    main()
