"""Repeat the attention numerics cases many times (timing-dependent faults show up as run-to-run variation).
    python tools/attn_stress.py [repeats]
Prints one JSON line: worst error per case over all repeats and the number of repeats above tolerance."""
import importlib.util
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
_spec = importlib.util.spec_from_file_location("gpu_selftest", os.path.join(ROOT, "tools", "gpu_selftest.py"))
st = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(st)


def main():
    reps = int(sys.argv[1]) if len(sys.argv) > 1 else 25
    cases = list(st._ATTN_CFGS) + [(1, 300, 4, 1, 128, True), (2, 200, 2, 2, 128, True), (1, 8192, 8, 2, 128, True),
                                   (1, 1000, 4, 4, 64, True)]
    out = {}
    for cfg in cases:
        worst, best, bad = 0.0, float("inf"), 0
        # element-wise bound of attention backward (tests/test_kernels_gpu.py): 5e-2; the maximum over more elements
        # grows with the sequence, a CPU emulation of the bf16 rounding gives 0.04-0.06 at 1k-4k tokens -> 0.15 at 8k
        tol = 5e-2 if cfg[1] < 4096 else 0.15
        for _ in range(reps if cfg[1] < 4096 else 3):
            r = st._attn_case(*cfg, bwd=True)
            e = max((v if isinstance(v, float) else 1e9) for v in r.values())
            worst, best = max(worst, e), min(best, e)
            bad += int(not (e < tol))
        # the kernels use no atomics: identical inputs must give identical errors, any spread is a timing-dependent fault
        out["x".join(str(int(v)) for v in cfg)] = {"worst": worst, "spread": worst - best, "bad": bad}
    print(json.dumps(out))
    return int(any(v["bad"] or v["spread"] > 0 for v in out.values()))


if __name__ == "__main__":
    sys.exit(main())
