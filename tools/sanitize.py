"""compute-sanitizer sweeps over the kernel checks: one subprocess per (tool, check).

    python tools/sanitize.py                      # memcheck + racecheck + synccheck + initcheck over every numeric check
    python tools/sanitize.py --tools racecheck --checks attn gemm

Each check of ``tools/gpu_selftest.py`` launches the kernel on small shapes, so a sanitizer pass takes seconds.  Results
(error counts per tool and check, first lines of any report) go to ``gpurun_out/sanitizer.json``; exit status is
non-zero if any tool reported an error.  tcgen05 / TMA kernels: ``racecheck`` covers shared-memory hazards between
generic-proxy accesses; mbarrier-ordered async-proxy traffic is outside what it models, which is why the kernels also
carry bounded waits and the stress tool (``tools/attn_stress.py``)."""
import argparse
import json
import os
import re
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TOOLS = ("memcheck", "racecheck", "synccheck", "initcheck")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tools", nargs="*", default=list(TOOLS))
    ap.add_argument("--checks", nargs="*", default=[], help="substrings of check names (default: all numeric checks)")
    ap.add_argument("--timeout", type=int, default=600)
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "sanitizer.json"))
    a = ap.parse_args()
    exe = shutil.which("compute-sanitizer") or "/usr/local/cuda/bin/compute-sanitizer"
    if not os.path.exists(exe):
        print("compute-sanitizer not found", file=sys.stderr)
        return 2
    sys.path.insert(0, ROOT)
    import importlib.util
    spec = importlib.util.spec_from_file_location("gpu_selftest", os.path.join(ROOT, "tools", "gpu_selftest.py"))
    st = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(st)
    names = [n for n in st.CHECKS if not n.endswith("_perf") and (not a.checks or any(c in n for c in a.checks))]
    results, bad = {}, 0
    for tool in a.tools:
        for name in names:
            cmd = [exe, "--tool", tool, "--error-exitcode", "86", "--print-limit", "20", sys.executable,
                   os.path.join(ROOT, "tools", "gpu_selftest.py"), "--one", name]
            try:
                p = subprocess.run(cmd, capture_output=True, text=True, timeout=a.timeout)
                text = p.stdout + p.stderr
                m = re.search(r"ERROR SUMMARY: (\d+) error", text)
                n_err = int(m.group(1)) if m else (0 if p.returncode == 0 else -1)
                rec = {"errors": n_err, "rc": p.returncode}
                if n_err != 0:
                    rec["report"] = [ln for ln in text.splitlines() if ln.startswith("=========")][:30]
                    bad += 1
            except subprocess.TimeoutExpired:
                rec = {"errors": -1, "rc": "timeout"}
                bad += 1
            results.setdefault(tool, {})[name] = rec
            print(tool, name, rec["errors"], flush=True)
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    json.dump(results, open(a.out, "w"), indent=1)
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
