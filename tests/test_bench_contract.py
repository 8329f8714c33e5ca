"""What the round driver relies on and can be checked without a GPU: ``bench.py``'s command line and the shape of what
it prints, the reference arm's "unavailable" answer, the entry points of ``__graft_entry__``."""
import importlib.util
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _load(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_bench_flags_and_defaults(monkeypatch):
    bench = _load("bench_under_test", os.path.join(ROOT, "bench.py"))
    monkeypatch.setattr(sys, "argv", ["bench.py"])
    a = bench.parse()
    assert (a.gpus, a.impl, a.model, a.seq, a.mbs) == (1, "ours", "llama3-8b", 8192, 1)
    assert a.warmup >= 3 and 1 <= a.steps <= 20          # timing rules: W >= 3; defaults finish within minutes
    monkeypatch.setattr(sys, "argv", ["bench.py", "--gpus", "8", "--steps", "7", "--warmup", "4", "--impl", "reference"])
    a = bench.parse()
    assert (a.gpus, a.steps, a.warmup, a.impl) == (8, 7, 4, "reference")
    # secondary configurations named by BASELINE.json stay reachable
    monkeypatch.setattr(sys, "argv", ["bench.py", "--model", "mixtral-8x7b", "--moe-nccl", "--sp", "8", "--seq", "131072"])
    a = bench.parse()
    assert a.model == "mixtral-8x7b" and a.moe_nccl and a.sp == 8 and a.seq == 131072


def test_reference_arm_answers_unavailable_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        return                                            # on a GPU box the arm runs for real (driver-measured)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                        "--warmup", "1"], capture_output=True, text=True, cwd=ROOT, timeout=600)
    assert r.returncode == 0, r.stderr[-1500:]
    line = [l for l in r.stdout.splitlines() if l.startswith("{")][-1]
    out = json.loads(line)
    assert out["impl"] == "reference" and isinstance(out.get("unavailable"), str) and out["unavailable"]


def test_graft_entry_points_exist_and_reference_is_staged_by_build():
    g = _load("graft_entry_under_test", os.path.join(ROOT, "__graft_entry__.py"))
    assert callable(g.build) and callable(g.smoke)
    make_ref = _load("make_ref_under_test", os.path.join(ROOT, "baseline", "make_ref.py"))
    assert callable(make_ref.is_staged) and callable(make_ref.main)
    src = open(os.path.join(ROOT, "__graft_entry__.py")).read()
    assert "_stage_reference" in src and "compute_100a" in open(
        os.path.join(ROOT, "shuffle_exchange_b200", "native", "build.py")).read()


def test_header_and_library_symbols_agree():
    """Every entry point the Python side registers exists in the built library (a renamed C function would otherwise
    only fail on the GPU box)."""
    import re
    so = os.path.join(ROOT, "shuffle_exchange_b200", "native", "libsxb200.so")
    if not os.path.exists(so):
        return
    # symbol table of the shared object, without loading it (loading needs the CUDA runtime's driver on some images)
    nm = subprocess.run(["nm", "-D", "--defined-only", so], capture_output=True, text=True).stdout
    defined = set(re.findall(r"\b(sxb_\w+)\b", nm))
    wanted = set()
    pkg = os.path.join(ROOT, "shuffle_exchange_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):
                wanted |= set(re.findall(r'_(?:n|lib)\.call\(\s*"(sxb_\w+)"', open(os.path.join(dirpath, f)).read()))
    missing = sorted(wanted - defined)
    assert not missing, missing
