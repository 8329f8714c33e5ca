"""In-tree build of the native libraries.

* ``libsxb200.so``      — every ``csrc/*.cu`` compiled for ``sm_100a`` (tcgen05/TMEM/TMA kernels,
  fused optimizers, norms, attention, quantisers, MoE/ragged kernels, symmetric-memory collectives).
* ``libsxb200_host.so`` — ``csrc/host/*.cpp``: AVX-512/AVX2 + OpenMP CPU optimizers, the async
  file-IO engine and host-side schedulers (g++ only, no CUDA).

The reference JIT-builds one pybind module per op through ``op_builder`` (absent from its tree,
see SURVEY.md §2.3).  Here there is a single C-ABI library per side, built with plain ``nvcc`` /
``g++`` in a few seconds, loaded through ``ctypes`` (``native/lib.py``) and declared in
``include/sxb200.h``.  Objects are cached on a content hash of the source, every header it can include
and the compiler flags (``<obj>.sha``), so a stale object cannot outlive a source change whatever the mtimes say.
"""
from __future__ import annotations

import concurrent.futures as cf
import hashlib
import os
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[2]
CSRC = ROOT / "csrc"
OUT_DIR = Path(__file__).resolve().parent
OBJ_DIR = ROOT / "build" / "obj"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "--expt-relaxed-constexpr",
    "-Xcompiler", "-fPIC",
    "-Xcompiler", "-fvisibility=hidden",
]
HOST_FLAGS = ["-O3", "-std=c++17", "-fPIC", "-fopenmp", "-march=x86-64-v3", "-mavx512f", "-mavx512bw",
              "-mavx512vl", "-mavx512dq", "-fvisibility=hidden", "-pthread"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found; cannot build libsxb200.so")


def _digest(src: Path, extra_deps, flags) -> str:
    h = hashlib.sha256()
    h.update("\0".join(flags).encode())
    for f in [src, *sorted(extra_deps)]:
        h.update(f.name.encode() + b"\0")
        h.update(f.read_bytes())
    return h.hexdigest()


def _stale(src: Path, obj: Path, extra_deps, flags) -> bool:
    sha = obj.with_suffix(obj.suffix + ".sha")
    if not obj.exists() or not sha.exists():
        return True
    return sha.read_text().strip() != _digest(src, extra_deps, flags)


def _stamp(src: Path, obj: Path, extra_deps, flags):
    obj.with_suffix(obj.suffix + ".sha").write_text(_digest(src, extra_deps, flags))


def _run(cmd):
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("build failed:\n" + " ".join(map(str, cmd)) + "\n" + r.stdout + r.stderr)
    return r.stdout + r.stderr


def build_cuda(verbose: bool = False, force: bool = False) -> Path:
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    nvcc = _nvcc()
    headers = list(CSRC.glob("*.cuh")) + list((ROOT / "include").glob("*.h"))
    srcs = sorted(CSRC.glob("*.cu"))
    jobs = []
    objs = []
    rebuilt = []
    for s in srcs:
        o = OBJ_DIR / (s.stem + ".o")
        objs.append(o)
        if force or _stale(s, o, headers, NVCC_FLAGS):
            cmd = [nvcc, *NVCC_FLAGS, "-I", str(ROOT / "include"), "-I", str(CSRC), "-c", str(s), "-o", str(o)]
            if verbose:
                cmd.insert(1, "-Xptxas=-v")
            jobs.append(cmd)
            rebuilt.append((s, o))
    if jobs:
        with cf.ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            for out in ex.map(_run, jobs):
                if verbose and out:
                    print(out)
        for s, o in rebuilt:
            _stamp(s, o, headers, NVCC_FLAGS)
    lib = OUT_DIR / "libsxb200.so"
    if jobs or not lib.exists():
        _run([nvcc, "-shared", "-o", str(lib), *map(str, objs), "-gencode", "arch=compute_100a,code=sm_100a",
              "-Xcompiler", "-fPIC", "-lcudart_static", "-lpthread", "-ldl", "-lrt"])
    return lib


def build_host(verbose: bool = False, force: bool = False) -> Path:
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    gxx = shutil.which("g++") or "g++"
    srcs = sorted((CSRC / "host").glob("*.cpp"))
    headers = list((CSRC / "host").glob("*.h")) + list((ROOT / "include").glob("*.h"))
    jobs, objs, rebuilt = [], [], []
    for s in srcs:
        o = OBJ_DIR / ("host_" + s.stem + ".o")
        objs.append(o)
        if force or _stale(s, o, headers, HOST_FLAGS):
            jobs.append([gxx, *HOST_FLAGS, "-I", str(ROOT / "include"), "-c", str(s), "-o", str(o)])
            rebuilt.append((s, o))
    if jobs:
        with cf.ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            list(ex.map(_run, jobs))
        for s, o in rebuilt:
            _stamp(s, o, headers, HOST_FLAGS)
    lib = OUT_DIR / "libsxb200_host.so"
    if srcs and (jobs or not lib.exists()):
        _run([gxx, "-shared", "-o", str(lib), *map(str, objs), "-fopenmp", "-pthread", "-ldl"])
    return lib


def build_all(verbose: bool = False, force: bool = False):
    return build_cuda(verbose, force), build_host(verbose, force)


if __name__ == "__main__":
    libs = build_all(verbose="-v" in sys.argv, force="-f" in sys.argv)
    for l in libs:
        print("built", l)
