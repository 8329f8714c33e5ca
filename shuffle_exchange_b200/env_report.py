"""``ds_report``-style environment report (reference: ``deepspeed/env_report.py:20-195``): native op table
(installed / compatible), toolchain and library versions, device facts.

    python -m shuffle_exchange_b200.env_report [--hide_operator_status] [--hide_errors_and_warnings]
"""
from __future__ import annotations

import argparse
import importlib
import os
import shutil
import subprocess
import sys

import torch

from . import git_version_info as gvi

GREEN, RED, YELLOW, END = "\033[92m", "\033[91m", "\033[93m", "\033[0m"
SUCCESS, WARNING, FAIL, INFO = f"{GREEN} [SUCCESS] {END}", f"{YELLOW} [WARNING] {END}", f"{RED} [FAIL] {END}", "[INFO]"
OKAY, NO, YES = f"{GREEN}[OKAY]{END}", f"{RED}[NO]{END}", f"{GREEN}[YES]{END}"
color_len = len(GREEN) + len(END)


def op_report(verbose=True):
    max_dots = 23
    print("-" * 50 + "\nDeepSpeed C++/CUDA extension op report\n" + "-" * 50)
    print("NOTE: every op family lives in ONE of two in-tree libraries (native/libsxb200.so for sm_100a kernels,\n"
          "      native/libsxb200_host.so for host code); `python -m shuffle_exchange_b200.native.build` rebuilds them.")
    print("-" * 50)
    h = ["op name", "installed", "compatible"]
    print(h[0], "." * (max_dots - len(h[0])), h[1], "..", h[2])
    print("-" * 50)
    status = gvi.op_status()
    for name in sorted(status):
        inst = YES if status[name] else NO
        print(name, "." * (max_dots - len(name)), inst, "." * (15 + color_len - len(inst)), OKAY)
    print("-" * 50)
    return status


def nvcc_version():
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        return f"{RED} [FAIL] cannot find nvcc {END}"
    try:
        out = subprocess.check_output([nvcc, "-V"], text=True)
        rel = out.split("release")[-1].split(",")[0].strip()
        return rel
    except Exception:
        return f"{RED} [FAIL] nvcc missing {END}"


def installed_cann_version():
    return None


def get_shm_size():
    try:
        st = os.statvfs("/dev/shm")
        size = st.f_frsize * st.f_blocks
        for unit in ("B", "KB", "MB", "GB", "TB"):
            if size < 1024:
                break
            size /= 1024
        return f"{size:.2f} {unit}", None
    except Exception:
        return "UNKNOWN", None


def debug_report():
    import shuffle_exchange_b200 as pkg
    rows = [("torch install path", torch.__path__), ("torch version", torch.__version__),
            ("shuffle_exchange_b200 install path", pkg.__path__),
            ("shuffle_exchange_b200 info", f"{gvi.version}, {gvi.git_hash}, {gvi.git_branch}"),
            ("torch cuda version", torch.version.cuda), ("nvcc version", nvcc_version()),
            ("target arch", "sm_100a (compute_100a)")]
    if torch.cuda.is_available():
        p = torch.cuda.get_device_properties(0)
        rows += [("device", f"{p.name}, cc {p.major}.{p.minor}, {p.multi_processor_count} SMs, {p.total_memory / 2**30:.0f} GiB"),
                 ("device count", torch.cuda.device_count())]
    else:
        rows.append(("device", "no CUDA device visible (build-only environment)"))
    try:
        rows.append(("nccl version", ".".join(map(str, torch.cuda.nccl.version()))))
    except Exception:
        rows.append(("nccl version", "unknown"))
    rows.append(("shared memory (/dev/shm) size", get_shm_size()[0]))
    for mod in ("pydantic", "numpy", "hjson", "psutil"):
        try:
            rows.append((f"{mod} version", importlib.import_module(mod).__version__))
        except Exception:
            rows.append((f"{mod} version", "not installed"))
    print("DeepSpeed general environment info:")
    for name, value in rows:
        print(name, "." * (38 - len(name)), value)


def parse_arguments():
    p = argparse.ArgumentParser()
    p.add_argument("--hide_operator_status", action="store_true")
    p.add_argument("--hide_errors_and_warnings", action="store_true")
    return p.parse_args()


def main(hide_operator_status=False, hide_errors_and_warnings=False):
    if not hide_operator_status:
        op_report(verbose=not hide_errors_and_warnings)
    debug_report()


def cli_main():
    a = parse_arguments()
    main(hide_operator_status=a.hide_operator_status, hide_errors_and_warnings=a.hide_errors_and_warnings)


if __name__ == "__main__":
    cli_main()


def ninja_installed() -> bool:
    try:
        import ninja  # noqa: F401
        return True
    except ImportError:
        return False


def human_readable_size(size) -> str:
    """``1536 -> '1.50 KB'`` (powers of 1024)."""
    units = ["B", "KB", "MB", "GB", "TB", "PB"]
    size = float(size)
    i = 0
    while size >= 1024 and i < len(units) - 1:
        size /= 1024
        i += 1
    return f"{size:.2f} {units[i]}"


def installed_cann_path():
    """Ascend toolkit location — never present on this platform."""
    return None
