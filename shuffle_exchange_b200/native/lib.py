"""ctypes binding of the C-ABI kernel libraries (declared in ``include/sxb200.h``).

One signature table instead of per-op pybind modules: ``p`` = ``void*``, ``i`` = ``int``,
``l`` = ``long long``, ``f`` = ``float``.  Every entry point returns an ``int`` status
(0 ok, >0 ``cudaError_t``, <0 argument error) that :func:`call` turns into an exception, so a
missing or failing kernel is loud — there is no silent PyTorch fallback on a GPU machine.
"""
from __future__ import annotations

import ctypes
import os
import threading
from pathlib import Path
from typing import Optional

import torch

_HERE = Path(__file__).resolve().parent
_LIB: Optional[ctypes.CDLL] = None
_HOST: Optional[ctypes.CDLL] = None
_LOCK = threading.Lock()

# number of native GPU entry-point calls since the last reset (bench.py reports this)
launch_count = 0

F32, F16, BF16, FP8E4M3, I8, U8, I32, I64 = range(8)
_DT = {
    torch.float32: F32, torch.float16: F16, torch.bfloat16: BF16, torch.int8: I8, torch.uint8: U8,
    torch.int32: I32, torch.int64: I64,
}
if hasattr(torch, "float8_e4m3fn"):
    _DT[torch.float8_e4m3fn] = FP8E4M3

_CT = {"p": ctypes.c_void_p, "i": ctypes.c_int, "l": ctypes.c_longlong, "f": ctypes.c_float,
       "d": ctypes.c_double}

# name -> argument signature
SIGNATURES = {
    "sxb_abi_version": "",
    "sxb_device_info": "pppp",
    "sxb_gemm_bf16": "pppiiiiiiiiifip",
    "sxb_adam_flat": "ppppp" "l" "ii" "fffff" "iii" "f" "ppp",
    "sxb_multi_tensor_adam": "ppii" "fffff" "iii" "pp",
    "sxb_lion_flat": "pppp" "l" "ii" "ffff" "f" "pp",
    "sxb_adagrad_flat": "pppp" "l" "ii" "fff" "f" "p",
    "sxb_lamb": "pppppppp" "l" "ii" "fffff" "ii" "fff" "p",
    "sxb_sumsq": "plipp",
    "sxb_norm_bwd_scratch_elems": "lii",
    "sxb_rmsnorm_fwd": "pppppp" "l" "i" "f" "i" "p",
    "sxb_rmsnorm_bwd": "pppppppp" "l" "iii" "p",
    "sxb_layernorm_fwd": "pppppppp" "l" "i" "f" "i" "p",
    "sxb_layernorm_bwd": "pppppppppp" "l" "iii" "p",
    "sxb_rope": "pppp" "l" "ii" "l" "ii" "p",
    "sxb_gated_act_fwd": "ppp" "l" "iii" "p",
    "sxb_gated_act_bwd": "ppp" "l" "iii" "p",
    "sxb_bias_act": "ppp" "l" "iii" "p",
    "sxb_bias_act_bwd": "pppp" "l" "iii" "p",
    "sxb_cross_entropy": "pppp" "l" "i" "l" "l" "p" "ii" "p",
    "sxb_embedding_fwd": "ppp" "l" "ii" "p",
    "sxb_embedding_bwd": "ppp" "l" "ii" "p",
    "sxb_scale_cast": "pp" "l" "ii" "f" "i" "p",
    "sxb_attn_fwd": "ppppp" "iiiii" "llll" "i" "f" "p",
    "sxb_attn_bwd": "pppppppppp" "iiiii" "llllllll" "i" "f" "p",
    "sxb_gemm_hang_info": "p",
    "sxb_gemm_set_mode": "i",
    "sxb_attn_hang_info": "p",
}
_LONG_RETURN = {"sxb_norm_bwd_scratch_elems"}


class NativeError(RuntimeError):
    pass


def lib_path() -> Path:
    # SXB_LIB_PATH: load an alternative build of the CUDA library (A/B experiments with different compile flags)
    alt = os.environ.get("SXB_LIB_PATH")
    return Path(alt) if alt else _HERE / "libsxb200.so"


def host_lib_path() -> Path:
    return _HERE / "libsxb200_host.so"


def available() -> bool:
    return lib_path().exists()


def _bind(lib: ctypes.CDLL, table) -> None:
    for name, sig in table.items():
        try:
            fn = getattr(lib, name)
        except AttributeError:
            continue
        fn.argtypes = [_CT[c] for c in sig]
        fn.restype = ctypes.c_longlong if name in _LONG_RETURN else ctypes.c_int


def load() -> ctypes.CDLL:
    """Load (building first if the sources are newer and a compiler is present)."""
    global _LIB
    if _LIB is not None:
        return _LIB
    with _LOCK:
        if _LIB is not None:
            return _LIB
        path = lib_path()
        if not path.exists() or os.environ.get("SXB_REBUILD") == "1":
            from . import build
            build.build_cuda()
        if not path.exists():
            raise NativeError(f"{path} is missing and could not be built; run `python -m "
                              "shuffle_exchange_b200.native.build`")
        lib = ctypes.CDLL(str(path), mode=ctypes.RTLD_GLOBAL)
        _bind(lib, SIGNATURES)
        _LIB = lib
        return lib


def register(table) -> None:
    """Let other modules (comm/symm, moe, inference) add entry-point signatures."""
    SIGNATURES.update(table)
    if _LIB is not None:
        _bind(_LIB, table)


def dtype_code(dt: torch.dtype) -> int:
    try:
        return _DT[dt]
    except KeyError as e:
        raise NativeError(f"dtype {dt} has no native kernel") from e


def ptr(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


def stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


def call(name: str, *args):
    """Invoke ``name`` and raise on a non-zero status."""
    global launch_count
    fn = getattr(load(), name)
    rc = fn(*args)
    launch_count += 1
    if name in _LONG_RETURN:
        return rc
    if rc != 0:
        if rc > 0:
            raise NativeError(f"{name} failed: cudaError {rc}")
        reason = {-1: "bad argument", -2: "unsupported shape/alignment", -3: "size beyond limit",
                  -4: "driver entry point unavailable"}.get(rc, "error")
        raise NativeError(f"{name} rejected the call: {reason} ({rc})")
    return 0


def hang_info(kind: str = "attn"):
    """Read-and-clear the pipeline watchdog record of a kernel family (``gemm`` | ``attn`` | ...).

    Returns ``None`` when no wait was abandoned since the last call, else a dict naming the
    barrier tag and the block/thread that gave up (see ``mbar_wait_to`` in csrc/sxb_common.cuh).
    """
    torch.cuda.synchronize()
    buf = (ctypes.c_uint * 8)()
    fn = getattr(load(), f"sxb_{kind}_hang_info")
    rc = fn(ctypes.cast(buf, ctypes.c_void_p))
    if rc != 0:
        raise NativeError(f"sxb_{kind}_hang_info failed: {rc}")
    if buf[0] == 0:
        return None
    return {"tag": buf[1], "block": (buf[2], buf[3], buf[4]), "thread": buf[5], "parity": buf[6], "aux": buf[7]}


def reset_launch_count() -> int:
    global launch_count
    n, launch_count = launch_count, 0
    return n


def use_native(t: torch.Tensor) -> bool:
    """True when ``t`` lives on a CUDA device (the native path is then mandatory)."""
    return t.is_cuda
