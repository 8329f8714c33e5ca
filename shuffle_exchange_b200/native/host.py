"""ctypes binding of ``libsxb200_host.so`` (CPU optimizers, async file IO, host schedulers)."""
from __future__ import annotations

import ctypes
import threading
from pathlib import Path

_HERE = Path(__file__).resolve().parent
_LIB = None
_LOCK = threading.Lock()

c_p, c_i, c_l, c_f = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_float

_SIGS = {
    "sxbh_create_adam": [c_i, c_f, c_f, c_f, c_f, c_f, c_i, c_i],
    "sxbh_destroy_adam": [c_i],
    "sxbh_adam_update": [c_i, c_l, c_f, c_f, c_f, c_f, c_f, c_i, c_p, c_p, c_p, c_p, c_l, c_p, c_i, c_f],
    "sxbh_create_adagrad": [c_i, c_f, c_f, c_f, c_i],
    "sxbh_destroy_adagrad": [c_i],
    "sxbh_adagrad_update": [c_i, c_l, c_f, c_f, c_f, c_p, c_p, c_p, c_l, c_p, c_i],
    "sxbh_create_lion": [c_i, c_f, c_f, c_f, c_f, c_i],
    "sxbh_destroy_lion": [c_i],
    "sxbh_lion_update": [c_i, c_l, c_f, c_f, c_f, c_f, c_p, c_p, c_p, c_l, c_p, c_i],
    "sxbh_bf16_to_f32": [c_p, c_p, c_l, c_f, c_i],
    "sxbh_num_threads": [],
    "sxbh_simd_width": [],
}


def register(table):
    _SIGS.update(table)
    if _LIB is not None:
        _bind(_LIB, table)


def _bind(lib, table):
    for name, args in table.items():
        try:
            fn = getattr(lib, name)
        except AttributeError:
            continue
        fn.argtypes = args
        fn.restype = getattr(fn, "_sxb_restype", ctypes.c_int)


def load():
    global _LIB
    if _LIB is not None:
        return _LIB
    with _LOCK:
        if _LIB is None:
            path = _HERE / "libsxb200_host.so"
            if not path.exists():
                from . import build
                build.build_host()
            lib = ctypes.CDLL(str(path))
            _bind(lib, _SIGS)
            _LIB = lib
    return _LIB
