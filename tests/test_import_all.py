"""Every module of the package imports on a machine without a GPU (dangling relative imports, missing names and
modules that touch CUDA at import time show up here rather than on the GPU box)."""
import importlib
import pkgutil


def test_every_module_imports():
    import shuffle_exchange_b200 as pkg
    bad = []
    for m in pkgutil.walk_packages(pkg.__path__, pkg.__name__ + "."):
        if m.name.rsplit(".", 1)[-1].startswith("libsxb200"):
            continue                                  # the shared objects themselves, loaded through ctypes
        try:
            importlib.import_module(m.name)
        except Exception as e:                        # noqa: BLE001 — the point is to list them all
            bad.append((m.name, f"{type(e).__name__}: {e}"[:160]))
    assert not bad, bad
