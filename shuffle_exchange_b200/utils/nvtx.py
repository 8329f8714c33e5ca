"""NVTX range decorator (reference: deepspeed/utils/nvtx.py:12-25)."""
import functools

from ..device import get_accelerator

enable_nvtx = True


def instrument_w_nvtx(func):
    @functools.wraps(func)
    def wrapped(*args, **kwargs):
        if enable_nvtx and get_accelerator().is_available():
            get_accelerator().range_push(func.__qualname__)
            try:
                return func(*args, **kwargs)
            finally:
                get_accelerator().range_pop()
        return func(*args, **kwargs)
    return wrapped
