"""``OnDevice`` — construct modules directly on a target device / as meta tensors, optionally casting
floating parameters (reference: deepspeed/utils/init_on_device.py:12-81).  Implemented with
``torch.device`` as a context manager plus a default-dtype switch instead of patching tensor factories."""
import torch


class OnDevice:
    def __init__(self, dtype=None, device="meta", enabled=True):
        self.dtype = dtype
        self.device = device
        self.enabled = enabled
        self._dev_ctx = None
        self._old_dtype = None

    def __enter__(self):
        if not self.enabled:
            return self
        self._dev_ctx = torch.device(self.device)
        self._dev_ctx.__enter__()
        if self.dtype is not None and self.dtype.is_floating_point:
            self._old_dtype = torch.get_default_dtype()
            torch.set_default_dtype(self.dtype)
        return self

    def __exit__(self, exc_type, exc_value, traceback):
        if not self.enabled:
            return False
        if self._old_dtype is not None:
            torch.set_default_dtype(self._old_dtype)
            self._old_dtype = None
        self._dev_ctx.__exit__(exc_type, exc_value, traceback)
        return False
