"""torch version shims (reference: deepspeed/utils/torch.py:25-31)."""
from packaging import version as pkg_version

import torch


def required_torch_version(min_version=None, max_version=None):
    v = pkg_version.parse(torch.__version__.split("+")[0])
    if min_version is not None and v < pkg_version.parse(str(min_version)):
        return False
    if max_version is not None and v > pkg_version.parse(str(max_version)):
        return False
    return True


def register_grad_hook(param, hook):
    """Run ``hook(param)`` after the gradient has been accumulated into ``param.grad``."""
    if hasattr(param, "register_post_accumulate_grad_hook"):
        return param.register_post_accumulate_grad_hook(hook)
    tmp = param.expand_as(param)
    acc = tmp.grad_fn.next_functions[0][0]
    return acc.register_hook(lambda *_: hook(param))
