"""``linear.Init``: build a model with every targeted ``nn.Linear`` replaced by an :class:`OptimizedLinear`
(reference: ``deepspeed/linear/context_manager.py:16-90``)."""
from __future__ import annotations

import torch.nn as nn

from .config import LoRAConfig, QuantizationConfig
from .optimized_linear import LoRAOptimizedLinear, OptimizedLinear

try:
    import transformers
except ImportError:  # pragma: no cover
    transformers = None


def init_lora(model):
    """Create the adapters of every ``LoRAOptimizedLinear`` built with ``delay_lora_init=True``."""
    model.requires_grad_(False)
    for m in model.modules():
        if isinstance(m, LoRAOptimizedLinear):
            m.init_lora()


class Init:
    """Inside the context, ``nn.Linear(in, out, bias=False)`` constructs an optimized layer; after model construction
    the layers whose names do not end with one of ``lora_config.target_mods`` are reverted to their frozen base."""

    def __init__(self, lora_config: LoRAConfig = None, quant_config: QuantizationConfig = None):
        from torch.nn.modules.linear import Linear as _TorchLinear
        self._orig_nn_linear = _TorchLinear
        self._orig_hooks = {}
        self.lora_config, self.quant_config = lora_config, quant_config
        self._post_init_hooked = False

    def __enter__(self):
        lora, quant, orig = self.lora_config, self.quant_config, self._orig_nn_linear

        class OptLinearWrapper:
            _orig_nn_linear = orig

            def __new__(cls, *args, **kwargs):
                kwargs = dict(kwargs)
                bias = kwargs.pop("bias", args[2] if len(args) > 2 else True)
                if bias:                                     # adapters only make sense for bias-free projections
                    return orig(*args[:2], bias=True, **{k: v for k, v in kwargs.items() if k in ("device", "dtype")})
                kw = {"dtype": kwargs["dtype"]} if "dtype" in kwargs and kwargs["dtype"] is not None else {}
                return OptimizedLinear(args[0], args[1], lora_config=lora, quantization_config=quant, **kw)

        nn.Linear = OptLinearWrapper
        if transformers is not None:
            outer = self

            def wrap(fn):
                def inner(*a, **k):
                    model = fn(*a, **k)
                    outer._finalize(model)
                    return model
                return inner
            for cls_name in ("AutoModelForCausalLM",):
                cls = getattr(transformers, cls_name, None)
                if cls is not None:
                    self._orig_hooks[cls] = (cls.from_pretrained, cls.from_config)
                    cls.from_pretrained = wrap(cls.from_pretrained)
                    cls.from_config = wrap(cls.from_config)
        return self

    def _finalize(self, model):
        targets = tuple(self.lora_config.target_mods) if self.lora_config is not None else ()
        for name, m in model.named_modules():
            if isinstance(m, LoRAOptimizedLinear) and not name.endswith(targets):
                m.disable()
        return model

    def __exit__(self, *exc):
        nn.Linear = self._orig_nn_linear
        for cls, (fp, fc) in self._orig_hooks.items():
            cls.from_pretrained, cls.from_config = fp, fc
        self._orig_hooks = {}
