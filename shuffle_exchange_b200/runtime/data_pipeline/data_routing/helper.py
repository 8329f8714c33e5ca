"""Model surgery helpers for random-LTD (reference: data_routing/helper.py:11-46)."""
from .basic_layer import RandomLayerTokenDrop


def convert_to_random_ltd(model, convert_type):
    """Wrap every sub-module of class ``convert_type`` in :class:`RandomLayerTokenDrop` (in place)."""
    target = model.module if hasattr(model, "module") else model
    for name, child in list(target.named_children()):
        if isinstance(child, convert_type):
            setattr(target, name, RandomLayerTokenDrop(child))
        else:
            convert_to_random_ltd(child, convert_type)
    return model


def remove_random_ltd_state_dict(state_dict):
    """Strip the wrapper's ``.random_ltd_layer`` name component so checkpoints load into the plain model."""
    out = type(state_dict)()
    for key, value in state_dict.items():
        out[key.replace(".random_ltd_layer", "")] = value
    return out


def save_without_random_ltd(model):
    target = model.module if hasattr(model, "module") else model
    return remove_random_ltd_state_dict(target.state_dict())
