"""Reference: ``module_inject/utils.py`` — policy → container lookup used by ``replace_transformer_layer``."""
from __future__ import annotations

from .containers import POLICY_TO_CONTAINER
from .containers.base import TransformerPolicy


def policy_to_ds_container(**kwargs):
    """``policy_to_ds_container(policy=<policy instance>, config=..., model_config=..., layer_id=..., child=...)`` →
    the architecture's container; an unknown policy type is an error (the reference logs and returns ``None``)."""
    policy = kwargs.get("policy")
    if policy is None:
        raise ValueError("policy_to_ds_container needs policy=<TransformerPolicy instance>")
    for pol_cls, container_cls in POLICY_TO_CONTAINER.items():
        if type(policy) is pol_cls:
            return container_cls(policy, kwargs.get("config"), kwargs.get("model_config"), kwargs.get("layer_id", 0),
                                 kwargs.get("child"))
    raise ValueError(f"no container registered for policy type {type(policy).__name__}")


def get_transformer_name(replaced_module) -> str:
    from .auto_tp import get_transformer_name as _g
    return _g(replaced_module)
