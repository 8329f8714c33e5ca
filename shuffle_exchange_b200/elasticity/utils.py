"""(reference: ``deepspeed/elasticity/utils.py``)"""
import torch


def is_torch_elastic_compatible() -> bool:
    """torch.distributed.elastic ships with torch >= 1.11."""
    major, minor = (int(x) for x in torch.__version__.split(".")[:2])
    return (major, minor) >= (1, 11)
