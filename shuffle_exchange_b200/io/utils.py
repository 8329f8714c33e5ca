"""Serialisation helpers (reference: ``deepspeed/io/utils.py``)."""
import numpy
import torch


def required_minimum_torch_version(major_version=1, minor_version=12):
    major, minor = (int(x) for x in torch.__version__.split(".")[:2])
    return (major, minor) >= (major_version, minor_version)


def tensor_to_bytes(tensor: torch.Tensor):
    """Raw bytes of a CPU tensor without an intermediate copy where numpy can view the dtype."""
    t = tensor.detach().contiguous()
    return t.view(torch.uint8).numpy().data if t.dtype == torch.bfloat16 else t.numpy().data


def bytes_to_tensor(buffer):
    return torch.from_numpy(numpy.array(numpy.frombuffer(buffer, dtype=numpy.uint8)))


def obj_serialization_details(obj):
    """(number of tensors, total tensor bytes) inside a nested container — sizing information for IO buffers."""
    count = size = 0
    stack = [obj]
    while stack:
        o = stack.pop()
        if torch.is_tensor(o):
            count += 1
            size += o.numel() * o.element_size()
        elif isinstance(o, dict):
            stack.extend(o.values())
        elif isinstance(o, (list, tuple)):
            stack.extend(o)
    return count, size


serialize_details = obj_serialization_details      # older name used by callers of the reference module
