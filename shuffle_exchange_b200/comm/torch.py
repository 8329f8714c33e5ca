"""``TorchBackend``: the torch.distributed transport as an object (reference: ``deepspeed/comm/torch.py:97-420``).

Every method forwards to the module-level wrapper in ``comm.comm`` (which adds the comms logger timing and the gloo
fall-backs), so code written against the reference's ``cdb`` object and code using the functions behave identically."""
from __future__ import annotations

import os

import torch
import torch.distributed as tdist

from . import comm as _c
from .backend import Backend
from .constants import NCCL_BACKEND
from .reduce_op import ReduceOp

DS_COMM_ALL_GATHER_OFF = False
DS_COMM_REDUCE_SCATTER_OFF = False
DS_COMM_BROADCAST_OFF = False
DS_COMM_ALL_REDUCE_OFF = False
DS_COMM_REDUCE_OFF = False


def disable_compiler_collective(func):
    """Kept for API parity: collectives here are never traced by a compiler."""
    return func


def build_shm_op():
    """Shared-memory all-reduce engine for CPU tensors (``comm/shm.py`` over ``csrc/host/shm_comm.cpp``); ``None`` when
    the host library is unavailable."""
    try:
        from .shm import get_shm_comm
        return get_shm_comm()
    except Exception:
        return None


def has_coalescing_manager() -> bool:
    return hasattr(tdist.distributed_c10d, "_coalescing_manager")


def has_all_reduce_coalesced() -> bool:
    return hasattr(tdist, "all_reduce_coalesced")


def get_coalescing_manager(group, device, reqs, async_op):
    return tdist.distributed_c10d._coalescing_manager(group, device=device, async_ops=async_op)


class TorchBackend(Backend):

    def __init__(self, backend=NCCL_BACKEND, timeout=None, init_method=None, rank=-1, world_size=-1, name="torch"):
        super().__init__(name=name)
        self.shm_comm_op = build_shm_op()
        self.has_all_reduce_coalesced = has_all_reduce_coalesced()
        self.has_coalescing_manager = has_coalescing_manager()
        self.all_gather_function = getattr(tdist, "all_gather_into_tensor", None)
        self.reduce_scatter_function = getattr(tdist, "reduce_scatter_tensor", None)
        self.using_mpi = False
        self.single_gpu_mode = True
        self.init_process_group(backend, timeout, init_method, rank, world_size)

    @classmethod
    def get_all_gather_function(cls):
        return getattr(tdist, "all_gather_into_tensor", getattr(tdist, "_all_gather_base", None))

    @classmethod
    def get_reduce_scatter_function(cls):
        return getattr(tdist, "reduce_scatter_tensor", getattr(tdist, "_reduce_scatter_base", None))

    def has_all_gather_into_tensor(self): return self.all_gather_function is not None
    def has_reduce_scatter_tensor(self): return self.reduce_scatter_function is not None

    def init_process_group(self, backend=NCCL_BACKEND, timeout=None, init_method=None, rank=-1, world_size=-1):
        if not tdist.is_initialized():
            kw = {}
            if timeout is not None:
                kw["timeout"] = timeout
            if init_method is not None:
                kw["init_method"] = init_method
            if rank >= 0 and world_size > 0:
                kw.update(rank=rank, world_size=world_size)
            tdist.init_process_group(backend, **kw)
        self.using_mpi = tdist.get_backend() == "mpi"
        self.world_group = tdist.group.WORLD
        self.world_rank, self.world_size = tdist.get_rank(), tdist.get_world_size()
        self.initialized = True

    # ---- collectives
    def all_reduce(self, tensor, op=ReduceOp.SUM, group=None, async_op=False): return _c.all_reduce(tensor, op, group, async_op)
    def inference_all_reduce(self, tensor, op=ReduceOp.SUM, group=None): return _c.inference_all_reduce(tensor, op, group)
    def all_reduce_coalesced(self, tensors, op=ReduceOp.SUM, group=None, async_op=False):
        return _c.all_reduce_coalesced(tensors, op, group, async_op)
    def reduce(self, tensor, dst, op=ReduceOp.SUM, group=None, async_op=False): return _c.reduce(tensor, dst, op, group, async_op)
    def reduce_scatter(self, output, input_list, op=ReduceOp.SUM, group=None, async_op=False):
        return _c.reduce_scatter(output, input_list, op, group, async_op)
    def broadcast(self, tensor, src, group=None, async_op=False): return _c.broadcast(tensor, src, group, async_op)
    def broadcast_object_list(self, object_list, src, group=None, device=None):
        return _c.broadcast_object_list(object_list, src, group, device)
    def all_gather(self, tensor_list, tensor, group=None, async_op=False): return _c.all_gather(tensor_list, tensor, group, async_op)
    def all_gather_into_tensor(self, output_tensor, input_tensor, group=None, async_op=False):
        return _c.all_gather_into_tensor(output_tensor, input_tensor, group, async_op)
    def all_gather_base(self, output_tensor, input_tensor, group=None, async_op=False):
        return _c.all_gather_into_tensor(output_tensor, input_tensor, group, async_op)
    def all_gather_coalesced(self, output_tensors, input_tensors, group=None, async_op=False):
        return _c.all_gather_coalesced(output_tensors, input_tensors, group, async_op)
    def all_gather_object(self, object_list, obj, group=None): return _c.all_gather_object(object_list, obj, group)
    def reduce_scatter_tensor(self, output_tensor, input_tensor, op=ReduceOp.SUM, group=None, async_op=False):
        return _c.reduce_scatter_tensor(output_tensor, input_tensor, op, group, async_op)
    def all_to_all_single(self, output, input, output_split_sizes=None, input_split_sizes=None, group=None, async_op=False):
        return _c.all_to_all_single(output, input, output_split_sizes, input_split_sizes, group, async_op)
    def all_to_all(self, output_tensor_list, input_tensor_list, group=None, async_op=False):
        return _c.all_to_all(output_tensor_list, input_tensor_list, group, async_op)
    def send(self, tensor, dst, group=None, tag=0): return _c.send(tensor, dst, group, tag)
    def recv(self, tensor, src=None, group=None, tag=0): return _c.recv(tensor, src, group, tag)
    def isend(self, tensor, dst, group=None, tag=0): return _c.isend(tensor, dst, group, tag)
    def irecv(self, tensor, src=None, group=None, tag=0): return _c.irecv(tensor, src, group, tag)
    def gather(self, tensor, gather_list=None, dst=0, group=None, async_op=False):
        return _c.gather(tensor, gather_list, dst, group, async_op)
    def scatter(self, tensor, scatter_list=None, src=0, group=None, async_op=False):
        return _c.scatter(tensor, scatter_list, src, group, async_op)
    def barrier(self, group=None, async_op=False, device_ids=None): return _c.barrier(group, async_op, device_ids)
    def monitored_barrier(self, group=None, timeout=None, wait_all_ranks=False):
        return _c.monitored_barrier(group, timeout, wait_all_ranks)

    # ---- queries / groups
    def get_rank(self, group=None): return tdist.get_rank(group=group)
    def get_world_size(self, group=None): return tdist.get_world_size(group=group)
    def is_initialized(self): return tdist.is_initialized()
    def get_backend(self, group=None): return tdist.get_backend(group=group)
    def new_group(self, ranks): return tdist.new_group(ranks)
    def get_global_rank(self, group, group_rank): return _c.get_global_rank(group, group_rank)
    def get_world_group(self): return tdist.group.WORLD
    def destroy_process_group(self, group=None): return tdist.destroy_process_group(group=group)
    def get_all_ranks_from_group(self, group): return _c.get_all_ranks_from_group(group)
    def init_device_mesh(self, mesh_shape, mesh_dim_names):
        return _c.initialize_mesh_device(mesh_shape, mesh_dim_names)
    def enable_symm_mem_for_group(self, group_name): return _c.enable_symm_mem_for_group(group_name)

    def _reduce_op(self, op):
        return _c._op(op) if not isinstance(op, tdist.ReduceOp) else op


# ---- collective kill switches (reference ``comm/torch.py:16-87``): run a step with a class of collectives disabled to
# measure compute-only time.  They set the environment flags the façade (``comm.comm._off``) consults, so the switch is
# seen by every code path, and ``flag=False`` re-enables.
from .comm import Noop  # noqa: E402,F401


def _switch(kind: str, flag: bool) -> None:
    global DS_COMM_ALL_GATHER_OFF, DS_COMM_REDUCE_SCATTER_OFF, DS_COMM_BROADCAST_OFF, DS_COMM_ALL_REDUCE_OFF, \
        DS_COMM_REDUCE_OFF
    os.environ[f"DS_COMM_{kind}_OFF"] = "1" if flag else "0"
    globals()[f"DS_COMM_{kind}_OFF"] = bool(flag)


def all_gather_comm_off(flag=False):
    _switch("ALL_GATHER", flag)


def reduce_scatter_comm_off(flag=False):
    _switch("REDUCE_SCATTER", flag)


def broadcast_comm_off(flag=False):
    _switch("BROADCAST", flag)


def all_reduce_comm_off(flag=False):
    _switch("ALL_REDUCE", flag)


def reduce_comm_off(flag=False):
    _switch("REDUCE", flag)


def backward_comm_off(flag=False):
    """Everything the backward pass communicates: gradient reduce-scatter / all-reduce and the ZeRO-3 re-gather."""
    all_gather_comm_off(flag)
    reduce_scatter_comm_off(flag)
    all_reduce_comm_off(flag)
