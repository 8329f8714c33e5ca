"""Functional communication façade.

Same call surface as the reference's ``deepspeed.comm`` (deepspeed/comm/comm.py: ``init_distributed``
:643-709, the ``@timed_op`` wrapper :102-137, every collective, ``initialize_mesh_device`` :616,
``enable_symm_mem_for_group`` :631, capability probes) but with exactly one backend:
``torch.distributed`` — NCCL over NVLink/NVSwitch on B200, gloo for CPU-only tests.  The fused
compute+collective kernels do NOT go through here; they use the symmetric heap in
``comm/symm.py``.  This layer is the bootstrap / scalar / pipeline-p2p plumbing and the baseline
path every fused op is measured against.

Kill switches (``SXB_COMM_*_OFF`` env vars, also ``DS_COMM_*_OFF`` for compatibility with
deepspeed/comm/torch.py:16-87) turn a class of collectives into no-ops so a step can be timed
compute-only — the mechanism used to measure exposed communication.
"""
from __future__ import annotations

import datetime
import os
import time
from typing import List, Optional

import torch
import torch.distributed as tdist

from ..device import get_accelerator
from ..utils.comms_logging import CommsLogger
from ..utils.logging import logger
from .reduce_op import ReduceOp

default_pg_timeout = datetime.timedelta(minutes=int(os.getenv("DEEPSPEED_TIMEOUT", os.getenv("SXB_TIMEOUT", 30))))

comms_logger = CommsLogger()
_timers_enabled = False
_mesh_device = None
_initialized_here = False


class Noop:
    """Handle returned by disabled collectives."""

    def wait(self):
        return None

    def is_completed(self):
        return True


def _off(kind: str) -> bool:
    for prefix in ("SXB_COMM_", "DS_COMM_"):
        v = os.environ.get(f"{prefix}{kind}_OFF")
        if v and v not in ("0", "false", "False"):
            return True
    return False


def _op(op):
    if isinstance(op, ReduceOp):
        return {ReduceOp.SUM: tdist.ReduceOp.SUM, ReduceOp.PRODUCT: tdist.ReduceOp.PRODUCT,
                ReduceOp.MIN: tdist.ReduceOp.MIN, ReduceOp.MAX: tdist.ReduceOp.MAX,
                ReduceOp.BAND: tdist.ReduceOp.BAND, ReduceOp.BOR: tdist.ReduceOp.BOR,
                ReduceOp.BXOR: tdist.ReduceOp.BXOR, ReduceOp.AVG: tdist.ReduceOp.AVG}[op]
    return op


def _msg_size(args, kwargs) -> int:
    for a in list(args) + list(kwargs.values()):
        if torch.is_tensor(a):
            return a.numel() * a.element_size()
        if isinstance(a, (list, tuple)) and a and torch.is_tensor(a[0]):
            return sum(t.numel() * t.element_size() for t in a)
    return 0


def timed_op(func):
    """Record latency/size of a collective when the comms logger is enabled (synchronising, as
    the reference does) — zero overhead otherwise."""
    name = func.__name__

    def wrapper(*args, **kwargs):
        prof = comms_logger.enabled and (comms_logger.prof_all or name in comms_logger.prof_ops
                                         or kwargs.get("prof", False))
        log_name = kwargs.pop("log_name", name)
        kwargs.pop("prof", None)
        kwargs.pop("debug", None)
        if not prof:
            return func(*args, **kwargs)
        acc = get_accelerator()
        acc.synchronize()
        t0 = time.perf_counter()
        out = func(*args, **kwargs)
        if hasattr(out, "wait") and kwargs.get("async_op"):
            out.wait()
        acc.synchronize()
        ms = (time.perf_counter() - t0) * 1e3
        comms_logger.append(name, log_name, ms, _msg_size(args, kwargs),
                            get_world_size(kwargs.get("group")))
        return out

    wrapper.__name__ = name
    wrapper.__doc__ = func.__doc__
    return wrapper


# ------------------------------------------------------------------------------ lifecycle
def is_initialized() -> bool:
    return tdist.is_available() and tdist.is_initialized()


def is_available() -> bool:
    return tdist.is_available()


def get_backend(group=None):
    return tdist.get_backend(group)


def mpi_discovery(distributed_port=29500, verbose=True):
    """Populate RANK/WORLD_SIZE/MASTER_* from an MPI launch (mpi4py optional)."""
    try:
        from mpi4py import MPI
    except ImportError:
        for k_rank, k_size, k_local in (("OMPI_COMM_WORLD_RANK", "OMPI_COMM_WORLD_SIZE", "OMPI_COMM_WORLD_LOCAL_RANK"),
                                        ("PMI_RANK", "PMI_SIZE", "MPI_LOCALRANKID"),
                                        ("SLURM_PROCID", "SLURM_NTASKS", "SLURM_LOCALID")):
            if k_rank in os.environ:
                os.environ.setdefault("RANK", os.environ[k_rank])
                os.environ.setdefault("WORLD_SIZE", os.environ[k_size])
                os.environ.setdefault("LOCAL_RANK", os.environ.get(k_local, "0"))
                os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
                os.environ.setdefault("MASTER_PORT", str(distributed_port))
                return
        return
    comm = MPI.COMM_WORLD
    rank, world = comm.Get_rank(), comm.Get_size()
    import socket
    master = comm.bcast(socket.gethostbyname(socket.gethostname()) if rank == 0 else None, root=0)
    names = comm.allgather(MPI.Get_processor_name())
    local_rank = sum(1 for n in names[:rank] if n == names[rank])
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(local_rank),
                      MASTER_ADDR=master, MASTER_PORT=str(distributed_port))
    if verbose:
        logger.info(f"MPI discovery: rank={rank} local_rank={local_rank} world={world} master={master}")


def init_distributed(dist_backend: Optional[str] = None, auto_mpi_discovery: bool = True,
                     distributed_port: int = 29500, verbose: bool = True, timeout=default_pg_timeout,
                     init_method: Optional[str] = None, dist_init_required: Optional[bool] = None,
                     config=None, rank: int = -1, world_size: int = -1):
    """Bring up the default process group (idempotent)."""
    global _initialized_here
    if config is not None:
        configure(config)
    if dist_init_required is None:
        dist_init_required = not is_initialized()
    if dist_init_required is False or is_initialized():
        return
    need = ("RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT")
    if auto_mpi_discovery and not all(k in os.environ for k in need):
        mpi_discovery(distributed_port, verbose)
    if not all(k in os.environ for k in need) and init_method is None and rank < 0:
        # single process convenience (reference: single_gpu_mode, comm/torch.py:114-119)
        os.environ.setdefault("RANK", "0")
        os.environ.setdefault("WORLD_SIZE", "1")
        os.environ.setdefault("LOCAL_RANK", "0")
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", str(distributed_port))
    acc = get_accelerator()
    backend = dist_backend or acc.communication_backend_name()
    if acc.is_available():
        acc.set_device(int(os.environ.get("LOCAL_RANK", "0")) % max(1, acc.device_count()))
    if verbose and int(os.environ.get("RANK", "0")) == 0:
        logger.info(f"Initializing torch.distributed with backend: {backend}")
    kw = dict(backend=backend, timeout=timeout)
    if init_method is not None:
        kw["init_method"] = init_method
    if rank >= 0 and world_size > 0:
        kw.update(rank=rank, world_size=world_size)
    if backend == "nccl" and acc.is_available():
        try:
            kw["device_id"] = torch.device("cuda", acc.current_device())
        except Exception:
            pass
    tdist.init_process_group(**kw)
    _initialized_here = True


def destroy_process_group(group=None):
    if is_initialized():
        tdist.destroy_process_group(group)


def configure(ds_config=None, enabled=None, prof_all=None, prof_ops=None, verbose=None, debug=None):
    if ds_config is not None and getattr(ds_config, "comms_config", None) is not None:
        comms_logger.configure(ds_config.comms_config.comms_logger)
    if enabled is not None:
        comms_logger.enabled = enabled
    if prof_all is not None:
        comms_logger.prof_all = prof_all
    if prof_ops is not None:
        comms_logger.prof_ops = prof_ops
    if verbose is not None:
        comms_logger.verbose = verbose
    if debug is not None:
        comms_logger.debug = debug


def log_summary(show_straggler=False):
    if is_initialized():
        barrier()
    rows = None
    if get_rank() == 0:
        rows = comms_logger.log_all(print_log=True, show_straggler=False)
    if show_straggler:
        comms_logger.log_all(print_log=get_rank() == 0, show_straggler=True)
    if is_initialized():
        barrier()
    return rows


# ------------------------------------------------------------------------------- topology
def get_rank(group=None) -> int:
    return tdist.get_rank(group) if is_initialized() else 0


def get_world_size(group=None) -> int:
    return tdist.get_world_size(group) if is_initialized() else 1


def get_local_rank() -> int:
    return int(os.environ.get("LOCAL_RANK", "0"))


def get_global_rank(group=None, group_rank=0) -> int:
    if group is None or not is_initialized():
        return group_rank
    return tdist.get_global_rank(group, group_rank)


def get_all_ranks_from_group(group=None) -> List[int]:
    if not is_initialized():
        return [0]
    if group is None:
        return list(range(get_world_size()))
    return list(tdist.get_process_group_ranks(group))


def get_world_group():
    return tdist.group.WORLD if is_initialized() else None


def new_group(ranks=None, **kw):
    return tdist.new_group(ranks=ranks, **kw)


def initialize_mesh_device(mesh_shape, mesh_dim_names):
    """(data_parallel, sequence_parallel) device mesh (reference: comm/torch.py:403-414)."""
    global _mesh_device
    from torch.distributed.device_mesh import init_device_mesh
    dev = "cuda" if get_accelerator().is_available() else "cpu"
    _mesh_device = init_device_mesh(dev, tuple(mesh_shape), mesh_dim_names=tuple(mesh_dim_names))
    return _mesh_device


def get_mesh_device():
    return _mesh_device


def enable_symm_mem_for_group(group_name: str):
    from torch.distributed import _symmetric_memory as symm
    return symm.enable_symm_mem_for_group(group_name)


# capability probes (reference: comm.py has_* helpers)
def has_all_gather_into_tensor() -> bool:
    return hasattr(tdist, "all_gather_into_tensor")


def has_reduce_scatter_tensor() -> bool:
    return hasattr(tdist, "reduce_scatter_tensor")


def has_coalescing_manager() -> bool:
    return hasattr(tdist.distributed_c10d, "_coalescing_manager")


def has_all_reduce_coalesced() -> bool:
    return hasattr(tdist, "all_reduce_coalesced")


# ----------------------------------------------------------------------------- collectives
@timed_op
def all_reduce(tensor, op=ReduceOp.SUM, group=None, async_op=False):
    if _off("ALL_REDUCE") or not is_initialized():
        return Noop()
    return tdist.all_reduce(tensor, op=_op(op), group=group, async_op=async_op)


@timed_op
def inference_all_reduce(tensor, op=ReduceOp.SUM, group=None):
    """Latency-oriented all-reduce of the inference paths.  CPU tensors of a single-host job go through the
    shared-memory engine (``comm/shm.py``, reference ``comm/torch.py:166-171``); everything else is ``all_reduce``."""
    if _off("ALL_REDUCE") or not is_initialized():
        return tensor
    if tensor.device.type == "cpu" and op == ReduceOp.SUM and (group is None or group is tdist.group.WORLD) \
            and os.environ.get("SXB_SHM_ALLREDUCE", "1") != "0":
        from . import shm
        if shm.all_ranks_on_this_host() and shm.ShmAllReduce.supports(tensor):
            eng = shm.get_shm_comm()
            if not eng.initialized:
                eng.initialize(tdist.get_world_size(), tdist.get_rank())
            eng.all_reduce(tensor)
            return tensor
    tdist.all_reduce(tensor, op=_op(op), group=group)
    return tensor


@timed_op
def all_reduce_coalesced(tensors, op=ReduceOp.SUM, group=None, async_op=False):
    if _off("ALL_REDUCE") or not is_initialized():
        return Noop()
    if len(tensors) == 0:
        return Noop()
    flat = torch._utils._flatten_dense_tensors(tensors)
    h = tdist.all_reduce(flat, op=_op(op), group=group, async_op=False)
    for t, s in zip(tensors, torch._utils._unflatten_dense_tensors(flat, tensors)):
        t.copy_(s)
    return h or Noop()


@timed_op
def reduce(tensor, dst, op=ReduceOp.SUM, group=None, async_op=False):
    if _off("REDUCE") or not is_initialized():
        return Noop()
    return tdist.reduce(tensor, dst=dst, op=_op(op), group=group, async_op=async_op)


@timed_op
def reduce_scatter(output, input_list, op=ReduceOp.SUM, group=None, async_op=False):
    if _off("REDUCE_SCATTER") or not is_initialized():
        if not is_initialized():
            output.copy_(input_list[0])
        return Noop()
    return tdist.reduce_scatter(output, input_list, op=_op(op), group=group, async_op=async_op)


@timed_op
def reduce_scatter_tensor(output_tensor, tensor, op=ReduceOp.SUM, group=None, async_op=False):
    if _off("REDUCE_SCATTER"):
        return Noop()
    if not is_initialized() or get_world_size(group) == 1 and tdist.get_backend(group) == "gloo":
        output_tensor.copy_(tensor.reshape(-1)[: output_tensor.numel()].view_as(output_tensor))
        return Noop()
    if tdist.get_backend(group) == "gloo":
        # gloo has no reduce_scatter: all_reduce then slice (test plumbing only)
        n, r = get_world_size(group), get_rank(group)
        tmp = tensor.clone()
        tdist.all_reduce(tmp, op=_op(op) if _op(op) != tdist.ReduceOp.AVG else tdist.ReduceOp.SUM, group=group)
        if _op(op) == tdist.ReduceOp.AVG:
            tmp /= n
        output_tensor.copy_(tmp.reshape(n, -1)[r].view_as(output_tensor))
        return Noop()
    return tdist.reduce_scatter_tensor(output_tensor, tensor, op=_op(op), group=group, async_op=async_op)


reduce_scatter_fn = reduce_scatter_tensor


@timed_op
def all_gather(tensor_list, tensor, group=None, async_op=False):
    if _off("ALL_GATHER"):
        return Noop()
    if not is_initialized():
        tensor_list[0].copy_(tensor)
        return Noop()
    return tdist.all_gather(tensor_list, tensor, group=group, async_op=async_op)


@timed_op
def all_gather_into_tensor(output_tensor, tensor, group=None, async_op=False):
    if _off("ALL_GATHER"):
        return Noop()
    if not is_initialized():
        output_tensor.reshape(-1)[: tensor.numel()].copy_(tensor.reshape(-1))
        return Noop()
    return tdist.all_gather_into_tensor(output_tensor, tensor, group=group, async_op=async_op)


allgather_fn = all_gather_into_tensor


def all_gather_coalesced(output_tensors, input_tensors, group=None, async_op=False):
    """output_tensors[i] (world * numel_i) <- gather(input_tensors[i]); one launch under NCCL."""
    if _off("ALL_GATHER"):
        return Noop()
    handles = []
    for o, i in zip(output_tensors, input_tensors):
        handles.append(all_gather_into_tensor(o, i, group=group, async_op=async_op))
    return handles[-1] if handles else Noop()


def all_gather_object(object_list, obj, group=None):
    if not is_initialized():
        object_list[0] = obj
        return
    return tdist.all_gather_object(object_list, obj, group=group)


@timed_op
def all_to_all_single(output, tensor, output_split_sizes=None, input_split_sizes=None, group=None,
                      async_op=False):
    if _off("ALL_TO_ALL"):
        return Noop()
    if not is_initialized() or get_world_size(group) == 1:
        output.copy_(tensor.view_as(output))
        return Noop()
    if tdist.get_backend(group) == "gloo":
        n = get_world_size(group)
        ins = list(tensor.split(input_split_sizes) if input_split_sizes else tensor.chunk(n))
        outs = list(output.split(output_split_sizes) if output_split_sizes else output.chunk(n))
        ins = [t.contiguous() for t in ins]
        outs_c = [torch.empty_like(o) for o in outs]
        _gloo_all_to_all(outs_c, ins, group)
        for o, c in zip(outs, outs_c):
            o.copy_(c)
        return Noop()
    return tdist.all_to_all_single(output, tensor, output_split_sizes, input_split_sizes, group=group,
                                   async_op=async_op)


def _gloo_all_to_all(outs, ins, group):
    n, r = get_world_size(group), get_rank(group)
    ranks = get_all_ranks_from_group(group)
    reqs = []
    for peer in range(n):
        if peer == r:
            outs[peer].copy_(ins[peer])
        else:
            reqs.append(tdist.isend(ins[peer], ranks[peer], group=group))
            reqs.append(tdist.irecv(outs[peer], ranks[peer], group=group))
    for q in reqs:
        q.wait()


@timed_op
def all_to_all(output_tensor_list, input_tensor_list, group=None, async_op=False):
    if _off("ALL_TO_ALL"):
        return Noop()
    if not is_initialized() or get_world_size(group) == 1:
        for o, i in zip(output_tensor_list, input_tensor_list):
            o.copy_(i)
        return Noop()
    if tdist.get_backend(group) == "gloo":
        _gloo_all_to_all(output_tensor_list, [t.contiguous() for t in input_tensor_list], group)
        return Noop()
    return tdist.all_to_all(output_tensor_list, input_tensor_list, group=group, async_op=async_op)


@timed_op
def broadcast(tensor, src, group=None, async_op=False):
    if _off("BROADCAST") or not is_initialized():
        return Noop()
    return tdist.broadcast(tensor, src=src, group=group, async_op=async_op)


def broadcast_object_list(object_list, src, group=None, device=None):
    if not is_initialized():
        return
    return tdist.broadcast_object_list(object_list, src=src, group=group, device=device)


@timed_op
def send(tensor, dst, group=None, tag=0):
    return tdist.send(tensor, dst=dst, group=group, tag=tag)


@timed_op
def recv(tensor, src=None, group=None, tag=0):
    return tdist.recv(tensor, src=src, group=group, tag=tag)


@timed_op
def isend(tensor, dst, group=None, tag=0):
    return tdist.isend(tensor, dst=dst, group=group, tag=tag)


@timed_op
def irecv(tensor, src=None, group=None, tag=0):
    return tdist.irecv(tensor, src=src, group=group, tag=tag)


@timed_op
def gather(tensor, gather_list=None, dst=0, group=None, async_op=False):
    if not is_initialized():
        if gather_list:
            gather_list[0].copy_(tensor)
        return Noop()
    return tdist.gather(tensor, gather_list=gather_list, dst=dst, group=group, async_op=async_op)


@timed_op
def scatter(tensor, scatter_list=None, src=0, group=None, async_op=False):
    if not is_initialized():
        if scatter_list:
            tensor.copy_(scatter_list[0])
        return Noop()
    return tdist.scatter(tensor, scatter_list=scatter_list, src=src, group=group, async_op=async_op)


@timed_op
def barrier(group=None, async_op=False, device_ids=None):
    if not is_initialized():
        return Noop()
    return tdist.barrier(group=group, async_op=async_op)


def monitored_barrier(group=None, timeout=None, wait_all_ranks=False):
    if not is_initialized():
        return
    if tdist.get_backend(group) == "gloo":
        return tdist.monitored_barrier(group=group, timeout=timeout, wait_all_ranks=wait_all_ranks)
    return tdist.barrier(group=group)


# ---- backend bookkeeping and cloud-launcher environment patches (reference ``comm/comm.py:42-120,760-840``)
ProcessGroup = tdist.ProcessGroup
cdb = None                      # the active TorchBackend object once ``init_deepspeed_backend`` / ``set_backend`` ran


def set_backend():
    """Make sure ``cdb`` names the backend object matching the running process group."""
    global cdb
    if cdb is None and tdist.is_initialized():
        from .torch import TorchBackend
        cdb = TorchBackend(backend=tdist.get_backend())
    return cdb


def init_deepspeed_backend(ds_backend, timeout=None, init_method=None):
    """Select a non-default transport by name.  This platform has one transport for plumbing (torch.distributed over
    NCCL / gloo); ``ccl`` / ``hccl`` / ``mpi`` are other vendors' or launchers' and report that they are unavailable."""
    global cdb
    name = str(ds_backend).lower()
    if name in ("nccl", "gloo"):
        from .torch import TorchBackend
        cdb = TorchBackend(backend=name, timeout=timeout, init_method=init_method)
        return cdb
    if name in ("ccl", "hccl", "sccl", "mpi"):
        logger.debug(f"communication backend {name!r} is not available on this platform")
        return None
    raise ValueError(f"unknown communication backend {ds_backend!r}")


def in_aml() -> bool:
    """Running as an Azure Machine Learning job."""
    return "AZUREML_EXPERIMENT_ID" in os.environ


def in_aws_sm() -> bool:
    """Running as an AWS SageMaker job."""
    return "SM_TRAINING_ENV" in os.environ


def in_dlts() -> bool:
    """Running on a DLTS cluster."""
    return "DLTS_JOB_ID" in os.environ


def patch_aml_env_for_torch_nccl_backend(master_port=6105, verbose=True):
    """AzureML exports Open MPI variables and its own master-node name; translate to torch.distributed's."""
    os.environ["RANK"] = os.environ["OMPI_COMM_WORLD_RANK"]
    os.environ["WORLD_SIZE"] = os.environ["OMPI_COMM_WORLD_SIZE"]
    single_node = int(os.environ["OMPI_COMM_WORLD_LOCAL_SIZE"]) == int(os.environ["WORLD_SIZE"])
    if not single_node:
        master = os.environ["AZ_BATCH_MASTER_NODE"].split(":")
        os.environ["MASTER_ADDR"] = master[0]
        os.environ.setdefault("MASTER_PORT", str(master_port))
    else:
        os.environ["MASTER_ADDR"] = os.environ["AZ_BATCHAI_MPI_MASTER_NODE"]
        os.environ["MASTER_PORT"] = os.environ.get("MASTER_PORT", "54965")
    os.environ["LOCAL_RANK"] = os.environ["OMPI_COMM_WORLD_LOCAL_RANK"]
    os.environ.setdefault("NCCL_SOCKET_IFNAME", "^docker0,lo")
    if verbose:
        logger.info("NCCL_SOCKET_IFNAME original value = {}".format(os.environ["NCCL_SOCKET_IFNAME"]))
        logger.info("Discovered AzureML settings of world_rank={}, local_rank={}, world_size={}, master_addr={}, "
                    "master_port={}".format(os.environ["RANK"], os.environ["LOCAL_RANK"], os.environ["WORLD_SIZE"],
                                            os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"]))


def patch_aws_sm_env_for_torch_nccl_backend(verbose=True):
    """SageMaker also starts ranks through Open MPI; MASTER_ADDR / MASTER_PORT are already exported."""
    os.environ["RANK"] = os.environ["OMPI_COMM_WORLD_RANK"]
    os.environ["LOCAL_RANK"] = os.environ["OMPI_COMM_WORLD_LOCAL_RANK"]
    os.environ["WORLD_SIZE"] = os.environ["OMPI_COMM_WORLD_SIZE"]
    if verbose:
        logger.info("Discovered AWS SageMaker settings of world_rank={}, local_rank={}, world_size={}, master_addr={}, "
                    "master_port={}".format(os.environ["RANK"], os.environ["LOCAL_RANK"], os.environ["WORLD_SIZE"],
                                            os.environ.get("MASTER_ADDR"), os.environ.get("MASTER_PORT")))
