"""``MpiBackend``: the 1-bit all-reduce for jobs started by ``mpirun`` (reference: ``deepspeed/runtime/comm/mpi.py``).

The reference moves the packed sign chunks with mpi4py (``Igather`` of cupy / numpy buffers, optionally CUDA-aware).
Here rank and size discovery accept the MPI environment (mpi4py when importable, else the launcher's variables) and the
data movement is :class:`CompressedBackend`'s — NCCL over NVLink on GPUs, gloo on hosts — because on NVSwitch systems a
host-staged MPI gather is strictly slower than the fabric; ``cuda_aware`` is accepted and recorded."""
from __future__ import annotations

import os

from ... import comm as dist
from .compressed import CompressedBackend


class MpiBackend(CompressedBackend):

    def __init__(self, cuda_aware: bool = False, mpu=None):
        self.cuda_aware = bool(cuda_aware)
        self.using_mpi = True
        try:
            from mpi4py import MPI
            self.comm = MPI.COMM_WORLD
            mpi_rank, mpi_size = self.comm.Get_rank(), self.comm.Get_size()
        except ImportError:
            self.comm = None
            mpi_rank = int(os.environ.get("OMPI_COMM_WORLD_RANK", os.environ.get("PMI_RANK", os.environ.get("RANK", 0))))
            mpi_size = int(os.environ.get("OMPI_COMM_WORLD_SIZE", os.environ.get("PMI_SIZE",
                                                                                 os.environ.get("WORLD_SIZE", 1))))
        if not dist.is_initialized():
            os.environ.setdefault("RANK", str(mpi_rank))
            os.environ.setdefault("WORLD_SIZE", str(mpi_size))
            dist.init_distributed(auto_mpi_discovery=self.comm is not None)
        super().__init__(mpu)
        if (self.rank, self.size) != (mpi_rank, mpi_size) and mpu is None:
            raise RuntimeError(f"MPI rank/size {(mpi_rank, mpi_size)} disagrees with torch.distributed "
                               f"{(self.rank, self.size)}")

    # reference-named transport pieces, expressed with the shared gather helpers
    def gather_cuda(self, rank, world_size, comm, cupy_sign_list_packed, cupy_recvbuf_sign, cupy_worker_scale,
                    cupy_recvbuf_scale):
        for root in range(world_size):
            self.my_gather(rank, world_size, self.world_group, cupy_sign_list_packed[root], cupy_recvbuf_sign, root)
            self.my_gather(rank, world_size, self.world_group, cupy_worker_scale, cupy_recvbuf_scale, root)

    gather_host = gather_cuda

    def allgather_cuda(self, comm, cupy_server_sign_packed, cupy_recvbuf_sign_server, cupy_server_scale,
                       cupy_recvbuf_scale_server):
        dist.all_gather_into_tensor(cupy_recvbuf_sign_server.view(-1), cupy_server_sign_packed.view(-1),
                                    group=self.world_group)
        dist.all_gather_into_tensor(cupy_recvbuf_scale_server.view(-1), cupy_server_scale.view(-1),
                                    group=self.world_group)

    allgather_host = allgather_cuda
