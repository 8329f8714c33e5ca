"""``NcclBackend``: the 1-bit all-reduce over NCCL (reference: ``deepspeed/runtime/comm/nccl.py:16-166``).  The
algorithm lives in :class:`CompressedBackend`; this subclass only pins the process group to the NCCL world / the
mpu's data-parallel group, which is what the reference class does besides using cupy for bit packing (here:
``sxb_packbits`` / ``sxb_unpackbits``)."""
from .compressed import CompressedBackend


class NcclBackend(CompressedBackend):
    def __init__(self, mpu=None):
        super().__init__(mpu)
        self.using_mpi = False
