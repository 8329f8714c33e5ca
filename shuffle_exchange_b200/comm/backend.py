"""Communication back-end object model (reference: ``deepspeed/comm/backend.py``).

The module-level functions of ``comm.comm`` are what the framework calls; a :class:`Backend` object is the reference's
seam for alternative transports.  On this platform there are two: :class:`~.torch.TorchBackend` (NCCL over NVLink /
NVSwitch, gloo on CPU) for plumbing, and the symmetric-memory kernels of ``comm.symm`` for the fused hot paths."""


class Backend:

    def __init__(self, name: str = "backend", rank: int = 0, size: int = 1):
        self.name = name
        # the world group is created lazily by init_process_group
        self.world_group = None
        self.world_size = size
        self.world_rank = rank
        self.process_groups = []
        self.initialized = False

    def is_initialized(self) -> bool:
        return self.initialized

    def new_group(self):
        raise NotImplementedError

    def init_process_group(self):
        # subclasses initialise their transport here and set ``initialized``
        self.initialized = True
