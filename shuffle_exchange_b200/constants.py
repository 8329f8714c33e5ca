"""Package-level constants (reference: ``deepspeed/constants.py``)."""
import os
from datetime import timedelta

# torch.distributed rendezvous defaults
TORCH_DISTRIBUTED_DEFAULT_PORT = 29500

# Process-group timeout: large models can spend a long time in the first collective (checkpoint load, JIT warm-up), so
# the default is 30 minutes; override with DEEPSPEED_TIMEOUT / SXB_TIMEOUT (minutes).
default_pg_timeout = timedelta(minutes=int(os.getenv("DEEPSPEED_TIMEOUT", os.getenv("SXB_TIMEOUT", default=30))))

INFERENCE_GENERIC_MODE = "generic"
INFERENCE_SPECIALIZED_MODE = "specialized"

CROSS_RANK = "CROSS_RANK"
CROSS_SIZE = "CROSS_SIZE"
LOCAL_RANK = "LOCAL_RANK"
