from .layer import DistributedAttention  # noqa: F401
