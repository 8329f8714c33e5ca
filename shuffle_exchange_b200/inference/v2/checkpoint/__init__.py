"""Checkpoint sources for the v2 engine (reference ``inference/v2/checkpoint/*``)."""
from .base_engine import CheckpointEngineBase
from .in_memory_engine import InMemoryModelEngine
from .huggingface_engine import HuggingFaceCheckpointEngine
