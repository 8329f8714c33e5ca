"""Common base of the per-architecture inference layers (reference ``transformers/ds_base.py``)."""
import torch.nn as nn


class DeepSpeedTransformerBase(nn.Module):

    def __init__(self):
        super().__init__()
