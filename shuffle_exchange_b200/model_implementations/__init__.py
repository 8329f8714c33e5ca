"""Kernel-backed model wrappers (reference ``deepspeed/model_implementations``)."""
from .transformers.ds_transformer import DeepSpeedTransformerInference
from .transformers.clip_encoder import DSClipEncoder
from .diffusers.unet import DSUNet
from .diffusers.vae import DSVAE
