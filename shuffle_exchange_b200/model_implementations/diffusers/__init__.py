from .unet import DSUNet
from .vae import DSVAE
