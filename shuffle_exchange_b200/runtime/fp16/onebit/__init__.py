from .adam import OnebitAdam
from .lamb import OnebitLamb
from .zoadam import ZeroOneAdam
