from . import *  # noqa: F401,F403  (reference path modules/configs/norm_config.py)
