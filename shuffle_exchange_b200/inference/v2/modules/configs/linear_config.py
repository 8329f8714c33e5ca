from . import *  # noqa: F401,F403  (reference path modules/configs/linear_config.py)
