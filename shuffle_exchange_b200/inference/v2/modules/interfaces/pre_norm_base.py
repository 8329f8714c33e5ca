from . import *  # noqa: F401,F403  (reference path modules/interfaces/pre_norm_base.py)
