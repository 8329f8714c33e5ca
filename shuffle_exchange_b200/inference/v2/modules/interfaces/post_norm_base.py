from . import *  # noqa: F401,F403  (reference path modules/interfaces/post_norm_base.py)
