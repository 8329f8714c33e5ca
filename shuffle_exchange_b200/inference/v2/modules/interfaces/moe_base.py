from . import *  # noqa: F401,F403  (reference path modules/interfaces/moe_base.py)
