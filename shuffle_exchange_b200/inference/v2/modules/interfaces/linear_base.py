from . import *  # noqa: F401,F403  (reference path modules/interfaces/linear_base.py)
