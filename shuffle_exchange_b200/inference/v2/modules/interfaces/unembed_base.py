from . import *  # noqa: F401,F403  (reference path modules/interfaces/unembed_base.py)
