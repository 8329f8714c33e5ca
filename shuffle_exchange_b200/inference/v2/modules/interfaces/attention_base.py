from . import *  # noqa: F401,F403  (reference path modules/interfaces/attention_base.py)
