"""Rank-aware logging (reference: deepspeed/utils/logging.py — ``logger``, ``log_dist``,
``print_json_dist``, ``warning_once``)."""
from __future__ import annotations

import functools
import json
import logging
import os
import sys

_FMT = "[%(asctime)s] [%(levelname)s] [%(filename)s:%(lineno)d:%(funcName)s] %(message)s"


def _create_logger(name="sxb200", level=logging.INFO):
    lg = logging.getLogger(name)
    if lg.handlers:
        return lg
    lg.setLevel(os.environ.get("SXB_LOG_LEVEL", "").upper() or level)
    lg.propagate = False
    h = logging.StreamHandler(stream=sys.stdout)
    h.setFormatter(logging.Formatter(_FMT))
    lg.addHandler(h)
    return lg


class LoggerFactory:
    """Reference spelling of the logger constructor (``LoggerFactory.create_logger(name, level)``): a named,
    non-propagating stdout logger with this package's line format; asking twice for a name returns the same logger."""

    @staticmethod
    def create_logger(name=None, level=logging.INFO):
        if name is None:
            raise ValueError("name for logger cannot be None")
        return _create_logger(name, level)


log_levels = {"debug": logging.DEBUG, "info": logging.INFO, "warning": logging.WARNING, "error": logging.ERROR,
              "critical": logging.CRITICAL}

logger = _create_logger()


def print_configuration(args, name):
    """Log every attribute of an argparse-style namespace, one aligned ``key ..... value`` line each."""
    logger.info(f"{name}:")
    for key in sorted(vars(args)):
        logger.info(f"  {key} {'.' * max(1, 29 - len(key))} {getattr(args, key)}")


def _my_rank() -> int:
    try:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            return dist.get_rank()
    except Exception:
        pass
    return int(os.environ.get("RANK", "0"))


def _should_log(ranks) -> bool:
    if ranks is None:
        return True
    r = _my_rank()
    return -1 in ranks or r in ranks


def log_dist(message, ranks=None, level=logging.INFO):
    """Log ``message`` only on ``ranks`` (``[-1]`` or ``None`` = every rank)."""
    if _should_log(ranks):
        logger.log(level, f"[Rank {_my_rank()}] {message}")


def print_json_dist(message, ranks=None, path=None):
    if _should_log(ranks):
        message = dict(message)
        message["rank"] = _my_rank()
        if path is None:
            print(json.dumps(message, sort_keys=True))
        else:
            with open(path, "w") as f:
                json.dump(message, f)
                os.fsync(f)


@functools.lru_cache(None)
def warning_once(msg: str):
    logger.warning(msg)


def get_current_level():
    return logger.getEffectiveLevel()


def should_log_le(max_log_level_str: str) -> bool:
    lvl = getattr(logging, max_log_level_str.upper(), None)
    if not isinstance(lvl, int):
        raise ValueError(f"{max_log_level_str} is not a logging level")
    return get_current_level() <= lvl
