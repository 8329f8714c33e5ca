"""Logger of the v2 inference stack (reference ``inference/v2/logging.py``)."""
import logging
import sys

inf_logger = None


def inference_logger(level: int = logging.INFO) -> logging.Logger:
    global inf_logger
    if inf_logger is None:
        inf_logger = logging.getLogger("DS-Inference")
        inf_logger.setLevel(level)
        if not inf_logger.handlers:
            h = logging.StreamHandler(stream=sys.stdout)
            h.setFormatter(logging.Formatter("[%(asctime)s] [%(levelname)s] [%(filename)s:%(lineno)d] %(message)s"))
            inf_logger.addHandler(h)
        inf_logger.propagate = False
        inf_logger.debug("Inference logger created.")
    return inf_logger
