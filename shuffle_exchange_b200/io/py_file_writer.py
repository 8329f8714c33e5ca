"""Ordinary buffered file as a checkpoint writer back end, instrumented like the others (reference:
``deepspeed/io/py_file_writer.py``) — the baseline the fast (aio, double-buffered) writer is compared with."""
import time

from . import constants as K
from .base_file_writer import BaseFileWriter


class PyFileWriter(BaseFileWriter):

    def __init__(self, file_path):
        super().__init__(file_path)
        self._fp = open(file_path, "wb")

    def write(self, buffer):
        began = time.time()
        written = self._fp.write(buffer)
        for key, amount in ((K.WRITE_SEC_KEY, time.time() - began), (K.WRITE_COUNT_KEY, 1),
                            (K.WRITE_BYTES_KEY, len(buffer))):
            self._incr_stats(key, amount)
        return written

    def flush(self):
        self._fp.flush()
        self._incr_stats(K.FLUSH_COUNT_KEY)

    def fileno(self):
        return self._fp.fileno()

    def close(self):
        self._fp.close()
        self._incr_stats(K.CLOSE_COUNT_KEY)
