"""Plain buffered-file writer with statistics (reference: ``deepspeed/io/py_file_writer.py``)."""
import time

from .base_file_writer import BaseFileWriter
from .constants import CLOSE_COUNT_KEY, FLUSH_COUNT_KEY, WRITE_BYTES_KEY, WRITE_COUNT_KEY, WRITE_SEC_KEY


class PyFileWriter(BaseFileWriter):
    def __init__(self, file_path):
        super().__init__(file_path)
        self._fp = open(file_path, "wb")

    def close(self):
        self._incr_stats(CLOSE_COUNT_KEY)
        self._fp.close()

    def fileno(self):
        return self._fp.fileno()

    def flush(self):
        self._incr_stats(FLUSH_COUNT_KEY)
        self._fp.flush()

    def write(self, buffer):
        t0 = time.time()
        n = self._fp.write(buffer)
        self._incr_stats(WRITE_SEC_KEY, time.time() - t0)
        self._incr_stats(WRITE_COUNT_KEY)
        self._incr_stats(WRITE_BYTES_KEY, len(buffer))
        return n
