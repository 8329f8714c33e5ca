"""Writer that discards everything — measures serialisation cost without IO (reference: ``deepspeed/io/mock_file_writer.py``)."""
from .base_file_writer import BaseFileWriter
from .constants import CLOSE_COUNT_KEY, FLUSH_COUNT_KEY, WRITE_BYTES_KEY, WRITE_COUNT_KEY


class MockFileWriter(BaseFileWriter):
    def close(self):
        self._incr_stats(CLOSE_COUNT_KEY)

    def fileno(self):
        return -1

    def flush(self):
        self._incr_stats(FLUSH_COUNT_KEY)

    def write(self, buffer):
        self._incr_stats(WRITE_COUNT_KEY)
        self._incr_stats(WRITE_BYTES_KEY, len(buffer))
        return len(buffer)
