"""File-like writer interface that ``torch.save`` can target (reference: ``deepspeed/io/base_file_writer.py``)."""
from .constants import *  # noqa: F401,F403

BASE_STAT_KEYS = [CLOSE_COUNT_KEY, FILE_PATH_KEY, FLUSH_COUNT_KEY, SAVE_STORAGE_BYTES_KEY, SAVE_STORAGE_KEY,
                  SAVE_STORAGE_SEC_KEY, WRITE_COUNT_KEY, WRITE_BYTES_KEY, WRITE_SEC_KEY]


class BaseFileWriter:
    def __init__(self, file_path):
        self._file_path = file_path
        self._stats = {k: 0 for k in BASE_STAT_KEYS}
        self._stats[FILE_PATH_KEY] = file_path

    def close(self):
        raise NotImplementedError

    def fileno(self):
        raise NotImplementedError

    def flush(self):
        raise NotImplementedError

    def write(self, buffer):
        raise NotImplementedError

    def file_path(self):
        return self._file_path

    def _incr_stats(self, key, incr=1):
        self._stats[key] = self._stats.get(key, 0) + incr

    def _dump_state(self):
        if self._stats[WRITE_SEC_KEY] > 0:
            self._stats["write_GB/s"] = self._stats[WRITE_BYTES_KEY] / self._stats[WRITE_SEC_KEY] / 1024 ** 3
        print(f"stats = {self._stats}")
