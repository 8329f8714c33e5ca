from .ds_aio_constants import (BENCH_LOG_DIR, READ_IO_DIR, READ_LOG_DIR, READ_OP_DESC, SCRIPT_PREFIX, WRITE_IO_DIR,
                               WRITE_LOG_DIR, WRITE_OP_DESC)  # noqa: F401
