"""Names shared by the IO benchmark modules."""
AIO_HANDLE, AIO_BASIC, TORCH_IO, TORCH_FAST_IO = "aio_handle", "aio_basic", "torch_io", "torch_fastio"
VALID_ENGINES = [AIO_HANDLE, AIO_BASIC, TORCH_IO, TORCH_FAST_IO]

BUFFER, BOUNCE_BUFFER, NUM_BYTES, FILE, HANDLE, ELAPSED_SEC = ("buffer", "bounce_buffer", "num_bytes", "file",
                                                               "handle", "elapsed_sec")
FAST_IO_BUFFER, USE_CPU_LOCKED_TENSOR, USE_GDS = "fast_io_buffer", "cpu_locked_tensor", "gds"

SCRIPT_PREFIX = "_aio_bench"
WRITE_OP_DESC, READ_OP_DESC = "write", "read"
READ_IO_DIR = f"{SCRIPT_PREFIX}_{READ_OP_DESC}_io"
WRITE_IO_DIR = f"{SCRIPT_PREFIX}_{WRITE_OP_DESC}_io"
BENCH_LOG_DIR = f"{SCRIPT_PREFIX}_logs"
READ_LOG_DIR = f"{SCRIPT_PREFIX}_{READ_OP_DESC}_logs"
WRITE_LOG_DIR = f"{SCRIPT_PREFIX}_{WRITE_OP_DESC}_logs"
