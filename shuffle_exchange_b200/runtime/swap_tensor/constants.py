"""``aio`` config block keys (reference: ``runtime/swap_tensor/constants.py``)."""
AIO = "aio"
AIO_BLOCK_SIZE, AIO_BLOCK_SIZE_DEFAULT = "block_size", 1048576
AIO_QUEUE_DEPTH, AIO_QUEUE_DEPTH_DEFAULT = "queue_depth", 8
AIO_INTRA_OP_PARALLELISM, AIO_INTRA_OP_PARALLELISM_DEFAULT = "intra_op_parallelism", 1
AIO_THREAD_COUNT, AIO_THREAD_COUNT_DEFAULT = "thread_count", 1
AIO_SINGLE_SUBMIT, AIO_SINGLE_SUBMIT_DEFAULT = "single_submit", False
AIO_OVERLAP_EVENTS, AIO_OVERLAP_EVENTS_DEFAULT = "overlap_events", True
AIO_USE_GDS, AIO_USE_GDS_DEFAULT = "use_gds", False
