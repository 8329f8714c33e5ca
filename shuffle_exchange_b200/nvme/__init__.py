"""NVMe / file-IO benchmarking and tuning tools (reference ``deepspeed/nvme``): single-run benchmark
(``test_ds_aio``), parameter sweep (``perf_run_sweep``), log parsing (``parse_nvme_stats``) and generation of the
``aio`` config block from sweep results (``perf_generate_param``)."""
from .perf_run_sweep import sweep_main, parse_sweep_arguments
from .perf_generate_param import generate_main
from .test_ds_aio import ds_io_main
