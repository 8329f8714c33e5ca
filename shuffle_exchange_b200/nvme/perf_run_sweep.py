"""Sweep the IO benchmark over block size / queue depth / submission mode / parallelism and log every point
(reference ``nvme/perf_run_sweep.py``).  ``sweep_main`` runs all read points then all write points; each point is one
``test_ds_aio`` sub-process whose stdout becomes ``<log_dir>/_aio_bench_{read,write}_logs/<tags>.txt``."""
from __future__ import annotations

import argparse
import itertools
import json
import os
import shutil
import sys
from typing import Dict, List

from .ds_aio_job import Job, run_job
from .perf_sweep_utils import (BENCH_LOG_DIR, READ_IO_DIR, READ_LOG_DIR, READ_OP_DESC, WRITE_IO_DIR, WRITE_LOG_DIR,
                               WRITE_OP_DESC)

OTHER_OPTIONS = "--engine aio_handle"
PERF_SCRIPT = "shuffle_exchange_b200.nvme.test_ds_aio"
DEFAULT_SWEEP_CONFIG = {
    "block_size": ["1M", "8M"],
    "queue_depth": [32, 128],
    "sequential_requests": [False],
    "single_submit": [False],
    "io_parallel": [1, 8],
}


class SweepConfig:

    def __init__(self, args):
        self.folder_to_device_mapping = get_ftd_map(args.nvme_dir)
        self.search_space = get_sweep_config_dict(args.sweep_config)
        self.search_space.update(self.folder_to_device_mapping)
        self.read = not args.no_read
        self.write = not args.no_write
        self.flush_cache = args.flush_page_cache
        self.log_dir = args.log_dir
        self.verbose = args.verbose
        self.other_options = f"{OTHER_OPTIONS} --loops {args.loops} --io_size {args.io_size}"
        if args.gpu:
            self.other_options += " --gpu"
        if args.gds:
            self.other_options += " --use_gds"


def validate_arguments(args) -> bool:
    if not args.nvme_dir:
        print("Missing --nvme_dir")
        return False
    for d in args.nvme_dir:
        if not os.path.isdir(d):
            print(f"Invalid folder in --nvme_dir: {d}")
            return False
    if args.no_read and args.no_write:
        print("--no_read and --no_write leave nothing to do")
        return False
    return True


def parse_sweep_arguments(argv=None):
    p = argparse.ArgumentParser()
    p.add_argument("--nvme_dir", nargs="+", required=True,
                   help="Directory in which to perform I/O tests. A writeable directory on a NVMe device.")
    p.add_argument("--sweep_config", type=str, default=None, help="Performance sweep configuration json file.")
    p.add_argument("--no_read", action="store_true", help="Disable read performance measurements.")
    p.add_argument("--no_write", action="store_true", help="Disable write performance measurements.")
    p.add_argument("--io_size", type=str, default="400M", help="Number of I/O bytes to read/write.")
    p.add_argument("--gpu", action="store_true", help="Test tensor transfers between GPU device and NVMe.")
    p.add_argument("--gds", action="store_true", help="Run the sweep over the device-tensor (GDS) path.")
    p.add_argument("--flush_page_cache", action="store_true",
                   help="Page cache will not be flushed and reported read speeds may be inflated unless set (needs sudo).")
    p.add_argument("--log_dir", type=str, default=BENCH_LOG_DIR,
                   help=f"Output directory for performance log files. Default is {os.path.join('.', BENCH_LOG_DIR)}")
    p.add_argument("--loops", type=int, default=1, help="Count of operation repetitions")
    p.add_argument("--verbose", action="store_true", help="Print debugging information.")
    args = p.parse_args(argv)
    print(f"args = {args}")
    return args


def dump_cmd_lines(cmd_lines):
    print(f"cmd line count = {len(cmd_lines)}")
    for i, cmd in enumerate(cmd_lines):
        print(f"{i}:  {cmd}")


def get_ftd_map(nvme_dir_list) -> Dict[str, List[str]]:
    """One process per folder; folder ``i`` is paired with device ``i``."""
    return {"folder_to_device_mapping": [" ".join(f"{d}:{i}" for i, d in enumerate(nvme_dir_list))]}


def get_sweep_config_dict(sweep_config_json) -> Dict[str, list]:
    if sweep_config_json is None:
        return dict(DEFAULT_SWEEP_CONFIG)
    with open(sweep_config_json) as f:
        return json.load(f)


def get_sweep_cmd_lines(sweep_config_dict) -> List[List[str]]:
    def options(key, values):
        out = []
        for v in values:
            if isinstance(v, bool):
                out.append(f"--{key}" if v else "")
            else:
                out.append(f"--{key} {v}")
        return out

    per_key = [options(k, v) for k, v in sweep_config_dict.items()]
    return [[o for o in combo if o] for combo in itertools.product(*per_key)]


def launch_sweep(sweep_jobs, sync_job, flush_cache_job, verbose):
    for job in sweep_jobs:
        if flush_cache_job is not None:
            run_job(sync_job, verbose)
            run_job(flush_cache_job, verbose)
        run_job(job, verbose)
        run_job(sync_job, verbose)


def create_cmd_tags(cmd_line: List[str]) -> Dict[str, str]:
    tags = {}
    for piece in cmd_line:
        for tok in piece.split("--"):
            fields = tok.split()
            if len(fields) == 1:
                tags[fields[0]] = None
            elif len(fields) >= 2:
                tags[fields[0]] = fields[1]
    return tags


def get_log_file(io_op_desc: str, cmd_line: List[str]) -> str:
    tags = create_cmd_tags(cmd_line)
    parts = [io_op_desc,
             "single" if "single_submit" in tags else "block",
             "sequential" if "sequential_requests" in tags else "overlap",
             f"p{tags.get('io_parallel') or 1}", f"d{tags.get('queue_depth') or 32}",
             f"bs{tags.get('block_size') or '1M'}"]
    return "_".join(parts) + ".txt"


def create_perf_jobs(io_op_desc, log_dir, cmd_lines) -> List[Job]:
    py = sys.executable
    return [Job(cmd_line=[py, "-m", PERF_SCRIPT] + cmd, output_file=os.path.join(log_dir, get_log_file(io_op_desc, cmd)))
            for cmd in cmd_lines]


def script_path():
    return os.path.dirname(os.path.realpath(sys.argv[0]))


def async_io_setup():
    from ..ops.aio import AsyncIOBuilder
    return AsyncIOBuilder().is_compatible()


def gds_io_setup():
    from ..ops.gds import GDSBuilder
    return GDSBuilder().is_compatible()


def remove_folder(folder):
    assert os.path.isdir(folder), f"Error: cannot remove {folder} - folder not found"
    shutil.rmtree(folder)


def _run(op_desc, io_dir_name, log_dir_name, extra, sweep_config, flush_cache_job, sync_job, cmd_lines):
    log_folder = os.path.join(sweep_config.log_dir, log_dir_name)
    os.makedirs(log_folder, exist_ok=True)
    cmds = [[extra] + cmd + [sweep_config.other_options] if extra else cmd + [sweep_config.other_options]
            for cmd in cmd_lines]
    launch_sweep(create_perf_jobs(op_desc, log_folder, cmds), sync_job, flush_cache_job, sweep_config.verbose)


def run_read_sweep(sweep_config, flush_cache_job, sync_job, cmd_lines):
    _run(READ_OP_DESC, READ_IO_DIR, READ_LOG_DIR, "--read", sweep_config, flush_cache_job, sync_job, cmd_lines)


def run_write_sweep(sweep_config, flush_cache_job, sync_job, cmd_lines):
    _run(WRITE_OP_DESC, WRITE_IO_DIR, WRITE_LOG_DIR, "", sweep_config, flush_cache_job, sync_job, cmd_lines)


def sweep_main(args):
    cfg = SweepConfig(args)
    cmd_lines = get_sweep_cmd_lines(cfg.search_space)
    if cfg.flush_cache:
        flush = Job(cmd_line=["sudo", "bash", "-c", "echo 1 > /proc/sys/vm/drop_caches"], split_pieces=False)
    else:
        flush = None
    sync_job = Job(cmd_line=["sync"])
    if cfg.verbose:
        dump_cmd_lines(cmd_lines)
    os.makedirs(cfg.log_dir, exist_ok=True)
    if cfg.read:
        run_read_sweep(cfg, flush, sync_job, cmd_lines)
    if cfg.write:
        run_write_sweep(cfg, flush, sync_job, cmd_lines)


def main(argv=None):
    print("Running performance sweep of file IO")
    args = parse_sweep_arguments(argv)
    if not validate_arguments(args):
        raise SystemExit(1)
    if not (gds_io_setup() if args.gds else async_io_setup()):
        raise SystemExit("the native IO engine is not available")
    sweep_main(args)


if __name__ == "__main__":
    main()
