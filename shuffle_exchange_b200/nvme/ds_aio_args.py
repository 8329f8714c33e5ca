"""Command line of the single-run IO benchmark (same option names as the reference ``nvme/ds_aio_args.py``)."""
from __future__ import annotations

import argparse
import json
import os

from .ds_aio_constants import AIO_BASIC, AIO_HANDLE, TORCH_FAST_IO, TORCH_IO, VALID_ENGINES
from .test_ds_aio_utils import refine_integer_value

MAPPING_DELIMITER = ":"


def _parse_mapping(items):
    """``["/mnt/nvme0:0,1", "/mnt/nvme1:2,3"]`` -> ``{"/mnt/nvme0": [0, 1], ...}``"""
    out = {}
    for it in items or []:
        folder, _, devs = it.partition(MAPPING_DELIMITER)
        out[folder] = [int(d) for d in devs.split(",") if d != ""]
    return out


def refine_args(args):
    if args.io_size is not None:
        args.io_size = refine_integer_value(args.io_size)
    args.block_size = refine_integer_value(args.block_size)
    args.fast_io_size = refine_integer_value(args.fast_io_size)
    args.folder_to_device_mapping = _parse_mapping(args.folder_to_device_mapping)
    return args


def validate_args(args) -> bool:
    if args.folder is None and not args.folder_to_device_mapping:
        print("Missing --folder or --folder_to_device_mapping")
        return False
    if args.folder is not None and args.folder_to_device_mapping:
        print("--folder and --folder_to_device_mapping cannot be specified together.")
        return False
    folders = [args.folder] if args.folder is not None else list(args.folder_to_device_mapping)
    for f in folders:
        if not os.path.isdir(f):
            print(f"Invalid folder: {f}")
            return False
    if args.engine not in VALID_ENGINES:
        print(f"Invalid engine {args.engine}; valid: {VALID_ENGINES}")
        return False
    if args.use_gds and args.engine != AIO_HANDLE:
        print("--use_gds needs --engine aio_handle")
        return False
    if args.read and args.engine in (TORCH_FAST_IO,):
        print(f"--engine {args.engine} supports writes only")
        return False
    return True


def parse_arguments(argv=None):
    p = argparse.ArgumentParser(description="file IO benchmark")
    p.add_argument("--folder", default=None, type=str, help="Folder to use for I/O.")
    p.add_argument("--folder_to_device_mapping", default=None, type=str, nargs="+",
                   help='Mapping of folders to device ids, e.g. "/mnt/nvme0:0,1"')
    p.add_argument("--io_size", type=str, default=None, required=True, help="Number of bytes to read or write.")
    p.add_argument("--fast_io_size", type=str, default="64M", help="Size of fast_io pinned buffer (bytes).")
    p.add_argument("--read", action="store_true", help="Perform read I/O (default is write)")
    p.add_argument("--multi_process", type=int, default=1, help="Number of parallel processes doing I/O.")
    p.add_argument("--block_size", type=str, default="1M", help="I/O block size; K, M or G suffix allowed.")
    p.add_argument("--queue_depth", type=int, default=32, help="I/O queue depth (default 32).")
    p.add_argument("--single_submit", action="store_true", help="Submit I/O requests one at a time.")
    p.add_argument("--sequential_requests", action="store_true",
                   help="Delay I/O request submission until completion of prior requests.")
    p.add_argument("--validate", action="store_true", help="Perform validation of I/O transfer in library.")
    p.add_argument("--engine", type=str, default=AIO_HANDLE, help=f"Engine to perform I/O: {VALID_ENGINES}")
    p.add_argument("--loops", type=int, default=3, help="Count of operation repetitions")
    p.add_argument("--io_parallel", type=int, default=None, help="Per iop parallelism")
    p.add_argument("--gpu", action="store_true", help="Use GPU memory")
    p.add_argument("--use_gds", action="store_true", help="Enable GDS AIO")
    p.add_argument("--slow_bounce_buffer", action="store_true", help="Use non-pinned bounce buffer for GPU I/O.")
    p.add_argument("--torch_legacy_save", action="store_true", help="Use torch legacy save approach")
    p.add_argument("--use_accelerator_pin_memory", action="store_true",
                   help="Obtain pinned (CPU page-locked) tensors from the accelerator")
    p.add_argument("--warmup_loops", type=int, default=1, help="Count of operation warmup repetitions")
    p.add_argument("--include_warmup_time", action="store_true", help="Include warmup latency in results")
    p.add_argument("--different_file_each_iteration", action="store_true",
                   help="Read/write a different file on each iteration.")
    args = p.parse_args(argv)
    print(f"args = {args}")
    return args


def get_validated_args(argv=None):
    args = refine_args(parse_arguments(argv))
    if not validate_args(args):
        raise SystemExit(1)
    print("Successful validation of command line arguments")
    return args


def dump_args(args) -> str:
    return json.dumps({k: v for k, v in vars(args).items()}, default=str)
