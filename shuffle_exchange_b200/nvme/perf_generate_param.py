"""Pick the best sweep point and emit the ``aio`` config block for it (reference ``nvme/perf_generate_param.py``)."""
from __future__ import annotations

import argparse
import json
import os

from .parse_nvme_stats import READ_SPEED, WRITE_SPEED, get_sorted_results
from .perf_sweep_utils import BENCH_LOG_DIR, READ_LOG_DIR, WRITE_LOG_DIR


def parse_arguments(argv=None):
    p = argparse.ArgumentParser()
    p.add_argument("--log_dir", type=str, default=BENCH_LOG_DIR,
                   help=f"Folder of performance sweep logs. Default is {os.path.join('.', BENCH_LOG_DIR)}")
    p.add_argument("--verbose", action="store_true", help="Print all sweep results")
    args = p.parse_args(argv)
    print(f"args = {args}")
    return args


def validate_args(args) -> bool:
    for d in (os.path.join(args.log_dir, READ_LOG_DIR), os.path.join(args.log_dir, WRITE_LOG_DIR)):
        if not os.path.isdir(d):
            print(f"{d} folder is not existent")
            return False
    return True


def convert_to_param(key) -> dict:
    """Config tags of a log-file key -> aio parameters.  Tags: ``single|block``, ``overlap|sequential``,
    ``p<io_parallel>``, ``d<queue_depth>``, ``bs<block_size>``."""
    out = {"single_submit": "false", "overlap_events": "true", "thread_count": 1, "queue_depth": 32,
           "block_size": 1 << 20}
    mult = {"K": 1 << 10, "M": 1 << 20, "G": 1 << 30}
    for tag in key:
        if tag == "single":
            out["single_submit"] = "true"
        elif tag == "block":
            out["single_submit"] = "false"
        elif tag == "sequential":
            out["overlap_events"] = "false"
        elif tag == "overlap":
            out["overlap_events"] = "true"
        elif tag.startswith("bs") and tag[2:3].isdigit():
            v = tag[2:]
            out["block_size"] = int(v[:-1]) * mult[v[-1].upper()] if v[-1].upper() in mult else int(v)
        elif tag.startswith("d") and tag[1:].isdigit():
            out["queue_depth"] = int(tag[1:])
        elif tag.startswith("p") and tag[1:].isdigit():
            out["thread_count"] = int(tag[1:])
    return out


def generate_aio_param(read_log_dir, write_log_dir):
    """Best combined config: the key maximising read + write speed among configs present in both sweeps."""
    reads = dict(get_sorted_results(read_log_dir, READ_SPEED))
    writes = dict(get_sorted_results(write_log_dir, WRITE_SPEED))
    strip = lambda k: tuple(t for t in k if t not in ("read", "write"))   # noqa: E731
    r = {strip(k): v for k, v in reads.items()}
    w = {strip(k): v for k, v in writes.items()}
    common = set(r) & set(w)
    if not common:
        raise RuntimeError("no sweep configuration has both a read and a write result")
    best = max(common, key=lambda k: r[k] + w[k])
    param = convert_to_param(best)
    print(f"Best performance (GB/sec): read = {r[best]:5.2f}, write = {w[best]:5.2f}")
    block = {"aio": param}
    print(json.dumps(block, indent=3))
    return block


def generate_main(log_dir):
    return generate_aio_param(os.path.join(log_dir, READ_LOG_DIR), os.path.join(log_dir, WRITE_LOG_DIR))


def main(argv=None):
    print("Generate aio param")
    args = parse_arguments(argv)
    if not validate_args(args):
        raise SystemExit(1)
    generate_main(args.log_dir)


if __name__ == "__main__":
    main()
