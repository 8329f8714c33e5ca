"""Extract latency / speed figures from benchmark logs (reference ``nvme/parse_nvme_stats.py``)."""
from __future__ import annotations

import argparse
import os
import re
from typing import Dict, List, Optional, Tuple

READ_SPEED, WRITE_SPEED = "read_speed", "write_speed"
PERF_METRICS = [READ_SPEED, WRITE_SPEED]
METRIC_SEARCH = {READ_SPEED: "E2E Read Speed", WRITE_SPEED: "E2E Write Speed"}
_NUM = re.compile(r"=\s*([0-9.]+)")


def parse_arguments(argv=None):
    p = argparse.ArgumentParser()
    p.add_argument("--log_dir", type=str, required=True, help="Folder of statistics logs")
    p.add_argument("--metric", type=str, required=True, help=f"Performance metric to report: {PERF_METRICS}")
    args = p.parse_args(argv)
    print(f"args = {args}")
    return args


def extract_value(key: str, file: str) -> Optional[float]:
    """Value after ``=`` on the first line containing ``key`` (``... Speed = 3.21 GB/sec``)."""
    with open(file) as f:
        for line in f:
            if key in line:
                m = _NUM.search(line)
                if m:
                    return float(m.group(1))
    return None


def get_file_key(file: str) -> Tuple[str, ...]:
    """``..._single_overlap_8_p1_d32_bs1M.txt`` -> the tuple of config tags (used as dict key / sort key)."""
    base = os.path.splitext(os.path.basename(file))[0]
    return tuple(base.split("_"))


def get_thread_count(file: str) -> int:
    for tag in get_file_key(file):
        if re.fullmatch(r"t\d+", tag):
            return int(tag[1:])
    return 1


def get_metric(file: str, metric: str) -> Optional[float]:
    return extract_value(METRIC_SEARCH[metric], file)


def validate_args(args) -> bool:
    if args.metric not in PERF_METRICS:
        print(f"{args.metric} is not a valid performance metrics")
        return False
    if not os.path.isdir(args.log_dir):
        print(f"{args.log_dir} folder is not existent")
        return False
    return True


def get_results(log_files: List[str], metric: str) -> Dict[Tuple[str, ...], float]:
    out = {}
    for f in log_files:
        v = get_metric(f, metric)
        if v is not None:
            out[get_file_key(f)] = v
    return out


def get_sorted_results(log_dir: str, metric: str):
    files = [os.path.join(log_dir, f) for f in os.listdir(log_dir) if os.path.isfile(os.path.join(log_dir, f))]
    res = get_results(files, metric)
    return sorted(res.items(), key=lambda kv: kv[0])


def main(argv=None):
    print("Parsing aio statistics")
    args = parse_arguments(argv)
    if not validate_args(args):
        raise SystemExit(1)
    for key, val in get_sorted_results(args.log_dir, args.metric):
        print(f"{key} = {val}")


if __name__ == "__main__":
    main()
