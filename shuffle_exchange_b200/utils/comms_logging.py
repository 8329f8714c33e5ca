"""Per-collective latency / bandwidth accounting.

Output schema follows the reference (deepspeed/utils/comms_logging.py:34-64 ``calc_bw_log``,
:126-181 ``log_all``): algbw = bytes/time, busbw = algbw x correction (all-gather / reduce-scatter /
all-to-all x (n-1)/n, all-reduce x 2(n-1)/n, point-to-point x 1).
"""
from __future__ import annotations

import math
from collections import defaultdict

from .logging import log_dist


def get_caller_func(frame=3):
    import sys
    return sys._getframe(frame).f_code.co_name


def convert_size(size_bytes: int) -> str:
    if size_bytes == 0:
        return "0B"
    names = ("B", "KB", "MB", "GB", "TB", "PB")
    i = int(math.floor(math.log(size_bytes, 1024)))
    return f"{round(size_bytes / math.pow(1024, i), 2)} {names[i]}"


def calc_bw_log(comm_op: str, size: int, duration_ms: float, n: int):
    """Returns (algbw_Gbps, busbw_Gbps, total_bytes)."""
    duration = max(duration_ms / 1e3, 1e-12)
    if comm_op in ("all_to_all_single", "all_to_all"):
        tput, busbw = size / duration, (size / duration) * ((n - 1) / max(n, 1))
    elif comm_op in ("all_gather", "all_gather_into_tensor", "reduce_scatter", "reduce_scatter_tensor",
                     "all_gather_base", "reduce_scatter_base", "all_gather_coalesced"):
        size *= n
        tput, busbw = size / duration, (size / duration) * ((n - 1) / max(n, 1))
    elif comm_op in ("all_reduce", "all_reduce_coalesced", "inference_all_reduce"):
        tput, busbw = size * 2 / duration, (size / duration) * (2 * (n - 1) / max(n, 1))
    else:  # send / recv / broadcast / reduce / gather / scatter / barrier
        tput = busbw = size / duration
    return tput * 8 / 1e9, busbw * 8 / 1e9, size


class CommsLogger:
    def __init__(self):
        self.comms_dict = defaultdict(dict)
        self.verbose = False
        self.debug = False
        self.prof_ops = []
        self.prof_all = True
        self.enabled = False

    def configure(self, cfg):
        self.enabled = cfg.enabled
        if self.enabled:
            self.verbose, self.debug = cfg.verbose, cfg.debug
            self.prof_ops, self.prof_all = list(cfg.prof_ops), cfg.prof_all

    def start_profiling_comms(self):
        self.prof_all = True

    def stop_profiling_comms(self):
        self.prof_all = False

    def start_profiling_op(self, names):
        self.prof_ops = list(set(self.prof_ops) | set(names))

    def stop_profiling_op(self, names):
        self.prof_ops = [o for o in self.prof_ops if o not in names]

    def append(self, raw_name, record_name, latency_ms, msg_size, world):
        algbw, busbw, size = calc_bw_log(raw_name, msg_size, latency_ms, world)
        rec = self.comms_dict[record_name]
        if size in rec:
            rec[size][0] += 1
            rec[size][1].append(latency_ms)
            rec[size][2].append(algbw)
            rec[size][3].append(busbw)
        else:
            rec[size] = [1, [latency_ms], [algbw], [busbw]]
        if self.verbose:
            log_dist(f"comm op: {record_name} | time (ms): {latency_ms:.2f} | msg size: {convert_size(size)} | "
                     f"algbw (Gbps): {algbw:.2f} | busbw (Gbps): {busbw:.2f}", ranks=[0])

    def log_all(self, print_log=True, show_straggler=False):
        from ..comm import comm as dist
        rows = []
        if print_log:
            print(f"{'Comm. Op': <20}{'Message Size': <20}{'Count': <20}{'Total Latency(ms)': <20}"
                  f"{'Avg Latency(ms)': <20}{'tput_avg (Gbps)': <20}{'busbw_avg (Gbps)': <20}")
        for name, sizes in self.comms_dict.items():
            if print_log:
                print(name)
            for size, (count, lats, algs, buss) in sorted(sizes.items()):
                total = sum(lats)
                row = (name, size, count, total, total / count, sum(algs) / count, sum(buss) / count)
                rows.append(row)
                if print_log:
                    print(f"{' ': <20}{convert_size(size): <20}{count: <20}{total: <20.2f}{row[4]: <20.2f}"
                          f"{row[5]: <20.2f}{row[6]: <20.2f}")
        if show_straggler and dist.is_initialized():
            import torch
            if print_log:
                print("_______________________________\nBreakdown with straggler effect\n"
                      "-------------------------------")
            for name, sizes in self.comms_dict.items():
                for size, (count, lats, _, _) in sorted(sizes.items()):
                    t = torch.tensor(lats, dtype=torch.float64)
                    mn = t.clone()
                    if dist.get_backend() == "nccl":
                        mn = mn.cuda()
                    dist.all_reduce(mn, op=dist.ReduceOp.MIN)
                    strag = float((t - mn.cpu()).sum())
                    if print_log:
                        print(f"{name: <20}{convert_size(size): <20}{count: <20}{sum(lats): <20.2f}{strag: <20.2f}")
        return rows
