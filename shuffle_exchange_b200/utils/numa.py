"""NUMA helpers for the launcher (reference: deepspeed/utils/numa.py): build a ``numactl`` prefix that
binds each local rank to its share of cores."""
import os
import shutil
import subprocess


def get_numa_cores():
    try:
        out = subprocess.check_output(["numactl", "--hardware"], text=True)
    except Exception:
        return []
    nodes = []
    for line in out.splitlines():
        if line.startswith("node") and "cpus:" in line:
            nodes.append([int(c) for c in line.split("cpus:")[1].split()])
    return nodes


def parse_range_list(s):
    cores = []
    for part in s.split(","):
        if "-" in part:
            a, b = part.split("-")
            cores.extend(range(int(a), int(b) + 1))
        elif part:
            cores.append(int(part))
    if sorted(set(cores)) != cores:
        raise ValueError(f"core list must be ascending without repeats: {s}")
    return cores


def get_numactl_cmd(bind_core_list, num_local_procs, local_rank):
    if shutil.which("numactl") is None:
        raise RuntimeError("--bind_cores_to_rank requires numactl")
    if bind_core_list:
        cores = parse_range_list(bind_core_list)
    else:
        cores = list(range(os.cpu_count() or 1))
    per = max(1, len(cores) // num_local_procs)
    mine = cores[local_rank * per:(local_rank + 1) * per]
    return per, ["numactl", "-C", ",".join(map(str, mine))]
