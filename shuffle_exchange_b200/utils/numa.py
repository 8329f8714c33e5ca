"""NUMA helpers for the launcher (reference: ``deepspeed/utils/numa.py:17-203``): build the ``numactl`` prefix that
binds each local rank to its share of cores and, when all of those cores sit in one NUMA domain, to that domain's
memory.  Three host layouts are told apart from ``numactl --hardware``:

* ordinary: every node has cores -> ``-m <node>``;
* flat HBM: as many core-less nodes (HBM) as nodes with cores -> prefer the HBM node paired with the core node
  (``-p``), falling back to DDR when it is full;
* "fake NUMA": several consecutive nodes list the same cores -> bind memory to all of them.
"""
import os
import shutil
import subprocess
from typing import List, Sequence, Tuple


def get_numa_cores() -> List[List[int]]:
    try:
        out = subprocess.check_output(["numactl", "--hardware"], text=True)
    except Exception:
        return []
    nodes = []
    for line in out.splitlines():
        if line.startswith("node") and "cpus:" in line:
            nodes.append([int(c) for c in line.split("cpus:")[1].split()])
    return nodes


def check_for_numactl_pkg() -> bool:
    """Whether the numactl package is installed according to the host's package manager (dpkg / pacman / rpm); prints
    the install hint when the manager says it is not."""
    for manager, query, installer in (("dpkg", "-l", "apt"), ("pacman", "-Q", "pacman"), ("rpm", "-q", "yum")):
        if shutil.which(manager) is None:
            continue
        ok = subprocess.run([manager, query, "numactl"], stdout=subprocess.DEVNULL,
                            stderr=subprocess.DEVNULL).returncode == 0
        if not ok:
            print(f"please install the numactl package with {installer}")
        return ok
    return False


def parse_range(rng: str) -> range:
    """``"5"`` -> ``range(5, 6)``; ``"2-4"`` -> ``range(2, 5)`` (both ends included)."""
    lo, dash, hi = rng.partition("-")
    try:
        start = int(lo)
        end = int(hi) if dash else start
    except ValueError:
        raise ValueError(f"Bad range: '{rng}', range must be either a number or two number separated by dash")
    if "-" in hi:
        raise ValueError(f"Bad range: '{rng}', range must be either a number or two number separated by dash")
    if start > end:
        raise ValueError(f"Bad range: '{rng}', range end must larger than or equal to start")
    return range(start, end + 1)


def parse_range_list(range_str: str) -> List[int]:
    """``"0,2-4,6"`` -> ``[0, 2, 3, 4, 6]``; sub-ranges ascending, no overlap, no spaces."""
    cores: List[int] = []
    for part in range_str.split(","):
        r = parse_range(part)
        if cores and r[0] <= cores[-1]:
            raise ValueError(f"Bad range: '{range_str}', sub ranges must not overlap with each other and should be "
                             "in ascend order")
        cores.extend(r)
    return cores


def _format_cores(cores: Sequence[int]) -> str:
    """Shortest ``-C`` argument: runs of consecutive ids become ``a-b``."""
    runs: List[Tuple[int, int]] = []
    for c in cores:
        if runs and c == runs[-1][1] + 1:
            runs[-1] = (runs[-1][0], c)
        else:
            runs.append((c, c))
    return ",".join(str(a) if a == b else f"{a}-{b}" for a, b in runs)


def _memory_policy(cores: Sequence[int], numa_cores: List[List[int]]) -> List[str]:
    """``numactl`` memory flags for a rank running on ``cores`` (empty when the cores span several domains)."""
    home = next((i for i, node in enumerate(numa_cores) if node and set(cores) <= set(node)), None)
    if home is None:
        return []
    with_cores = [i for i, node in enumerate(numa_cores) if node]
    without = [i for i, node in enumerate(numa_cores) if not node]
    if without and len(without) == len(with_cores):                 # flat HBM: k-th core node <-> k-th HBM node
        return ["-p", str(without[with_cores.index(home)])]
    twins = [i for i, node in enumerate(numa_cores) if node == numa_cores[home]]
    if len(twins) > 1 and twins == list(range(twins[0], twins[-1] + 1)):   # fake NUMA: same cores listed repeatedly
        return ["-m", ",".join(map(str, twins))]
    return ["-m", str(home)]


def get_numactl_cmd(bind_core_list, num_local_procs, local_rank):
    """-> ``(cores_per_rank, ["numactl", <memory flags>, "-C", "<cores>"])`` for this local rank."""
    if shutil.which("numactl") is None:
        check_for_numactl_pkg()
        raise RuntimeError("--bind_cores_to_rank requires numactl")
    if "KMP_AFFINITY" in os.environ:
        raise ValueError("Environment variable KMP_AFFINITY conflicts with numactl because it interferes with how many "
                         "CPU cores numactl can set. Unset KMP_AFFINITY before launching.")
    if bind_core_list:
        cores = parse_range_list(bind_core_list)
    else:
        try:
            import psutil
            n = psutil.cpu_count(logical=False) or os.cpu_count() or 1
        except ImportError:
            n = os.cpu_count() or 1
        cores = list(range(n))
    per = len(cores) // num_local_procs
    assert per >= 1, "At least one core needs to be assigned to each rank"
    mine = cores[local_rank * per:(local_rank + 1) * per]
    return per, ["numactl", *_memory_policy(mine, get_numa_cores()), "-C", _format_cores(mine)]
