"""``sxb200`` / ``deepspeed``-style job launcher front end.

Reference: deepspeed/launcher/runner.py — hostfile parsing :218 (``hostname slots=N``), ``--include`` /
``--exclude`` resource filters ``host1:0,2@host2`` :298, single-node fast path that execs
``launcher.launch`` locally, multi-node dispatch through PDSH/OpenMPI/MPICH/IMPI/Slurm/MVAPICH.
"""
from __future__ import annotations

import argparse
import base64
import collections
import json
import os
import re
import shutil
import signal
import subprocess
import sys
from copy import deepcopy

import torch

from ..utils.logging import logger
from .constants import (IMPI_LAUNCHER, MPICH_LAUNCHER, MVAPICH_LAUNCHER, OPENMPI_LAUNCHER, PDSH_LAUNCHER, SLURM_LAUNCHER,
                        TORCH_DISTRIBUTED_DEFAULT_PORT)
from .multinode_runner import IMPIRunner, MPICHRunner, MVAPICHRunner, OpenMPIRunner, PDSHRunner, SlurmRunner

DLTS_HOSTFILE = "/job/hostfile"
EXPORT_ENVS = ["NCCL", "PYTHON", "CUDA", "SXB", "LD_LIBRARY_PATH", "PATH"]
DEEPSPEED_ENVIRONMENT_NAME = os.getenv("DS_ENV_FILE", ".deepspeed_env")
DEEPSPEED_ENVIRONMENT_PATHS = [os.path.expanduser("~"), "."]


def parse_args(args=None):
    p = argparse.ArgumentParser(description="sxb200 runner: launch distributed multi-node / multi-GPU training jobs")
    p.add_argument("-H", "--hostfile", type=str, default=DLTS_HOSTFILE)
    p.add_argument("-i", "--include", type=str, default="")
    p.add_argument("-e", "--exclude", type=str, default="")
    p.add_argument("--num_nodes", type=int, default=-1)
    p.add_argument("--min_elastic_nodes", type=int, default=-1)
    p.add_argument("--max_elastic_nodes", type=int, default=-1)
    p.add_argument("--num_gpus", "--num_accelerators", type=int, default=-1)
    p.add_argument("--master_port", default=TORCH_DISTRIBUTED_DEFAULT_PORT, type=int)
    p.add_argument("--master_addr", default="", type=str)
    p.add_argument("--launcher", default=PDSH_LAUNCHER, type=str)
    p.add_argument("--launcher_args", default="", type=str)
    p.add_argument("--module", action="store_true")
    p.add_argument("--no_python", action="store_true")
    p.add_argument("--no_local_rank", action="store_true")
    p.add_argument("--no_ssh", action="store_true")
    p.add_argument("--no_ssh_check", action="store_true")
    p.add_argument("--node_rank", type=int, default=-1)
    p.add_argument("--force_multi", action="store_true")
    p.add_argument("--save_pid", action="store_true")
    p.add_argument("--enable_each_rank_log", default="None", type=str)
    p.add_argument("--autotuning", default="", choices=["tune", "run", ""], type=str)
    p.add_argument("--elastic_training", action="store_true")
    p.add_argument("--bind_cores_to_rank", action="store_true")
    p.add_argument("--bind_core_list", type=str, default=None)
    p.add_argument("--ssh_port", type=int, default=None)
    p.add_argument("--venv_script", type=str, default=None)
    p.add_argument("user_script", type=str)
    p.add_argument("user_args", nargs=argparse.REMAINDER)
    return p.parse_args(args=args)


def fetch_hostfile(hostfile_path):
    if not os.path.isfile(hostfile_path):
        logger.warning("Unable to find hostfile, will proceed with training with local resources only.")
        return None
    with open(hostfile_path, "r") as fd:
        return _parse_hostfile(fd.readlines())


def _parse_hostfile(lines):
    pattern = re.compile(r"^(\S+)\s+slots=(\d+)")
    pool = collections.OrderedDict()
    for line in lines:
        line = line.strip()
        if not line or line.startswith("#"):
            continue
        m = pattern.search(line)
        if not m:
            raise ValueError(f"Hostfile contains a bad entry: {line}, unable to proceed with training.")
        host, slots = m.group(1), int(m.group(2))
        if host in pool:
            raise ValueError(f"Hostfile contains multiple entries for {host}, unable to proceed with training.")
        pool[host] = slots
    if not pool:
        raise ValueError("Hostfile is empty or not formatted correctly, unable to proceed with training.")
    return pool


def _stable_unique(xs):
    seen = set()
    return [x for x in xs if not (x in seen or seen.add(x))]


def parse_resource_filter(host_info, include_str="", exclude_str=""):
    """``host_info``: {host: [slot ids]}.  Filters look like ``worker-0:0,2@worker-1`` (``@`` separates
    nodes, ``:`` introduces a slot list).  include and exclude are mutually exclusive."""
    if include_str and exclude_str:
        raise ValueError("include_str and exclude_str are mutually exclusive.")
    if not include_str and not exclude_str:
        return host_info
    filtered = collections.OrderedDict()
    if include_str:
        spec = include_str
    else:
        filtered = deepcopy(host_info)
        spec = exclude_str
    for node in spec.split("@"):
        if ":" in node:
            host, slots = node.split(":")
            slots = [int(x) for x in slots.split(",")]
            if host not in host_info:
                raise ValueError(f"Hostname '{host}' not found in hostfile")
            for s in slots:
                if s not in host_info[host]:
                    raise ValueError(f"No slot '{s}' specified on host '{host}'")
            if include_str:
                filtered[host] = filtered.get(host, []) + slots
            else:
                for s in slots:
                    filtered[host].remove(s)
        else:
            host = node
            if host not in host_info:
                raise ValueError(f"Hostname '{host}' not found in hostfile")
            if include_str:
                filtered[host] = host_info[host]
            else:
                filtered[host] = []
    out = collections.OrderedDict()
    for host in host_info:
        if host in filtered and filtered[host]:
            out[host] = sorted(_stable_unique(filtered[host]))
    return out


def parse_inclusion_exclusion(resource_pool, inclusion, exclusion):
    active = collections.OrderedDict((h, list(range(s))) for h, s in resource_pool.items())
    return parse_resource_filter(active, include_str=inclusion, exclude_str=exclusion)


def encode_world_info(world_info):
    return base64.urlsafe_b64encode(json.dumps(world_info).encode("utf-8")).decode("utf-8")


def parse_num_nodes(str_num_nodes: str, elastic_training: bool):
    parts = str_num_nodes.split(":")
    if len(parts) == 1:
        return int(parts[0]), -1
    if len(parts) == 2 and elastic_training:
        mn, mx = int(parts[0]), int(parts[1])
        if mn > mx:
            raise RuntimeError("MIN:MAX format is required with MIN <= MAX")
        return mn, mx
    raise RuntimeError("num_nodes {} is not in MIN:MAX format".format(str_num_nodes))


def parse_node_config(node_config: str):
    """``"host"`` -> ``("host", [])`` (all slots); ``"host:0,2,5"`` -> ``("host", [0, 2, 5])``."""
    SLOT_LIST_START, SLOT_SEP = ":", ","
    if SLOT_LIST_START not in node_config:
        return node_config, []
    hostname, slots = node_config.split(SLOT_LIST_START)
    return hostname, [int(x) for x in slots.split(SLOT_SEP)]


def parse_node_config_list(node_config_list):
    """``["a:0,1", "b"]`` (the ``@``-separated pieces of an include / exclude string) -> ``{host: [slots]}``."""
    out = collections.OrderedDict()
    for node_config in node_config_list:
        hostname, slots = parse_node_config(node_config)
        if hostname in out:
            raise ValueError(f"Duplicate hostname found in resource filter: {hostname}")
        out[hostname] = slots
    return out


def run_autotuning(args, active_resources):
    """``--autotuning {tune,run}``: explore ZeRO stage / micro-batch configurations with the Autotuner; ``run`` also
    launches the job with the best configuration found (reference ``launcher/runner.py:356-382``)."""
    from ..autotuning import Autotuner
    tuner = Autotuner(args, active_resources)
    logger.info("[Start] Running autotuning")
    tuner.tune()
    tuner.print_tuning_results()
    logger.info("[End] Running autotuning")
    tuner.write_optimal_config()
    if args.autotuning == "run":
        tuner.run_after_tuning()


def main(args=None):
    args = parse_args(args)
    resource_pool = fetch_hostfile(args.hostfile)
    cuda_visible = os.environ.get("CUDA_VISIBLE_DEVICES", "")
    if not resource_pool and cuda_visible and not (args.include or args.exclude):
        args.include = f"localhost:{cuda_visible}"
        del os.environ["CUDA_VISIBLE_DEVICES"]
    if args.num_nodes >= 0 or args.num_gpus >= 0:
        if args.include or args.exclude:
            raise ValueError("Cannot specify num_nodes/gpus with include/exclude")
    multi_node_exec = True
    if not resource_pool:
        n = torch.cuda.device_count() if torch.cuda.is_available() else 0
        if n == 0:
            raise RuntimeError("Unable to proceed, no GPU resources available")
        resource_pool = collections.OrderedDict(localhost=n)
        args.master_addr = "127.0.0.1"
        multi_node_exec = False
    active = parse_inclusion_exclusion(resource_pool, args.include, args.exclude)
    if args.num_nodes > 0:
        active = collections.OrderedDict(list(active.items())[: args.num_nodes])
    if args.num_gpus > 0:
        active = collections.OrderedDict((h, list(range(args.num_gpus))) for h in active)
    if not args.master_addr:
        first = list(active.keys())[0]
        try:
            out = subprocess.check_output(["ssh", first, "hostname -I"], text=True)
            args.master_addr = out.split()[0]
        except Exception:
            args.master_addr = first
    if args.autotuning != "":
        run_autotuning(args, active)
        return
    world_info = encode_world_info(active)
    multi_node_exec = args.force_multi or len(active) > 1
    env = os.environ.copy()
    if not multi_node_exec:
        cmd = [sys.executable, "-u", "-m", "shuffle_exchange_b200.launcher.launch", f"--world_info={world_info}",
               f"--master_addr={args.master_addr}", f"--master_port={args.master_port}"]
        for flag in ("module", "no_python", "no_local_rank", "bind_cores_to_rank"):
            if getattr(args, flag):
                cmd.append(f"--{flag}")
        if args.save_pid:
            cmd += ["--save_pid", f"{os.getpid()}"]
        if args.enable_each_rank_log != "None":
            cmd.append(f"--enable_each_rank_log={args.enable_each_rank_log}")
        if args.bind_core_list:
            cmd.append(f"--bind_core_list={args.bind_core_list}")
        cmd += [args.user_script] + args.user_args
    else:
        name = args.launcher.lower()
        pool = collections.OrderedDict((h, len(s)) for h, s in active.items())
        runner = {PDSH_LAUNCHER: lambda: PDSHRunner(args, world_info),
                  OPENMPI_LAUNCHER: lambda: OpenMPIRunner(args, world_info, pool),
                  MPICH_LAUNCHER: lambda: MPICHRunner(args, world_info, pool),
                  IMPI_LAUNCHER: lambda: IMPIRunner(args, world_info, pool),
                  SLURM_LAUNCHER: lambda: SlurmRunner(args, world_info, pool),
                  MVAPICH_LAUNCHER: lambda: MVAPICHRunner(args, world_info, pool)}.get(name)
        if runner is None:
            raise NotImplementedError(f"Unknown launcher {args.launcher}")
        runner = runner()
        if not runner.backend_exists():
            raise RuntimeError(f"launcher '{args.launcher}' not installed.")
        curr = os.path.abspath(".")
        env["PYTHONPATH"] = curr + (":" + env["PYTHONPATH"] if "PYTHONPATH" in env else "")
        for var in env:
            if any(var.startswith(n) for n in EXPORT_ENVS) and var not in ("PATH",):
                runner.add_export(var, env[var])
        for path in DEEPSPEED_ENVIRONMENT_PATHS:
            f = os.path.join(path, DEEPSPEED_ENVIRONMENT_NAME)
            if os.path.isfile(f):
                with open(f) as fd:
                    for line in fd:
                        if "=" in line:
                            k, v = line.strip().split("=", 1)
                            runner.add_export(k, v)
        got = runner.get_cmd(env, active)
        cmd = got[0] if isinstance(got, tuple) else got
        if isinstance(got, tuple) and len(got) == 3:
            env = got[2]
    logger.info(f"cmd = {' '.join(cmd)}")
    result = subprocess.Popen(cmd, env=env)

    def sigkill_handler(signum, frame):
        result.send_signal(signal.SIGINT)
        import time
        time.sleep(0.1)
        result.send_signal(signal.SIGTERM)
        result.wait()
        sys.exit(1)

    signal.signal(signal.SIGINT, sigkill_handler)
    signal.signal(signal.SIGTERM, sigkill_handler)
    result.wait()
    if result.returncode > 0:
        sys.exit(result.returncode)


if __name__ == "__main__":
    main()
