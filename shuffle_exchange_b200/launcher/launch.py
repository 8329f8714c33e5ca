"""Per-node process spawner: one training process per local GPU.

Reference: deepspeed/launcher/launch.py — decodes the world layout, exports ``RANK / LOCAL_RANK /
WORLD_SIZE / MASTER_ADDR / MASTER_PORT``, ``Popen``s one process per GPU (:133-250), polls the children
every second and, on any non-zero exit, kills the whole tree and exits with that code (:317-355),
forwards SIGINT/SIGTERM, optional torch-elastic agent path (:252-305).
"""
from __future__ import annotations

import argparse
import base64
import json
import os
import signal
import subprocess
import sys
import time
from collections import defaultdict

from ..utils.logging import logger
from .constants import ELASTIC_TRAINING_ID_DEFAULT, PID_FILE_BASEPATH

try:
    import psutil
except ImportError:  # pragma: no cover
    psutil = None


def parse_args(args=None):
    p = argparse.ArgumentParser(description="sxb200 per-node launcher: spawns one process per local accelerator")
    p.add_argument("--node_rank", type=int, default=0)
    p.add_argument("--master_addr", default="127.0.0.1", type=str)
    p.add_argument("--master_port", default=29500, type=int)
    p.add_argument("--world_info", default="None", type=str, help="base64 json {host: [gpu ids]}")
    p.add_argument("--module", action="store_true", help="run the script as `python -m`")
    p.add_argument("--no_python", action="store_true", help="exec the script directly")
    p.add_argument("--enable_elastic_training", action="store_true")
    p.add_argument("--min_elastic_nodes", type=int, default=-1)
    p.add_argument("--max_elastic_nodes", type=int, default=-1)
    p.add_argument("--no_local_rank", action="store_true", help="do not pass --local_rank to the script")
    p.add_argument("--save_pid", type=int, default=0)
    p.add_argument("--enable_each_rank_log", default="None", type=str, help="directory for per-rank logs")
    p.add_argument("--bind_cores_to_rank", action="store_true")
    p.add_argument("--bind_core_list", type=str, default=None)
    p.add_argument("training_script", type=str)
    p.add_argument("training_script_args", nargs=argparse.REMAINDER)
    return p.parse_args(args=args)


def decode_world_info(encoded: str) -> dict:
    return json.loads(base64.urlsafe_b64decode(encoded))


def global_rank_mapping(world_info: dict):
    """{host: [global ranks]} with ranks assigned host by host in dict order."""
    mapping, nxt = defaultdict(list), 0
    for host, gpus in world_info.items():
        for _ in gpus:
            mapping[host].append(nxt)
            nxt += 1
    return mapping, nxt


def terminate_process_tree(pid: int):
    if psutil is None:
        try:
            os.kill(pid, signal.SIGTERM)
        except ProcessLookupError:
            pass
        return
    try:
        proc = psutil.Process(pid)
    except psutil.NoSuchProcess:
        return
    procs = proc.children(recursive=True) + [proc]
    for p in procs:
        try:
            p.terminate()
        except psutil.NoSuchProcess:
            pass
    _, alive = psutil.wait_procs(procs, timeout=30)
    for p in alive:
        try:
            p.kill()
        except psutil.NoSuchProcess:
            pass


def build_command(args, local_rank: int):
    cmd = []
    if args.bind_cores_to_rank:
        from ..utils.numa import get_numactl_cmd
        _, numactl = get_numactl_cmd(args.bind_core_list, args._num_local_procs, local_rank)
        cmd += numactl
    if not args.no_python:
        cmd += [sys.executable, "-u"]
        if args.module:
            cmd.append("-m")
    elif args.module:
        raise ValueError("Don't use both the '--no_python' flag and the '--module' flag at the same time.")
    cmd.append(args.training_script)
    if not args.no_local_rank:
        cmd.append(f"--local_rank={local_rank}")
    cmd += args.training_script_args
    return cmd


def main(argv=None):
    args = parse_args(argv)
    env = os.environ.copy()
    for k in list(env):
        if "NCCL" in k:
            logger.info(f"{args.node_rank} {k}={env[k]}")
    if args.world_info == "None":
        raise ValueError("world_info can not be None")
    world_info = decode_world_info(args.world_info)
    hosts = list(world_info.keys())
    node = hosts[args.node_rank]
    local_gpus = world_info[node]
    mapping, world_size = global_rank_mapping(world_info)
    args._num_local_procs = len(local_gpus)
    logger.info(f"WORLD INFO DICT: {world_info}")
    logger.info(f"nnodes={len(hosts)}, num_local_procs={len(local_gpus)}, node_rank={args.node_rank}")
    logger.info(f"global_rank_mapping={dict(mapping)}, dist_world_size={world_size}")
    env["CUDA_VISIBLE_DEVICES"] = ",".join(map(str, local_gpus))
    env["MASTER_ADDR"], env["MASTER_PORT"] = args.master_addr, str(args.master_port)
    env["WORLD_SIZE"], env["CROSS_RANK"], env["CROSS_SIZE"] = str(world_size), str(args.node_rank), str(len(hosts))
    env["LOCAL_SIZE"] = str(len(local_gpus))
    if args.save_pid:
        pid_file = os.path.join(PID_FILE_BASEPATH, f"{args.save_pid}.sxb200")
        with open(pid_file, "w") as fd:
            fd.write(f"{os.getpid()}")
    if args.enable_elastic_training:
        from ..elasticity.elastic_agent import run_elastic
        return run_elastic(args, env, local_gpus, world_size)
    log_dir = None if args.enable_each_rank_log == "None" else args.enable_each_rank_log
    if log_dir:
        os.makedirs(log_dir, exist_ok=True)
    processes, last_return_code = [], None
    for local_rank, global_rank in enumerate(mapping[node]):
        e = env.copy()
        e["RANK"], e["LOCAL_RANK"] = str(global_rank), str(local_rank)
        e.setdefault("OMP_NUM_THREADS", str(max(1, (os.cpu_count() or 1) // max(1, len(local_gpus)))))
        cmd = build_command(args, local_rank)
        if log_dir:
            f = open(os.path.join(log_dir, f"rank{global_rank}.log"), "w")
            processes.append(subprocess.Popen(cmd, env=e, stdout=f, stderr=f))
        else:
            processes.append(subprocess.Popen(cmd, env=e))
        logger.info(f"process {processes[-1].pid} spawned with command: {cmd}")

    def on_signal(signum, frame):
        for p in processes:
            logger.info(f"Killing subprocess {p.pid}")
            terminate_process_tree(p.pid)
        if last_return_code is not None:
            logger.error(f"{cmd} exits with return code = {last_return_code}")
            sys.exit(last_return_code)
        sys.exit(1 if signum in (signal.SIGINT, signal.SIGTERM) else 0)

    signal.signal(signal.SIGINT, on_signal)
    signal.signal(signal.SIGTERM, on_signal)
    alive = set(processes)
    while alive:
        done = set()
        for p in alive:
            rc = p.poll()
            if rc is None:
                continue
            if rc != 0:
                last_return_code = rc
                on_signal(signal.SIGTERM, None)       # a failing rank takes the whole job down
            done.add(p)
            logger.info(f"Process {p.pid} exits successfully.")
        alive -= done
        time.sleep(1)
    return 0


if __name__ == "__main__":
    sys.exit(main())
