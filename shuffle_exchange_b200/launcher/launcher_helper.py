"""Per-process shim used by MPI-style multinode launchers (reference: ``launcher/launcher_helper.py``).

``mpirun`` starts one copy of this helper per rank; it translates the MPI implementation's rank / size environment
variables into the names the framework reads (``RANK``, ``LOCAL_RANK``, ``WORLD_SIZE``, ``CROSS_RANK`` …), optionally
binds the process to its share of cores, and then executes the user script in-process."""
from __future__ import annotations

import argparse
import os
import runpy
import sys

from ..utils.logging import logger
from ..utils.numa import get_numactl_cmd

_ENV_SOURCES = {
    "openmpi": ("OMPI_COMM_WORLD_RANK", "OMPI_COMM_WORLD_LOCAL_RANK", "OMPI_COMM_WORLD_SIZE", "OMPI_COMM_WORLD_LOCAL_SIZE"),
    "mpich": ("PMI_RANK", "MPI_LOCALRANKID", "PMI_SIZE", "MPI_LOCALNRANKS"),
    "impi": ("PMI_RANK", "MPI_LOCALRANKID", "PMI_SIZE", "MPI_LOCALNRANKS"),
    "mvapich": ("MV2_COMM_WORLD_RANK", "MV2_COMM_WORLD_LOCAL_RANK", "MV2_COMM_WORLD_SIZE", "MV2_COMM_WORLD_LOCAL_SIZE"),
    "slurm": ("SLURM_PROCID", "SLURM_LOCALID", "SLURM_NTASKS", "SLURM_NTASKS_PER_NODE"),
}


def parse_args(args=None):
    p = argparse.ArgumentParser(description="launcher helper: map the MPI environment and run the user script")
    p.add_argument("--launcher", default="openmpi", type=str, help="multi-node launcher backend: " + ", ".join(_ENV_SOURCES))
    p.add_argument("--module", action="store_true", help="treat user_script as a python module (python -m)")
    p.add_argument("--no_python", action="store_true", help="exec user_script directly instead of through python")
    p.add_argument("--bind_cores_to_rank", action="store_true", help="bind the process to its share of the host's cores")
    p.add_argument("--bind_core_list", type=str, default=None, help="cores to divide among local ranks, e.g. 0-27,32-59")
    p.add_argument("user_script", type=str)
    p.add_argument("user_args", nargs=argparse.REMAINDER)
    return p.parse_args(args=args)


def env_mapping(env: dict, rank_name_list=None, local_rank_name_list=None) -> dict:
    """Set RANK / LOCAL_RANK from the first variable of each list that is present; all listed variables that are
    present must agree."""
    def pick(names, what):
        found = {n: env[n] for n in (names or ()) if n in env}
        if not found:
            raise EnvironmentError(f"no {what} variable found among {list(names or ())}")
        if len(set(found.values())) != 1:
            raise EnvironmentError(f"{what} variables disagree: {found}")
        return next(iter(found.values()))

    env["RANK"] = pick(rank_name_list, "rank")
    env["LOCAL_RANK"] = pick(local_rank_name_list, "local rank")
    return env


def main(args=None):
    args = parse_args(args)
    env = os.environ
    key = args.launcher.lower()
    if key not in _ENV_SOURCES:
        raise NotImplementedError(f"Unknown launcher {args.launcher}")
    rank_v, local_v, size_v, local_size_v = _ENV_SOURCES[key]
    env_mapping(env, rank_name_list=[rank_v], local_rank_name_list=[local_v])
    env["WORLD_SIZE"] = env.get(size_v, env.get("WORLD_SIZE", "1"))
    local_size = int(env.get(local_size_v, env.get("LOCAL_SIZE", "1")))
    env["LOCAL_SIZE"] = str(local_size)
    env["CROSS_SIZE"] = str(max(1, int(env["WORLD_SIZE"]) // local_size))
    env["CROSS_RANK"] = str(int(env["RANK"]) // local_size)
    env.setdefault("MASTER_ADDR", "127.0.0.1")
    env.setdefault("MASTER_PORT", "29500")
    cmd_prefix = []
    if args.bind_cores_to_rank:
        cores_per_rank, numactl = get_numactl_cmd(args.bind_core_list, local_size, int(env["LOCAL_RANK"]))
        env["OMP_NUM_THREADS"] = str(cores_per_rank)
        cmd_prefix = numactl
    logger.info(f"launcher_helper rank={env['RANK']} local_rank={env['LOCAL_RANK']} world={env['WORLD_SIZE']}")
    if args.no_python or cmd_prefix:
        cmd = cmd_prefix + ([] if args.no_python else [sys.executable, "-u"] + (["-m"] if args.module else []))
        cmd += [args.user_script] + args.user_args
        os.execvpe(cmd[0], cmd, env)
    sys.argv = [args.user_script] + args.user_args
    if args.module:
        runpy.run_module(args.user_script, run_name="__main__", alter_sys=True)
    else:
        runpy.run_path(args.user_script, run_name="__main__")


if __name__ == "__main__":
    main()
