"""Multi-node launch back-ends: build the command line that starts ``launcher.launch`` on every node.

Reference: deepspeed/launcher/multinode_runner.py — PDSH :55, OpenMPI :126, MPICH :206, IMPI :278,
Slurm :363, MVAPICH :411.  Each runner exposes ``backend_exists()`` and ``get_cmd(environment,
active_resources)``; exports are forwarded in the runner's native syntax."""
from __future__ import annotations

import os
import shutil
import sys
from abc import ABC, abstractmethod
from shlex import quote, split

from .constants import MVAPICH_TMP_HOSTFILE, PDSH_MAX_FAN_OUT


class MultiNodeRunner(ABC):
    def __init__(self, args, world_info_base64):
        self.args = args
        self.user_arguments = list(map(quote, args.user_args)) if getattr(args, "user_args", None) else []
        self.user_script = args.user_script
        self.world_info_base64 = world_info_base64
        self.exports = {}

    @abstractmethod
    def backend_exists(self): ...

    @abstractmethod
    def get_cmd(self, environment, active_resources): ...

    def add_export(self, key, var):
        self.exports[key.strip()] = var.strip()

    def parse_user_args(self):
        return self.user_arguments

    @property
    def name(self):
        return self.__class__.__name__

    def validate_args(self):
        pass

    def _launch_module(self):
        return [sys.executable, "-u", "-m", "shuffle_exchange_b200.launcher.launch",
                f"--world_info={self.world_info_base64}", f"--master_addr={self.args.master_addr}",
                f"--master_port={self.args.master_port}"]

    def _script(self):
        flags = []
        if getattr(self.args, "no_python", False):
            flags.append("--no_python")
        if getattr(self.args, "module", False):
            flags.append("--module")
        if getattr(self.args, "no_local_rank", False):
            flags.append("--no_local_rank")
        if getattr(self.args, "save_pid", False):
            flags += ["--save_pid", f"{os.getpid()}"]
        if getattr(self.args, "elastic_training", False):
            flags += ["--enable_elastic_training", f"--max_elastic_nodes={self.args.max_elastic_nodes}",
                      f"--min_elastic_nodes={self.args.min_elastic_nodes}"]
        return flags + [self.user_script] + self.user_arguments


class PDSHRunner(MultiNodeRunner):
    def backend_exists(self):
        return shutil.which("pdsh") is not None

    def get_cmd(self, environment, active_resources):
        environment["PDSH_RCMD_TYPE"] = "ssh"
        if getattr(self.args, "ssh_port", None) is not None:
            environment["PDSH_SSH_ARGS_APPEND"] = f"{environment.get('PDSH_SSH_ARGS_APPEND', '')} -p {self.args.ssh_port}"
        workers = ",".join(active_resources.keys())
        pdsh = ["pdsh", "-S", "-f", str(PDSH_MAX_FAN_OUT), "-w", workers] + split(getattr(self.args, "launcher_args", "") or "")
        exports = "".join(f"export {k}={quote(v)}; " for k, v in self.exports.items())
        launch = [exports, f"cd {os.path.abspath('.')};"] + self._launch_module() + ["--node_rank=%n"]
        return pdsh + launch + self._script(), ["pkill", "-f", " ".join(launch[2:-1])], environment


class OpenMPIRunner(MultiNodeRunner):
    def __init__(self, args, world_info_base64, resource_pool):
        super().__init__(args, world_info_base64)
        self.resource_pool = resource_pool
        self.add_export("UCX_TLS", "tcp")

    def backend_exists(self):
        return shutil.which("ompi_info") is not None

    def validate_args(self):
        if getattr(self.args, "include", "") or getattr(self.args, "exclude", ""):
            raise ValueError(f"{self.name} backend does not support worker include/exclusion")
        if getattr(self.args, "num_nodes", -1) != -1 or getattr(self.args, "num_gpus", -1) != -1:
            raise ValueError(f"{self.name} backend does not support limiting num nodes/gpus")

    def get_cmd(self, environment, active_resources):
        total = sum(self.resource_pool.values())
        cmd = ["mpirun", "-n", f"{total}", "-hostfile", f"{self.args.hostfile}", "--mca", "btl", "^openib",
               "--mca", "btl_tcp_if_include", "eth0"] + split(getattr(self.args, "launcher_args", "") or "")
        for k, v in self.exports.items():
            cmd += ["-x", f"{k}={quote(v)}"]
        py = [] if getattr(self.args, "no_python", False) else [sys.executable, "-u"] + (["-m"] if getattr(self.args, "module", False) else [])
        return cmd + py + [self.user_script] + self.user_arguments


class MPICHRunner(OpenMPIRunner):
    def backend_exists(self):
        return shutil.which("mpirun") is not None

    def get_cmd(self, environment, active_resources):
        total = sum(self.resource_pool.values())
        per_node = list(self.resource_pool.values())[0]
        cmd = ["mpirun", "-n", f"{total}", "-ppn", f"{per_node}"] + split(getattr(self.args, "launcher_args", "") or "")
        for k, v in self.exports.items():
            cmd += ["-genv", k, quote(v)]
        cmd += ["-genv", "MASTER_ADDR", str(self.args.master_addr), "-genv", "MASTER_PORT", str(self.args.master_port),
                "-genv", "WORLD_SIZE", str(total), "-hosts", ",".join(self.resource_pool.keys())]
        py = [] if getattr(self.args, "no_python", False) else [sys.executable, "-u"] + (["-m"] if getattr(self.args, "module", False) else [])
        return cmd + py + [self.user_script] + self.user_arguments


class IMPIRunner(MPICHRunner):
    pass


class SlurmRunner(MultiNodeRunner):
    def __init__(self, args, world_info_base64, resource_pool):
        super().__init__(args, world_info_base64)
        self.resource_pool = resource_pool

    def backend_exists(self):
        return shutil.which("sinfo") is not None

    def get_cmd(self, environment, active_resources):
        total = sum(self.resource_pool.values())
        cmd = ["srun", "-n", f"{total}"] + split(getattr(self.args, "launcher_args", "") or "")
        if getattr(self.args, "comment", ""):
            cmd += ["--comment", self.args.comment]
        if getattr(self.args, "include", ""):
            cmd += ["--include", f"{self.args.include}"]
        if getattr(self.args, "exclude", ""):
            cmd += ["--exclude", f"{self.args.exclude}"]
        if getattr(self.args, "num_nodes", -1) > 0:
            cmd += ["--nodes", f"{self.args.num_nodes}"]
        if getattr(self.args, "num_gpus", -1) > 0:
            cmd += ["--gpus", f"{self.args.num_gpus}"]
        exports = "--export=ALL" + "".join(f",{k}={quote(v)}" for k, v in self.exports.items())
        return cmd + [exports, sys.executable, "-u", self.user_script] + self.user_arguments


class MVAPICHRunner(MultiNodeRunner):
    def __init__(self, args, world_info_base64, resource_pool):
        super().__init__(args, world_info_base64)
        self.resource_pool = resource_pool
        for k, v in (("MV2_SMP_USE_CMA", "0"), ("MV2_DEBUG_SHOW_BACKTRACE", "1"), ("MV2_USE_CUDA", "1"),
                     ("MV2_SUPPORT_DL", "1"), ("MV2_ENABLE_AFFINITY", "0"), ("MV2_INTER_ALLGATHER_TUNING", "5"),
                     ("MV2_CUDA_USE_NAIVE", "0")):
            self.add_export(k, v)

    def backend_exists(self):
        return shutil.which("mpiname") is not None

    def get_cmd(self, environment, active_resources):
        total = sum(self.resource_pool.values())
        per_node = list(self.resource_pool.values())[0]
        with open(MVAPICH_TMP_HOSTFILE, "w") as fd:
            for host in self.resource_pool:
                fd.write(f"{host}\n")
        cmd = ["mpirun", "-np", f"{total}", "-ppn", f"{per_node}", "--hostfile", MVAPICH_TMP_HOSTFILE] + \
            split(getattr(self.args, "launcher_args", "") or "")
        for k, v in self.exports.items():
            cmd += ["-env", f"{k}={quote(v)}"]
        return cmd + [sys.executable, "-u", self.user_script] + self.user_arguments
