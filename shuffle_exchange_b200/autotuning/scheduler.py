"""Experiment scheduler of the autotuner (reference: ``deepspeed/autotuning/scheduler.py:32-440``): a pool of
``Node``s with GPU slots, a queue of experiments (each a DeepSpeed config + the user's launch command), reservations,
one launcher subprocess per experiment, result parsing (the engine writes ``metrics.json`` at ``autotuning.metric_path``).
"""
from __future__ import annotations

import base64
import copy
import json
import os
import subprocess
import sys
import threading
import time
from typing import Dict, List, Optional, Tuple

from ..utils.logging import logger
from .constants import *  # noqa: F401,F403
from .utils import find_replace, get_val_by_key, search_error, set_val_by_key, was_interruptted

TIMEOUT = 5


class Node:
    def __init__(self, host, max_slots):
        self.host, self.max_slots = host, max_slots
        self.idle_slots = list(range(max_slots))

    def reserve_slots(self, slot_request: int) -> Optional[list]:
        if len(self.idle_slots) >= slot_request:
            got, self.idle_slots = self.idle_slots[:slot_request], self.idle_slots[slot_request:]
            return got
        return None

    def restore_slots(self, slots: list):
        self.idle_slots += slots


class Reservation:
    def __init__(self, node, slots):
        self.node, self.slots = node, slots

    def restore_slots(self):
        self.node.restore_slots(self.slots)

    def desc(self):
        return f"{self.node.host}:{','.join(map(str, self.slots))}@"


class Job:
    def __init__(self, exp, reservations, process):
        self.exp, self.reservations, self.process = exp, reservations, process


class ResourceManager:
    def __init__(self, args, hosts, num_gpus_per_node, results_dir, exps_dir, arg_mappings):
        self.results_dir, self.exps_dir = results_dir, exps_dir
        self.nodes = [Node(h, num_gpus_per_node) for h in hosts]
        self.experiment_queue: List[dict] = []
        self.running_experiments: Dict[int, Tuple] = {}
        self.finished_experiments: Dict[int, Tuple[dict, Optional[str]]] = {}
        self.experiment_count = 0
        self.exp_paths = set()
        self.args = args
        self.arg_mappings = {}
        if arg_mappings is not None:
            for k, v in arg_mappings.items():
                k, v = k.strip(), v.strip()
                self.arg_mappings.setdefault(k, v)

    def schedule_experiments(self, exp_paths):
        for path in exp_paths:
            if path in self.exp_paths:
                continue
            self.exp_paths.add(path)
            with open(path) as fd:
                exp = json.load(fd)
            exp["exp_id"] = self.experiment_count
            self.experiment_count += 1
            result_dir = exp["result_dir"] = os.path.join(self.results_dir, exp["name"])
            if AUTOTUNING in exp["ds_config"]:
                exp["ds_config"][AUTOTUNING][AUTOTUNING_METRIC_PATH] = os.path.join(result_dir, "metrics.json")
            stderr = os.path.join(result_dir, "stderr.log")
            model_info_file = os.path.join(result_dir, "model_info.json")
            metric_file = os.path.join(result_dir, "metrics.json")
            if os.path.exists(result_dir) and os.path.exists(stderr) and not was_interruptted(stderr):
                err = search_error(stderr)
                exp_id = exp["exp_id"]
                if os.path.exists(metric_file) or os.path.exists(model_info_file):
                    logger.info(f"Skipping exp {exp['name']} whose result already exists")
                    self.finished_experiments[exp_id] = (exp, err)
                    continue
            self.experiment_queue.append(exp)

    def run_job(self, exp: dict, reservations):
        exp_id = exp["exp_id"]
        exp["master_port"] = self.args.master_port + exp_id if hasattr(self.args, "master_port") else 29500 + exp_id
        exp["result_dir"] = os.path.join(self.results_dir, exp["name"])
        user_script, user_args = self.args.user_script, list(self.args.user_args)
        for k, v in self.arg_mappings.items():       # forward tuned values into user command-line arguments
            val = get_val_by_key(exp["ds_config"], k)
            if val is not None and str(val) != "auto" and v in user_args:
                user_args[user_args.index(v) + 1] = str(val)
        exp["user_script"], exp["user_args"] = user_script, user_args
        exp["hostfile"] = getattr(self.args, "hostfile", None)
        exp["reservations"] = [r.desc() for r in reservations]
        t = threading.Thread(target=run_experiment, args=(exp, reservations, user_script, user_args))
        t.start()
        self.running_experiments[exp_id] = (t, exp, reservations, time.time())

    def experiment_check(self, pbar=None):
        done = []
        for exp_id, (t, exp, reservations, start) in self.running_experiments.items():
            t.join(timeout=TIMEOUT)
            if not t.is_alive():
                err = search_error(os.path.join(exp["result_dir"], "stderr.log"))
                done.append((exp_id, exp, err))
                for r in reservations:
                    r.restore_slots()
        for exp_id, exp, err in done:
            del self.running_experiments[exp_id]
            self.finished_experiments[exp_id] = (exp, err)
            logger.info(f"Done running exp_id = {exp_id}, exp_name = {exp['name']}" + (f", error = {err}" if err else ""))

    def resource_request(self, exp):
        num_nodes, num_gpus = exp.get("num_nodes", 1), exp.get("num_gpus", 1)
        reservations = []
        for node in self.nodes:
            if len(reservations) == num_nodes:
                break
            slots = node.reserve_slots(slot_request=num_gpus)
            if slots:
                reservations.append(Reservation(node=node, slots=slots))
        if len(reservations) == num_nodes:
            return reservations
        for r in reservations:
            r.restore_slots()
        return None

    def status(self):
        return ", ".join(f"{n.host} ({len(n.idle_slots)} idle gpus)" for n in self.nodes)

    def run(self):
        while self.experiment_queue:
            exp = self.experiment_queue.pop(0)
            reservations = self.resource_request(exp)
            if not reservations:
                self.experiment_queue.insert(0, exp)
                logger.debug(f"Unable to schedule exp {exp['name']}; waiting for resources: {self.status()}")
                self.experiment_check()
            else:
                logger.info(f"Scheduling exp_id = {exp['exp_id']}, exp_name = {exp['name']}, reservation = {[r.desc() for r in reservations]}")
                self.run_job(exp, reservations)
        while self.running_experiments:
            self.experiment_check()

    def save_exp_results_to_database(self, message, ranks=None, path=None):
        if path is None:
            return
        with open(path, "a") as fd:
            fd.write(message + "\n")

    def read_metric(self, exp, metric):
        f = exp["ds_config"].get(AUTOTUNING, {}).get(AUTOTUNING_METRIC_PATH)
        if f and os.path.exists(f):
            with open(f) as fh:
                return json.load(fh).get(metric)
        return None

    def parse_results(self, metric):
        """Best finished experiment by ``metric`` (higher is better; latency is inverted by the engine when written)."""
        best_val, best_exp = None, None
        for exp_id, (exp, err) in self.finished_experiments.items():
            if err:
                logger.info(f"The experiment exp_id = {exp_id}, exp_name = {exp['name']}, did not run successfully with error = {err}")
                continue
            val = self.read_metric(exp, metric)
            exp["results"] = {metric: val} if val is not None else {}
            if val is not None and (best_val is None or val > best_val):
                best_val, best_exp = val, exp
        return best_exp, best_val

    def clear(self):
        self.experiment_queue = []
        for exp_id, (t, exp, reservations, _) in self.running_experiments.items():
            clean_up(exp, reservations)
        self.running_experiments = {}
        self.finished_experiments = {}
        self.exp_paths = set()


def run_experiment(exp: dict, reservations, user_script, user_args):
    include = "".join(r.desc() for r in reservations)[:-1]
    master_port = exp["master_port"]
    result_dir = exp["result_dir"]
    os.makedirs(result_dir, exist_ok=True)
    ds_config_path = os.path.join(result_dir, "ds_config.json")
    exp["ds_config_path"] = ds_config_path
    ds_config = copy.deepcopy(exp["ds_config"])
    ds_config_b64 = base64.urlsafe_b64encode(json.dumps(ds_config).encode("utf-8")).decode("utf-8")
    exp["ds_config_base64"] = ds_config_b64
    with open(ds_config_path, "w", buffering=BUFSIZE) as fd:
        json.dump(ds_config, fd)
        fd.flush()
        os.fsync(fd)
    # point the user's --deepspeed_config / --deepspeed argument at the experiment's config
    new_args = list(user_args)
    for flag in ("--deepspeed_config", "--deepspeed"):
        if flag in new_args:
            idx = new_args.index(flag)
            if idx + 1 < len(new_args) and not new_args[idx + 1].startswith("--"):
                new_args[idx + 1] = ds_config_path
    with open(os.path.join(result_dir, "exp.json"), "w", buffering=BUFSIZE) as fd:
        json.dump(exp, fd)
        fd.flush()
        os.fsync(fd)
    cmd = [sys.executable, "-m", "shuffle_exchange_b200.launcher.runner", "--include", include, "--master_port", str(master_port)]
    if exp.get("hostfile"):
        cmd += ["--hostfile", exp["hostfile"]]
    cmd += [user_script] + new_args
    with open(os.path.join(result_dir, "cmd.txt"), "w", buffering=BUFSIZE) as fd:
        fd.write(" ".join(cmd) + "\n")
    with open(os.path.join(result_dir, "stdout.log"), "wb") as out, open(os.path.join(result_dir, "stderr.log"), "wb") as err:
        proc = subprocess.Popen(cmd, stdout=out, stderr=err)
        proc.wait()
    clean_up(exp, reservations)


def clean_up(exp: dict, reservations):
    """Nothing survives a finished launcher: its process tree is torn down by ``launcher.launch`` itself.  (The reference
    pdsh-kills stragglers by command-line pattern on every host; this framework never kills by pattern.)"""
    return None
