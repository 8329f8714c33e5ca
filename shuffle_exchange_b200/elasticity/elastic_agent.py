"""Per-node elastic agent (reference: ``deepspeed/elasticity/elastic_agent.py:32-191``): a
``torch.distributed.elastic`` ``LocalElasticAgent`` whose workers get the launcher's environment (so every restart sees
the same DeepSpeed variables) and which restarts the worker group on membership changes or failures up to
``max_restarts`` times."""
from __future__ import annotations

import os
import time
from typing import Any, Dict

from torch.distributed.elastic.agent.server.api import RunResult, WorkerGroup, WorkerSpec, WorkerState
from torch.distributed.elastic.agent.server.local_elastic_agent import LocalElasticAgent
from torch.distributed.elastic.metrics import put_metric
from torch.distributed.elastic.multiprocessing import Std

from ..utils.logging import logger


class DSElasticAgent(LocalElasticAgent):
    def __init__(self, spec: WorkerSpec, env: Dict, start_method="spawn", exit_barrier_timeout: float = 300,
                 log_dir=None):
        try:
            super().__init__(spec, start_method=start_method, exit_barrier_timeout=exit_barrier_timeout, log_dir=log_dir)
        except TypeError:      # newer torch: log_dir replaced by logs_specs
            super().__init__(spec, start_method=start_method, exit_barrier_timeout=exit_barrier_timeout)
        self.ds_env = env

    def _start_workers(self, worker_group: WorkerGroup) -> Dict[int, Any]:
        # inject the launcher environment into the entrypoint's environment before delegating
        saved = {}
        for k, v in (self.ds_env or {}).items():
            saved[k] = os.environ.get(k)
            os.environ[k] = str(v)
        try:
            return super()._start_workers(worker_group)
        finally:
            for k, v in saved.items():
                if v is None:
                    os.environ.pop(k, None)
                else:
                    os.environ[k] = v

    def _invoke_run(self, role: str = "default") -> RunResult:
        spec = self._worker_group.spec
        role = spec.role
        logger.info(f"[{role}] starting workers for entrypoint: {spec.get_entrypoint_name()}")
        self._initialize_workers(self._worker_group)
        monitor_interval = spec.monitor_interval
        rdzv_handler = spec.rdzv_handler
        participants = rdzv_handler._state_holder.state.participants if hasattr(rdzv_handler, "_state_holder") else None
        while True:
            assert self._worker_group.state != WorkerState.INIT
            time.sleep(monitor_interval)
            run_result = self._monitor_workers(self._worker_group)
            state = run_result.state
            self._worker_group.state = state
            expire = getattr(getattr(rdzv_handler, "_settings", None), "keep_alive_interval", None)
            put_metric(f"workers.{role}.remaining_restarts", self._remaining_restarts)
            put_metric(f"workers.{role}.{state.name.lower()}", 1)
            if state == WorkerState.SUCCEEDED:
                logger.info(f"[{role}] worker group successfully finished. Waiting {self._exit_barrier_timeout} seconds for other agents to finish.")
                self._exit_barrier()
                return run_result
            if state in {WorkerState.UNHEALTHY, WorkerState.FAILED}:
                if self._remaining_restarts > 0:
                    logger.info(f"[{role}] Worker group {state.name}. {self._remaining_restarts}/{spec.max_restarts} attempts left; will restart worker group")
                    self._remaining_restarts -= 1
                    self._restart_workers(self._worker_group)
                else:
                    self._stop_workers(self._worker_group)
                    self._worker_group.state = WorkerState.FAILED
                    return run_result
            elif state == WorkerState.HEALTHY:
                # membership change (nodes joined / left): restart on the new world
                changed = False
                try:
                    changed = rdzv_handler.num_nodes_waiting() > 0
                except Exception:
                    pass
                if changed:
                    logger.info(f"[{role}] Detected membership change, will restart worker group")
                    self._restart_workers(self._worker_group)
            else:
                raise Exception(f"[{role}] Worker group in {state.name} state")


def run_elastic(args, env, local_gpus, world_size):
    """Entry used by ``launcher.launch --enable_elastic_training``: run the training script under a
    :class:`DSElasticAgent` with a c10d rendezvous on ``MASTER_ADDR:MASTER_PORT``."""
    import sys
    import torch.distributed.elastic.rendezvous.registry as rdzv_registry
    from torch.distributed.elastic.rendezvous import RendezvousParameters
    cmd = ([] if args.no_python else [sys.executable, "-u"] + (["-m"] if args.module else [])) + [args.training_script] + args.training_script_args
    min_nodes = args.min_elastic_nodes if args.min_elastic_nodes > 0 else 1
    max_nodes = args.max_elastic_nodes if args.max_elastic_nodes > 0 else max(1, world_size // max(1, len(local_gpus)))
    params = RendezvousParameters(backend="c10d", endpoint=f"{args.master_addr}:{args.master_port}",
                                  run_id=os.environ.get("ELASTIC_RUN_ID", "sxb200_elastic"), min_nodes=min_nodes,
                                  max_nodes=max_nodes)
    spec = WorkerSpec(role="trainer", local_world_size=len(local_gpus), entrypoint=cmd[0], args=tuple(cmd[1:]),
                      rdzv_handler=rdzv_registry.get_rendezvous_handler(params), max_restarts=100, monitor_interval=5,
                      redirects=Std.from_str("0"), tee=Std.from_str("0"), master_addr=None, master_port=None)
    agent = DSElasticAgent(spec, env)
    result = agent.run()
    return 0 if not result.is_failed() else 1
