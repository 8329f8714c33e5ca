"""Per-node elastic agent (reference: ``deepspeed/elasticity/elastic_agent.py:32-191``): a
``torch.distributed.elastic`` ``LocalElasticAgent`` whose workers get the launcher's environment (so every restart sees
the same DeepSpeed variables) and which restarts the worker group on membership changes or failures up to
``max_restarts`` times."""
from __future__ import annotations

import os
from typing import Any, Dict

from torch.distributed.elastic.agent.server.api import RunResult, WorkerGroup, WorkerSpec, WorkerState
from torch.distributed.elastic.agent.server.local_elastic_agent import LocalElasticAgent

from ..utils.logging import logger


def _participants(rdzv_handler) -> int:
    """Number of nodes in the current rendezvous round (dynamic / c10d rendezvous keeps it in its state holder)."""
    holder = getattr(rdzv_handler, "_state_holder", None)
    state = getattr(holder, "state", None)
    return len(getattr(state, "participants", ()) or ())


class DSElasticAgent(LocalElasticAgent):
    """``LocalElasticAgent`` with two additions:

    * workers start with the launcher's environment ``env`` on top of the agent's own, so every restart sees the same
      launcher variables;
    * a node LEAVING the rendezvous makes the group unhealthy.  torch's agent loop restarts the group when workers
      fail or when nodes are waiting to join; the reference's copy of that loop (``elastic_agent.py:127-191``) adds the
      scale-down case by comparing the participant count with the one at the last (re)start.  Here the same rule sits in
      ``_monitor_workers``, so the stock loop does the restarting (and counts it against ``max_restarts``)."""

    def __init__(self, spec: WorkerSpec, env: Dict, start_method="spawn", exit_barrier_timeout: float = 300,
                 log_dir=None):
        try:
            from torch.distributed.elastic.multiprocessing import DefaultLogsSpecs
            super().__init__(spec, logs_specs=DefaultLogsSpecs(log_dir=log_dir), start_method=start_method,
                             exit_barrier_timeout=exit_barrier_timeout)
        except (ImportError, TypeError):       # torch < 2.2: the agent takes log_dir itself
            super().__init__(spec, start_method=start_method, exit_barrier_timeout=exit_barrier_timeout,
                             log_dir=log_dir)
        self.ds_env = env
        self._participants_at_start = 0

    def _start_workers(self, worker_group: WorkerGroup) -> Dict[int, Any]:
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

    def _initialize_workers(self, worker_group: WorkerGroup) -> None:
        super()._initialize_workers(worker_group)
        self._participants_at_start = _participants(worker_group.spec.rdzv_handler)

    def _monitor_workers(self, worker_group: WorkerGroup) -> RunResult:
        result = super()._monitor_workers(worker_group)
        if result.state == WorkerState.HEALTHY:
            now = _participants(worker_group.spec.rdzv_handler)
            if 0 < now < self._participants_at_start:
                logger.info(f"[{worker_group.spec.role}] {self._participants_at_start - now} node(s) left the "
                            "rendezvous; the worker group will be restarted on the smaller world")
                return RunResult(state=WorkerState.UNHEALTHY)
        return result


def run_elastic(args, env, local_gpus, world_size):
    """Entry used by ``launcher.launch --enable_elastic_training``: run the training script under a
    :class:`DSElasticAgent` with a c10d rendezvous on ``MASTER_ADDR:MASTER_PORT``."""
    import inspect
    import sys
    import torch.distributed.elastic.rendezvous.registry as rdzv_registry
    from torch.distributed.elastic.rendezvous import RendezvousParameters
    cmd = ([] if args.no_python else [sys.executable, "-u"] + (["-m"] if args.module else [])) + [args.training_script] + args.training_script_args
    min_nodes = args.min_elastic_nodes if args.min_elastic_nodes > 0 else 1
    max_nodes = args.max_elastic_nodes if args.max_elastic_nodes > 0 else max(1, world_size // max(1, len(local_gpus)))
    params = RendezvousParameters(backend="c10d", endpoint=f"{args.master_addr}:{args.master_port}",
                                  run_id=os.environ.get("ELASTIC_RUN_ID", "sxb200_elastic"), min_nodes=min_nodes,
                                  max_nodes=max_nodes)
    kw = dict(role="trainer", local_world_size=len(local_gpus), entrypoint=cmd[0], args=tuple(cmd[1:]),
              rdzv_handler=rdzv_registry.get_rendezvous_handler(params),
              max_restarts=int(getattr(args, "max_elastic_restarts", 100)),
              monitor_interval=float(getattr(args, "elastic_monitor_interval", 5)), master_addr=None, master_port=None)
    if "redirects" in inspect.signature(WorkerSpec.__init__).parameters:      # torch < 2.2 kept log routing here
        from torch.distributed.elastic.multiprocessing import Std
        kw.update(redirects=Std.from_str("0"), tee=Std.from_str("0"))
    agent = DSElasticAgent(WorkerSpec(**kw), env)
    result = agent.run()
    return 0 if not result.is_failed() else 1
