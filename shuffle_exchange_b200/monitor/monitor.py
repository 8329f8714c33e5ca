"""Experiment monitors (reference: deepspeed/monitor/{monitor,tensorboard,wandb,csv_monitor,comet}.py).
``MonitorMaster`` fans ``(tag, value, step)`` events out to every enabled backend on rank 0."""
from __future__ import annotations

import csv
import os
from abc import ABC, abstractmethod

from ..comm import comm as dist
from ..utils.logging import logger


class Monitor(ABC):
    def __init__(self, cfg):
        self.cfg = cfg
        self.enabled = cfg.enabled

    @abstractmethod
    def write_events(self, event_list):
        ...


def _rank0():
    return dist.get_rank() == 0


class TensorBoardMonitor(Monitor):
    def __init__(self, cfg):
        super().__init__(cfg)
        self.summary_writer = None
        if self.enabled and _rank0():
            try:
                from torch.utils.tensorboard import SummaryWriter
                base = cfg.output_path or os.path.join(os.path.expanduser("~"), "tensorboard")
                path = os.path.join(base, cfg.job_name)
                os.makedirs(path, exist_ok=True)
                self.summary_writer = SummaryWriter(log_dir=path)
            except Exception as e:
                logger.warning(f"tensorboard unavailable ({e}); disabling the monitor")
                self.enabled = False

    def write_events(self, event_list, flush=True):
        if self.summary_writer is None:
            return
        for ev in event_list:
            self.summary_writer.add_scalar(*ev)
        if flush:
            self.summary_writer.flush()

    def flush(self):
        if self.summary_writer is not None:
            self.summary_writer.flush()


class WandbMonitor(Monitor):
    def __init__(self, cfg):
        super().__init__(cfg)
        self._wandb = None
        if self.enabled and _rank0():
            try:
                import wandb
                wandb.init(project=cfg.project, group=cfg.group, entity=cfg.team)
                self._wandb = wandb
            except Exception as e:
                logger.warning(f"wandb unavailable ({e}); disabling the monitor")
                self.enabled = False

    def write_events(self, event_list):
        if self._wandb is None:
            return
        for label, value, step in event_list:
            self._wandb.log({label: value}, step=step)


class CometMonitor(Monitor):
    def __init__(self, cfg):
        super().__init__(cfg)
        self._exp = None
        if self.enabled and _rank0():
            try:
                import comet_ml
                self._exp = comet_ml.start(api_key=cfg.api_key, project=cfg.project, workspace=cfg.workspace,
                                           experiment_key=cfg.experiment_key, mode=cfg.mode, online=cfg.online)
                if cfg.experiment_name:
                    self._exp.set_name(cfg.experiment_name)
            except Exception as e:
                logger.warning(f"comet_ml unavailable ({e}); disabling the monitor")
                self.enabled = False
        self._interval = cfg.samples_log_interval

    def write_events(self, event_list):
        if self._exp is None:
            return
        for label, value, step in event_list:
            self._exp.log_metric(label, value, step=step)


class csvMonitor(Monitor):
    def __init__(self, cfg):
        super().__init__(cfg)
        self.filenames = {}
        self.log_dir = None
        if self.enabled and _rank0():
            base = cfg.output_path or os.path.join(os.path.expanduser("~"), "csv_monitor")
            self.log_dir = os.path.join(base, cfg.job_name)
            os.makedirs(self.log_dir, exist_ok=True)

    def write_events(self, event_list):
        if self.log_dir is None:
            return
        for label, value, step in event_list:
            header = label.split("/")[-1]
            fname = os.path.join(self.log_dir, label.replace("/", "_") + ".csv")
            new = fname not in self.filenames and not os.path.exists(fname)
            self.filenames[fname] = True
            with open(fname, "a+", newline="") as f:
                w = csv.writer(f)
                if new:
                    w.writerow(["step", header])
                w.writerow([step, value])


class MonitorMaster(Monitor):
    def __init__(self, monitor_config):
        self.cfg = monitor_config
        self.enabled = monitor_config.enabled
        self.tb_monitor = self.wandb_monitor = self.csv_monitor = self.comet_monitor = None
        if _rank0():
            if monitor_config.tensorboard.enabled:
                self.tb_monitor = TensorBoardMonitor(monitor_config.tensorboard)
            if monitor_config.wandb.enabled:
                self.wandb_monitor = WandbMonitor(monitor_config.wandb)
            if monitor_config.csv_monitor.enabled:
                self.csv_monitor = csvMonitor(monitor_config.csv_monitor)
            if monitor_config.comet.enabled:
                self.comet_monitor = CometMonitor(monitor_config.comet)

    def write_events(self, event_list):
        if not _rank0():
            return
        for m in (self.tb_monitor, self.wandb_monitor, self.csv_monitor, self.comet_monitor):
            if m is not None and m.enabled:
                m.write_events(event_list)
