"""Turns techniques on when training reaches their ``schedule_offset`` (reference: ``deepspeed/compression/scheduler.py``)."""
from ..utils.logging import logger
from .compress import get_module_name
from .constants import *  # noqa: F401,F403
from .helper import recursive_getattr

_FLAG = {WEIGHT_QUANTIZATION: "weight_quantization_enabled", ACTIVATION_QUANTIZATION: "activation_quantization_enabled",
         SPARSE_PRUNING: "sparse_pruning_enabled", HEAD_PRUNING: "head_pruning_enabled", ROW_PRUNING: "row_pruning_enabled",
         CHANNEL_PRUNING: "channel_pruning_enabled"}


class compression_scheduler:
    def __init__(self, model, compression_config):
        self.model = model
        self.compression_config = compression_config
        self.make_init()
        self.training_steps = 0
        self.weight_quantization_enabled = False
        self.verbose = {k: False for k in _FLAG}

    def make_init(self):
        self.different_compression_methods = {}
        for method, content in self.compression_config.items():
            if method == LAYER_REDUCTION:
                continue
            shared = content[SHARED_PARAMETERS]
            entry = {TECHNIQUE_ENABLED: shared[TECHNIQUE_ENABLED], SHARED_PARAMETERS: shared, DIFFERENT_GROUPS: []}
            for gname, g in content[DIFFERENT_GROUPS].items():
                names = []
                scopes = g[DIFFERENT_GROUPS_MODULE_SCOPE]
                for kw in (scopes if isinstance(scopes, (list, tuple)) else [scopes]):
                    found, _ = get_module_name(gname, self.model, kw, set(), verbose=False)
                    names.extend(found)
                if names:
                    entry[DIFFERENT_GROUPS].append([gname, names, dict(g[DIFFERENT_GROUPS_PARAMETERS])])
            self.different_compression_methods[method] = entry

    def _check(self, method):
        entry = self.different_compression_methods[method]
        if not entry[TECHNIQUE_ENABLED]:
            return
        shared = entry[SHARED_PARAMETERS]
        if self.training_steps < shared[TECHNIQUE_SCHEDULE_OFFSET]:
            return
        if method == SPARSE_PRUNING and self.training_steps > shared.get(TECHNIQUE_SCHEDULE_OFFSET_END, 1 << 62) \
                and shared[SPARSE_PRUNING_METHOD] == SPARSE_PRUNING_METHOD_SNIP_MOMENTUM:
            return
        for _g, names, _p in entry[DIFFERENT_GROUPS]:
            for name in names:
                setattr(recursive_getattr(self.model, name), _FLAG[method], True)
        if not self.verbose[method]:
            logger.info(f"{method} is enabled at step {self.training_steps}")
            self.verbose[method] = True
            if method == WEIGHT_QUANTIZATION:
                self.weight_quantization_enabled = True

    def check_weight_quantization(self): self._check(WEIGHT_QUANTIZATION)
    def check_activation_quantization(self): self._check(ACTIVATION_QUANTIZATION)
    def check_sparse_pruning(self): self._check(SPARSE_PRUNING)
    def check_head_pruning(self): self._check(HEAD_PRUNING)
    def check_row_pruning(self): self._check(ROW_PRUNING)
    def check_channel_pruning(self): self._check(CHANNEL_PRUNING)

    def check_all_modules(self):
        for method in _FLAG:
            self._check(method)

    def step(self, step_zero_check=False):
        if not step_zero_check:
            self.training_steps += 1
        self.check_all_modules()
