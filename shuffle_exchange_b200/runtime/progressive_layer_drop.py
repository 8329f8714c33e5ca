"""Progressive layer dropping schedule (reference: deepspeed/runtime/progressive_layer_drop.py:10-40).

``theta(t) = (1 - theta_bar) * exp(-gamma * t) + theta_bar``; the engine passes ``get_state()`` into the module's
forward as kwargs each step (reference engine.py:2126-2127) and calls ``update_state`` at every optimizer step.
"""
import math

from ..utils.logging import log_dist


class ProgressiveLayerDrop:
    def __init__(self, theta: float = 0.5, gamma: float = 0.001):
        self.theta, self.gamma = float(theta), float(gamma)
        self.current_theta = 1.0
        log_dist(f"Enabled progressive layer dropping (theta = {self.theta})", ranks=[0])

    def get_theta(self) -> float:
        return self.current_theta

    def get_state(self) -> dict:
        return {"progressive_layer_drop": True, "pld_theta": self.get_theta()}

    def update_state(self, global_step: int) -> None:
        keep_floor = self.theta
        self.current_theta = (1.0 - keep_floor) * math.exp(-self.gamma * global_step) + keep_floor
