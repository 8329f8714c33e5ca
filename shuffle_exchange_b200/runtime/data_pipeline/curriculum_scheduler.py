"""Difficulty schedule for curriculum learning.

Reference: deepspeed/runtime/data_pipeline/curriculum_scheduler.py:11-158.  Four schedule types:
``fixed_discrete`` (step table), ``fixed_linear`` / ``fixed_root`` (difficulty grows as
``(t/T)^(1/degree)`` between min and max, floored to a multiple of ``difficulty_step``) and ``custom``
(user callback).  State lives in a plain dict so the engine can checkpoint it.
"""
from __future__ import annotations

import math

from ...utils.logging import logger
from .constants import *  # noqa: F401,F403

_REQUIRED = {
    CURRICULUM_LEARNING_SCHEDULE_FIXED_DISCRETE: (CURRICULUM_LEARNING_SCHEDULE_DIFFICULTY,
                                                  CURRICULUM_LEARNING_SCHEDULE_MAX_STEP),
    CURRICULUM_LEARNING_SCHEDULE_FIXED_ROOT: (CURRICULUM_LEARNING_SCHEDULE_TOTAL_STEP,
                                              CURRICULUM_LEARNING_SCHEDULE_DIFFICULTY_STEP,
                                              CURRICULUM_LEARNING_SCHEDULE_ROOT_DEGREE),
    CURRICULUM_LEARNING_SCHEDULE_FIXED_LINEAR: (CURRICULUM_LEARNING_SCHEDULE_TOTAL_STEP,
                                                CURRICULUM_LEARNING_SCHEDULE_DIFFICULTY_STEP),
    CURRICULUM_LEARNING_SCHEDULE_CUSTOM: (),
}


class CurriculumScheduler:
    def __init__(self, config: dict):
        for key in (CURRICULUM_LEARNING_MIN_DIFFICULTY, CURRICULUM_LEARNING_MAX_DIFFICULTY,
                    CURRICULUM_LEARNING_SCHEDULE_TYPE):
            assert key in config, f"Curriculum learning requires the config '{key}'"
        kind = config[CURRICULUM_LEARNING_SCHEDULE_TYPE]
        if kind not in _REQUIRED:
            raise RuntimeError("Unsupported curriculum schedule type")
        self.state = {
            CURRICULUM_LEARNING_MIN_DIFFICULTY: config[CURRICULUM_LEARNING_MIN_DIFFICULTY],
            CURRICULUM_LEARNING_MAX_DIFFICULTY: config[CURRICULUM_LEARNING_MAX_DIFFICULTY],
            CURRICULUM_LEARNING_CURRENT_DIFFICULTY: config[CURRICULUM_LEARNING_MIN_DIFFICULTY],
            CURRICULUM_LEARNING_SCHEDULE_TYPE: kind,
        }
        self.first_step = True
        self.custom_get_difficulty = None
        if kind != CURRICULUM_LEARNING_SCHEDULE_CUSTOM:
            sc = config.get(CURRICULUM_LEARNING_SCHEDULE_CONFIG, {})
            for key in _REQUIRED[kind]:
                assert key in sc, f"Curriculum learning with {kind} schedule requires the schedule_config '{key}'"
            if kind == CURRICULUM_LEARNING_SCHEDULE_FIXED_DISCRETE:
                diffs, steps = sc[CURRICULUM_LEARNING_SCHEDULE_DIFFICULTY], sc[CURRICULUM_LEARNING_SCHEDULE_MAX_STEP]
                assert len(steps) > 0 and len(diffs) == len(steps) + 1, \
                    "fixed_discrete needs one more difficulty than max_step entries"
            elif sc[CURRICULUM_LEARNING_SCHEDULE_DIFFICULTY_STEP] % 8 != 0:
                logger.warning("difficulty_step should be a multiple of 8 when the metric is sequence length so that "
                               "GEMM shapes stay tensor-core friendly; ignore if unrelated to your metric.")
            self.state[CURRICULUM_LEARNING_SCHEDULE_CONFIG] = sc

    # -- state ---------------------------------------------------------------------------------------
    def get_current_difficulty(self):
        return self.state[CURRICULUM_LEARNING_CURRENT_DIFFICULTY]

    def set_current_difficulty(self, difficulty):
        self.state[CURRICULUM_LEARNING_CURRENT_DIFFICULTY] = difficulty

    def set_custom_get_difficulty(self, schedule_function):
        self.custom_get_difficulty = schedule_function

    def get_state(self):
        return self.state

    def set_state(self, state):
        self.state = state

    # -- schedules -------------------------------------------------------------------------------------
    def _discrete(self, step):
        sc = self.state[CURRICULUM_LEARNING_SCHEDULE_CONFIG]
        for limit, diff in zip(sc[CURRICULUM_LEARNING_SCHEDULE_MAX_STEP], sc[CURRICULUM_LEARNING_SCHEDULE_DIFFICULTY]):
            if step <= limit:
                return diff
        return sc[CURRICULUM_LEARNING_SCHEDULE_DIFFICULTY][-1]

    def _root(self, step, degree=None):
        sc = self.state[CURRICULUM_LEARNING_SCHEDULE_CONFIG]
        degree = sc[CURRICULUM_LEARNING_SCHEDULE_ROOT_DEGREE] if degree is None else degree
        lo, hi = self.state[CURRICULUM_LEARNING_MIN_DIFFICULTY], self.state[CURRICULUM_LEARNING_MAX_DIFFICULTY]
        frac = (float(step) / sc[CURRICULUM_LEARNING_SCHEDULE_TOTAL_STEP]) ** (1.0 / degree)
        d = math.floor(frac * (hi - lo) + lo)
        d -= d % sc[CURRICULUM_LEARNING_SCHEDULE_DIFFICULTY_STEP]
        return min(d, hi)

    def get_difficulty(self, global_steps):
        kind = self.state[CURRICULUM_LEARNING_SCHEDULE_TYPE]
        if kind == CURRICULUM_LEARNING_SCHEDULE_FIXED_DISCRETE:
            return self._discrete(global_steps)
        if kind == CURRICULUM_LEARNING_SCHEDULE_FIXED_LINEAR:
            return self._root(global_steps, 1)
        if kind == CURRICULUM_LEARNING_SCHEDULE_FIXED_ROOT:
            return self._root(global_steps)
        if kind == CURRICULUM_LEARNING_SCHEDULE_CUSTOM:
            return self.custom_get_difficulty(global_steps)
        raise RuntimeError("Unsupported curriculum schedule type")

    def update_difficulty(self, global_steps):
        if self.state[CURRICULUM_LEARNING_CURRENT_DIFFICULTY] < self.state[CURRICULUM_LEARNING_MAX_DIFFICULTY]:
            self.state[CURRICULUM_LEARNING_CURRENT_DIFFICULTY] = self.get_difficulty(global_steps)
        return self.state[CURRICULUM_LEARNING_CURRENT_DIFFICULTY]
