"""``"timers"`` section (reference: ``utils/config.py``): ``{"timers": {"throughput": {"enabled": true,
"synchronized": true}}}``."""
from ..runtime.config import TimersConfig

TIMERS = "timers"
TIMERS_THROUGHPUT = "throughput"

DeepSpeedThroughputTimerConfig = TimersConfig.Throughput


def get_timers_config(param_dict):
    d = param_dict.get(TIMERS, {}).get(TIMERS_THROUGHPUT, {}) if param_dict else {}
    return DeepSpeedThroughputTimerConfig(**d)
