"""Is an optional monitor back end importable?  (reference: ``monitor/utils.py``.)  Each check re-raises the
``ImportError`` after printing what to install, so enabling a monitor in the config without its package fails at
start-up rather than at the first write."""
import importlib

_HINTS = {
    "tensorboard": "If you want to use tensorboard logging, please `pip install tensorboard`",
    "wandb": "If you want to use wandb logging, please `pip install wandb` and follow the instructions at "
             "https://docs.wandb.ai/quickstart",
    "comet_ml": 'If you want to use comet logging, please `pip install "comet_ml>=3.41.0"`',
}
_COMET_MIN = (3, 41)


def _require(package: str, also=()):
    try:
        mod = importlib.import_module(package)
        for extra in also:
            importlib.import_module(extra)
        return mod
    except ImportError:
        print(_HINTS[package])
        raise


def check_tb_availability():
    _require("tensorboard", also=("torch.utils.tensorboard",))


def check_wandb_availability():
    _require("wandb")


def check_comet_availability():
    comet = _require("comet_ml")
    found = tuple(int(x) for x in comet.__version__.split(".")[:2])
    if found < _COMET_MIN:
        print(_HINTS["comet_ml"])
        raise ImportError(f"`comet_ml` must have at least version {_COMET_MIN[0]}.{_COMET_MIN[1]}.0 (found {comet.__version__})")
