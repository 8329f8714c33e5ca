"""Availability checks for the optional monitor back ends (reference: ``monitor/utils.py``): each raises an
``ImportError`` with the install hint when the package is missing."""


def check_tb_availability():
    try:
        import tensorboard  # noqa: F401
        from torch.utils.tensorboard import SummaryWriter  # noqa: F401
    except ImportError:
        print('If you want to use tensorboard logging, please `pip install tensorboard`')
        raise


def check_wandb_availability():
    try:
        import wandb  # noqa: F401
    except ImportError:
        print('If you want to use wandb logging, please `pip install wandb` and follow the instructions at '
              'https://docs.wandb.ai/quickstart')
        raise


def check_comet_availability():
    try:
        import comet_ml
        version = tuple(int(x) for x in comet_ml.__version__.split(".")[:2])
        if version < (3, 41):
            raise ImportError("`comet_ml` must have at least version 3.41.0")
    except ImportError:
        print('If you want to use comet logging, please `pip install "comet_ml>=3.41.0"`')
        raise
