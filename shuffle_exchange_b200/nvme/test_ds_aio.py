"""Single-run benchmark: ``python -m shuffle_exchange_b200.nvme.test_ds_aio --folder /mnt/nvme --io_size 1G --read``."""
from __future__ import annotations

import multiprocessing as mp

from .ds_aio_args import get_validated_args
from .io_engine import io_engine_multiprocessing


def ds_io_main(argv=None):
    print("Testing file IO performance")
    args = get_validated_args(argv)
    try:
        mp.set_start_method("spawn", force=True)
    except RuntimeError:
        pass
    return io_engine_multiprocessing(args, args.read)


if __name__ == "__main__":
    ds_io_main()
