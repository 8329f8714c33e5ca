"""Name maps for debugging ZeRO state (reference: deepspeed/utils/debug.py)."""
module_names = {}
param_names = {}


def debug_extract_module_and_param_names(model):
    global module_names, param_names
    module_names = {m: n for n, m in model.named_modules()}
    param_names = {p: n for n, p in model.named_parameters()}


def debug_module2name(module):
    return module_names.get(module, "unknown")


def debug_param2name(param):
    return param_names.get(param, "unknown")


def debug_param2name_id_shape(param):
    return f"name={debug_param2name(param)} id={getattr(param, 'ds_id', None)} shape={tuple(param.shape)}"


def debug_param2name_id_shape_status(param):
    return f"{debug_param2name_id_shape(param)} status={getattr(param, 'ds_status', None)}"


def printflock(*msgs):
    import fcntl
    with open(__file__, "r") as fh:
        fcntl.flock(fh, fcntl.LOCK_EX)
        try:
            print(*msgs)
        finally:
            fcntl.flock(fh, fcntl.LOCK_UN)


def debug_clear_module_and_param_names():
    global module_names, param_names
    module_names, param_names = {}, {}


def debug_module2name_id(module):
    return f"name={debug_module2name(module)} id={getattr(module, 'id', None)}"


def debug_module2name_class(module):
    return f"name={debug_module2name(module)} {module.__class__.__name__}"


def debug_param2name_id(param):
    return f"name={debug_param2name(param)} id={getattr(param, 'ds_id', None)}"


def debug_param2name_id_shape_device(param):
    return f"{debug_param2name_id_shape(param)} device={param.device}"


def debug_param2name_id_numel(param):
    return f"name={debug_param2name(param)} id={getattr(param, 'ds_id', None)} numel={param.numel()}"


def _rank() -> int:
    try:
        import torch.distributed as td
        return td.get_rank() if td.is_initialized() else 0
    except Exception:
        return 0


def log_rank_file(rank, *msgs):
    """Append ``msgs`` to ``log_rank_<rank>.txt`` when called on ``rank`` (rank -1: every rank, its own file)."""
    me = _rank()
    if rank != -1 and rank != me:
        return
    with open(f"log_rank_{me}.txt", "a") as fh:
        for m in msgs:
            fh.write(f"{m}\n")
            fh.flush()


def print_rank(rank, *msgs, force=False):
    """Print on one rank (``-1``: all ranks), prefixed with the rank."""
    me = _rank()
    if force or rank == -1 or rank == me:
        print(f"[Rank {me}]", *msgs, flush=True)


def print_rank0(*msgs, force=False):
    print_rank(0, *msgs, force=force)


def print_backward_tensors(tensor):
    """Walk the autograd graph behind ``tensor`` and print the leaf tensors each backward node touches."""
    seen = set()

    def walk(fn, depth=0):
        if fn is None or fn in seen:
            return
        seen.add(fn)
        for nxt, _ in getattr(fn, "next_functions", ()):
            if nxt is None:
                continue
            var = getattr(nxt, "variable", None)
            if var is not None:
                print(f"{'  ' * depth}{type(fn).__name__} -> leaf {tuple(var.shape)} "
                      f"{getattr(var, 'ds_id', debug_param2name(var))}")
            walk(nxt, depth + 1)

    if getattr(tensor, "grad_fn", None) is not None:
        walk(tensor.grad_fn)
