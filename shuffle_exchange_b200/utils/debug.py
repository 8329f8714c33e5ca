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
