"""Build/version facts (reference: ``deepspeed/git_version_info.py:8-33``): version string, git hash/branch when the
package sits in a git checkout, and the table of native op families with their build status."""
import os
import subprocess

from .version import __version__ as version

_root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _git(*args):
    try:
        return subprocess.check_output(["git", "-C", _root, *args], stderr=subprocess.DEVNULL, text=True).strip()
    except Exception:
        return "unknown"


git_hash = _git("rev-parse", "--short", "HEAD")
git_branch = _git("rev-parse", "--abbrev-ref", "HEAD")

# op family -> C entry point that proves it is present in libsxb200.so / libsxb200_host.so
ALL_OPS = {
    "gemm_tcgen05": ("cuda", "sxb_gemm_bf16"), "flash_attention": ("cuda", "sxb_attn_fwd"),
    "fused_adam": ("cuda", "sxb_adam_flat"), "fused_lion": ("cuda", "sxb_lion_flat"), "fused_lamb": ("cuda", "sxb_lamb"),
    "fused_adagrad": ("cuda", "sxb_adagrad_flat"), "norms": ("cuda", "sxb_rmsnorm_fwd"),
    "quantizer": ("cuda", "sxb_quantize_int"), "fp_quantizer": ("cuda", "sxb_fp_quantize"),
    "packbits": ("cuda", "sxb_packbits"), "random_ltd": ("cuda", "sxb_token_move"),
    "moe_routing": ("cuda", "sxb_top_k_gating"), "ragged_inference": ("cuda", "sxb_paged_attention"),
    "symm_collectives": ("cuda", "sxb_symm_rs_adam"),
    "cpu_adam": ("host", "sxbh_adam_update"), "cpu_adagrad": ("host", "sxbh_adagrad_update"),
    "cpu_lion": ("host", "sxbh_lion_update"), "async_io": ("host", "sxbh_aio_pread"),
}


def _probe():
    import ctypes
    here = os.path.join(os.path.dirname(os.path.abspath(__file__)), "native")
    libs = {}
    for kind, name in (("cuda", "libsxb200.so"), ("host", "libsxb200_host.so")):
        path = os.path.join(here, name)
        try:
            libs[kind] = ctypes.CDLL(path) if os.path.exists(path) else None
        except OSError:
            libs[kind] = None
    out = {}
    for op, (kind, sym) in ALL_OPS.items():
        lib = libs.get(kind)
        out[op] = bool(lib is not None and hasattr(lib, sym))
    return out


installed_ops = compatible_ops = None


def op_status():
    global installed_ops, compatible_ops
    if installed_ops is None:
        installed_ops = _probe()
        compatible_ops = dict.fromkeys(ALL_OPS, True)
    return installed_ops
