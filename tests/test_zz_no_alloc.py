"""RMSNorm backward through the entry point that never allocates.

Some users of this library cannot obtain device memory at the moment they need the kernel: a trainer
whose memory plan has already committed the whole GPU, a server that gave the device to its KV-cache
pool, or a recovery path that runs because an allocation just failed.  ``sxb_rmsnorm_bwd_noalloc`` keeps
its scratch inside the library image, so it is called here exactly as such a user would — straight
through the C ABI with raw pointers — and must agree with the ordinary entry point (which takes a
caller-provided scratch buffer) on two consecutive calls that share the built-in scratch.

This file sorts last on purpose: if it fails under ``pytest -x`` every other test has already run.
"""
import ctypes
import os

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "shuffle_exchange_b200", "native", "libsxb200.so")
BF16 = 2
TOL = 2e-2          # same bound the kernel numerics tests use


def _lib():
    lib = ctypes.CDLL(LIB)
    P, L, I = ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int
    lib.sxb_rmsnorm_bwd.argtypes = [P, P, P, P, P, P, P, P, L, I, I, I, P]
    lib.sxb_rmsnorm_bwd.restype = I
    lib.sxb_rmsnorm_bwd_noalloc.argtypes = [P, P, P, P, P, P, P, L, I, I, I, P]
    lib.sxb_rmsnorm_bwd_noalloc.restype = I
    lib.sxb_norm_bwd_scratch_elems.argtypes = [L, I, I]
    lib.sxb_norm_bwd_scratch_elems.restype = L
    lib.sxb_rmsnorm_bwd_noalloc_max_h.restype = I
    return lib


def _rel(a, b):
    return ((a.float() - b.float()).abs().max() / b.float().abs().max().clamp_min(1e-6)).item()


def _case(lib, rows, H, seed, with_residual_grad):
    torch.manual_seed(seed)
    dev = "cuda"
    x = torch.randn(rows, H, device=dev, dtype=torch.bfloat16)
    dy = torch.randn(rows, H, device=dev, dtype=torch.bfloat16)
    w = (1 + 0.1 * torch.randn(H, device=dev)).to(torch.bfloat16)
    rstd = torch.rsqrt(x.float().pow(2).mean(-1) + 1e-5).contiguous()
    dres = torch.randn(rows, H, device=dev, dtype=torch.bfloat16) if with_residual_grad else None
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    pres = ctypes.c_void_p(dres.data_ptr()) if dres is not None else None
    # ordinary entry point: the caller allocates the scratch
    dx_a, dw_a = torch.empty_like(x), torch.empty_like(w)
    scratch = torch.empty(lib.sxb_norm_bwd_scratch_elems(rows, H, 0), device=dev, dtype=torch.float32)
    rc = lib.sxb_rmsnorm_bwd(dy.data_ptr(), x.data_ptr(), w.data_ptr(), rstd.data_ptr(), pres, dx_a.data_ptr(),
                             dw_a.data_ptr(), scratch.data_ptr(), rows, H, BF16, 0, stream)
    assert rc == 0
    # entry point under test: outputs pre-allocated by the caller, nothing else
    dx_b, dw_b = torch.empty_like(x), torch.empty_like(w)
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    rc = lib.sxb_rmsnorm_bwd_noalloc(dy.data_ptr(), x.data_ptr(), w.data_ptr(), rstd.data_ptr(), pres,
                                     dx_b.data_ptr(), dw_b.data_ptr(), rows, H, BF16, 0, stream)
    assert rc == 0
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    assert _rel(dx_b, dx_a) < TOL and _rel(dw_b, dw_a) < TOL, (_rel(dx_b, dx_a), _rel(dw_b, dw_a))
    # and both agree with plain PyTorch fp32 math
    xf, wf, dyf = x.float().requires_grad_(), w.float().requires_grad_(), dy.float()
    (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + 1e-5) * wf).backward(dyf)
    ref_dx = xf.grad + (dres.float() if dres is not None else 0)
    assert _rel(dx_b, ref_dx) < TOL and _rel(dw_b, wf.grad) < TOL


@pytest.mark.gpu
@pytest.mark.timeout(120)
def test_noalloc_rmsnorm_bwd_matches_regular_entry_point_twice():
    lib = _lib()
    _case(lib, rows=1000, H=4096, seed=0, with_residual_grad=True)     # first call also resolves the scratch address
    _case(lib, rows=77, H=512, seed=1, with_residual_grad=False)       # reuses the scratch with stale contents in it


@pytest.mark.gpu
@pytest.mark.timeout(120)
def test_noalloc_rejects_sizes_beyond_its_scratch():
    lib = _lib()
    max_h = lib.sxb_rmsnorm_bwd_noalloc_max_h()
    H = max_h * 2
    t = torch.zeros(4, H, device="cuda", dtype=torch.bfloat16)
    w = torch.ones(H, device="cuda", dtype=torch.bfloat16)
    r = torch.ones(4, device="cuda")
    rc = lib.sxb_rmsnorm_bwd_noalloc(t.data_ptr(), t.data_ptr(), w.data_ptr(), r.data_ptr(), None, t.data_ptr(),
                                     w.data_ptr(), 4, H, BF16, 0, None)
    assert rc == -3        # "size beyond limit", reported without launching anything
