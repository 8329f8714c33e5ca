"""Autograd-aware front ends of the sm_100a kernels.

Each op has two implementations behind one function:

* CUDA tensors  -> the hand-written kernel in ``libsxb200.so`` (mandatory; errors are raised,
  never swallowed).
* CPU tensors   -> a plain PyTorch composite with identical semantics (used by the CPU/gloo test
  suite and as the numerical oracle of the GPU tests).

Weight gradients support the "main grad" protocol used by the ZeRO engines: if a parameter
carries ``_sxb_main_grad`` (a view into the unit's flat gradient buffer) the wgrad GEMM writes
(or accumulates) straight into it and ``_sxb_grad_ready(param)`` is invoked, so no per-parameter
``.grad`` tensor is ever materialised and the reduce-scatter can start tile-exact.
"""
from __future__ import annotations

import math
from typing import Optional, Tuple

import torch
import torch.nn.functional as F

from ..native import lib as _n

ACT_IDENTITY, ACT_RELU, ACT_GELU, ACT_SILU = 0, 1, 2, 3
_ACT_CODES = {"identity": 0, "none": 0, "relu": 1, "gelu": 2, "silu": 3, "swish": 3}


def _act_code(act) -> int:
    return act if isinstance(act, int) else _ACT_CODES[str(act).lower()]


def _torch_act(x, code):
    if code == ACT_RELU:
        return F.relu(x)
    if code == ACT_GELU:
        return F.gelu(x, approximate="tanh")
    if code == ACT_SILU:
        return F.silu(x)
    return x


# ----------------------------------------------------------------------------- GEMM
GEMM_NT, GEMM_NN, GEMM_TN = 0, 1, 2

# Set by profiling.flops_profiler while a profile is running: callable(kind: str, flops: float).  The native ops
# are invisible to torch's dispatcher, so they report their own work.
flop_sink = None


def gemm(a: torch.Tensor, b: torch.Tensor, layout: int = GEMM_NT, out: Optional[torch.Tensor] = None,
         accumulate: bool = False, alpha: float = 1.0, out_dtype: Optional[torch.dtype] = None,
         max_ctas: int = 0) -> torch.Tensor:
    """``out (+)= alpha * op(a) op(b)`` on the tcgen05 GEMM (bf16 in, bf16/fp32 out).

    NT: a[M,K] b[N,K] ; NN: a[M,K] b[K,N] ; TN: a[K,M] b[K,N].
    """
    if layout == GEMM_NT:
        M, K = a.shape
        N = b.shape[0]
    elif layout == GEMM_NN:
        M, K = a.shape
        N = b.shape[1]
    else:
        K, M = a.shape
        N = b.shape[1]
    if out is None:
        out = torch.empty((M, N), dtype=out_dtype or a.dtype, device=a.device)
        accumulate = False
    if flop_sink is not None:
        flop_sink("gemm", 2.0 * M * N * K)
    if not a.is_cuda:
        fa, fb = a.float(), b.float()
        r = fa @ fb.t() if layout == GEMM_NT else (fa @ fb if layout == GEMM_NN else fa.t() @ fb)
        r = alpha * r
        if accumulate:
            r = r + out.float()
        out.copy_(r.to(out.dtype))
        return out
    if a.dtype != torch.bfloat16 or b.dtype != torch.bfloat16:
        raise _n.NativeError("sxb gemm expects bf16 operands")
    if a.stride(-1) != 1 or b.stride(-1) != 1 or out.stride(-1) != 1:
        raise _n.NativeError("sxb gemm expects unit inner stride")
    _n.call("sxb_gemm_bf16", a.data_ptr(), b.data_ptr(), out.data_ptr(), M, N, K, layout,
            a.stride(0), b.stride(0), out.stride(0), _n.dtype_code(out.dtype), int(accumulate),
            float(alpha), int(max_ctas), _n.stream_ptr())
    return out


# ------------------------------------------------------------------------- matmul front ends
def _tc_ok(*ts) -> bool:
    return all(t.is_cuda and t.dtype == torch.bfloat16 for t in ts)


def matmul_nt(x: torch.Tensor, w: torch.Tensor) -> torch.Tensor:
    """``x[..., K] @ w[N, K]^T`` — the tcgen05 GEMM for bf16 CUDA operands; PyTorch for CPU tensors (oracle / gloo
    tests) and for dtypes the kernel does not take."""
    if _tc_ok(x, w) and x.shape[-1] % 8 == 0:
        x2 = x.reshape(-1, x.shape[-1])
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        return gemm(x2, w if w.stride(-1) == 1 else w.contiguous(), GEMM_NT).view(*x.shape[:-1], w.shape[0])
    return torch.matmul(x, w.t())


def matmul_nn(x: torch.Tensor, w: torch.Tensor) -> torch.Tensor:
    """``x[..., K] @ w[K, N]`` (dgrad form)."""
    if _tc_ok(x, w) and x.shape[-1] % 8 == 0 and w.shape[-1] % 8 == 0:
        x2 = x.reshape(-1, x.shape[-1])
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        return gemm(x2, w if w.stride(-1) == 1 else w.contiguous(), GEMM_NN).view(*x.shape[:-1], w.shape[1])
    return torch.matmul(x, w)


def matmul_tn(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """``a[R, M]^T @ b[R, N]`` (wgrad form); leading dimensions of both are flattened into ``R``."""
    a2, b2 = a.reshape(-1, a.shape[-1]), b.reshape(-1, b.shape[-1])
    if _tc_ok(a2, b2) and a2.shape[1] % 8 == 0 and b2.shape[1] % 8 == 0:
        return gemm(a2 if a2.is_contiguous() else a2.contiguous(), b2 if b2.is_contiguous() else b2.contiguous(),
                    GEMM_TN)
    return torch.matmul(a2.t(), b2)


def accumulate_tn(acc: torch.Tensor, a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """``acc[M, N] += a[R, M]^T @ b[R, N]`` with ``acc`` in fp32 — chunked weight gradients accumulate in the GEMM
    epilogue instead of a separate add pass."""
    a2, b2 = a.reshape(-1, a.shape[-1]), b.reshape(-1, b.shape[-1])
    if _tc_ok(a2, b2) and a2.shape[1] % 8 == 0 and b2.shape[1] % 8 == 0 and acc.is_contiguous():
        gemm(a2 if a2.is_contiguous() else a2.contiguous(), b2 if b2.is_contiguous() else b2.contiguous(), GEMM_TN,
             out=acc, accumulate=True)
        return acc
    return acc.add_(torch.matmul(a2.t(), b2).to(acc.dtype))


# ---------------------------------------------------------------------- gradient-use accounting
# A parameter used k times in one forward (tied embedding / LM head, a shared projection) delivers k
# contributions through the main-grad protocol, one per use.  The sharded optimizers must not start a
# unit's reduction before the LAST one landed, so every protocol op notes its use of the parameter in
# forward; the optimizer compares deliveries against that count (runtime/zero/sharded_optimizer.py:
# ``GradUseCounter``).  Uses are counted only while the window is open: it closes when backward starts
# (recomputation inside backward replays forwards that must not count) and a new one opens after it.
_use_window = {"gen": 0, "open": True}


def close_use_window():
    _use_window["open"] = False


def open_use_window():
    _use_window["gen"] += 1
    _use_window["open"] = True


def expected_uses(param) -> int:
    """How many main-grad deliveries this backward will make for ``param`` (>= 1)."""
    if getattr(param, "_sxb_use_gen", -1) != _use_window["gen"]:
        return 1
    return max(1, getattr(param, "_sxb_uses", 1))


def _note_use(param):
    if not _use_window["open"] or getattr(param, "_sxb_grad_ready", None) is None:
        return
    if getattr(param, "_sxb_use_gen", -1) != _use_window["gen"]:
        param._sxb_use_gen = _use_window["gen"]
        param._sxb_uses = 0
    param._sxb_uses += 1


def _deliver_wgrad(weight, dy2, x2):
    """dW = dy^T x, either returned or written into the parameter's main-grad view."""
    mg = getattr(weight, "_sxb_main_grad", None)
    if mg is not None:
        first = not getattr(weight, "_sxb_grad_written", False)
        gemm(dy2, x2, GEMM_TN, out=mg, accumulate=not first)
        weight._sxb_grad_written = True
        hook = getattr(weight, "_sxb_grad_ready", None)
        if hook is not None:
            hook(weight)
        return None
    return gemm(dy2, x2, GEMM_TN, out_dtype=weight.dtype)


class _Linear(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias):
        x2 = x.reshape(-1, x.shape[-1])
        # allocate with the final shape so the returned tensor is not a view (in-place consumers
        # such as RoPE may then modify it)
        y = torch.empty((*x.shape[:-1], weight.shape[0]), dtype=x.dtype, device=x.device)
        gate = getattr(weight, "_sxb_gate", None)
        if gate is not None:
            # K2: the weight is still being gathered; the kernel waits for the tiles it needs (ops/gemm_ext.py)
            from .gemm_ext import gemm_gated
            gemm_gated(x2, weight, GEMM_NT, gate, out=y.view(-1, weight.shape[0]))
        else:
            gemm(x2, weight, GEMM_NT, out=y.view(-1, weight.shape[0]))
        if bias is not None:
            y.add_(bias)
        _note_use(weight)
        ctx.save_for_backward(x2, weight)
        ctx.has_bias = bias is not None
        ctx.xshape = x.shape
        return y

    @staticmethod
    def backward(ctx, dy):
        x2, weight = ctx.saved_tensors
        dy2 = dy.reshape(-1, dy.shape[-1])
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            gate = getattr(weight, "_sxb_gate", None)
            if gate is not None:
                from .gemm_ext import gemm_gated
                dx = gemm_gated(dy2, weight, GEMM_NN, gate).view(ctx.xshape)
            else:
                dx = gemm(dy2, weight, GEMM_NN).view(ctx.xshape)
        if ctx.needs_input_grad[1]:
            dw = _deliver_wgrad(weight, dy2, x2)
        if ctx.has_bias and ctx.needs_input_grad[2]:
            db = dy2.sum(0)
        return dx, dw, db


_warned = set()


def _warn_once(key: str, message: str) -> None:
    if key not in _warned:
        _warned.add(key)
        import warnings
        warnings.warn(message, stacklevel=3)


def linear(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``x @ weight.T + bias`` — tcgen05 GEMM for bf16 CUDA tensors (fwd, dgrad and wgrad)."""
    if x.is_cuda and x.dtype == torch.bfloat16 and weight.dtype == torch.bfloat16 \
            and x.shape[-1] % 8 == 0 and weight.shape[0] % 8 == 0:
        if not x.is_contiguous():
            x = x.contiguous()
        return _Linear.apply(x, weight, bias)
    if x.is_cuda:
        # fp16 / fp32 operands or rows that are not 16-byte multiples: the tcgen05 GEMM takes bf16 only, the library
        # GEMM serves the rest (the reference's fp16 + loss-scaling mode keeps working; said once, not silently)
        _warn_once("linear", f"sxb linear: {x.dtype} {tuple(x.shape)} x {tuple(weight.shape)} runs on the library GEMM "
                             "(the tcgen05 path takes bf16 operands with 8-element-aligned sizes)")
    return F.linear(x, weight, bias)


# ----------------------------------------------------------------------------- norms
_scratch_cache = {}


def _scratch(device, nelem: int) -> torch.Tensor:
    key = (device.type, device.index)
    buf = _scratch_cache.get(key)
    if buf is None or buf.numel() < nelem:
        buf = torch.empty(max(nelem, 1 << 20), dtype=torch.float32, device=device)
        _scratch_cache[key] = buf
    return buf


def rms_norm_ref(x, weight, eps, residual=None):
    h = x if residual is None else (x.float() + residual.float()).to(x.dtype)
    hf = h.float()
    y = hf * torch.rsqrt(hf.pow(2).mean(-1, keepdim=True) + eps) * weight.float()
    return y.to(x.dtype), h


class _RMSNorm(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, eps, residual):
        H = x.shape[-1]
        x2 = x.reshape(-1, H)
        rows = x2.shape[0]
        y = torch.empty_like(x2)
        rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
        if residual is not None:
            r2 = residual.reshape(-1, H)
            h = torch.empty_like(x2)
        else:
            r2, h = None, None
        _n.call("sxb_rmsnorm_fwd", x2.data_ptr(), _n.ptr(r2), weight.data_ptr(), y.data_ptr(),
                _n.ptr(h), rstd.data_ptr(), rows, H, float(eps), _n.dtype_code(x.dtype), _n.stream_ptr())
        src = h if h is not None else x2
        _note_use(weight)
        ctx.save_for_backward(src, weight, rstd)
        ctx.has_res = residual is not None
        ctx.shape = x.shape
        ctx.set_materialize_grads(False)
        if residual is not None:
            return y.view(x.shape), h.view(x.shape)
        return y.view(x.shape), x.new_empty(0)

    @staticmethod
    def backward(ctx, dy, dh):
        src, weight, rstd = ctx.saved_tensors
        H = src.shape[-1]
        rows = src.shape[0]
        if dy is None:
            dy = torch.zeros(ctx.shape, dtype=src.dtype, device=src.device)
        dy2 = dy.reshape(-1, H)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dres = None
        if dh is not None and ctx.has_res:
            dres = dh.reshape(-1, H)
            if not dres.is_contiguous():
                dres = dres.contiguous()
        dx = torch.empty_like(src)
        mg = getattr(weight, "_sxb_main_grad", None)
        first = not getattr(weight, "_sxb_grad_written", False)
        dw = mg if mg is not None else torch.empty_like(weight)
        scratch = _scratch(src.device, _n.call("sxb_norm_bwd_scratch_elems", rows, H, 0))
        _n.call("sxb_rmsnorm_bwd", dy2.data_ptr(), src.data_ptr(), weight.data_ptr(), rstd.data_ptr(),
                _n.ptr(dres), dx.data_ptr(), dw.data_ptr(), scratch.data_ptr(), rows, H,
                _n.dtype_code(src.dtype), int(mg is not None and not first), _n.stream_ptr())
        if mg is not None:
            weight._sxb_grad_written = True
            hook = getattr(weight, "_sxb_grad_ready", None)
            if hook is not None:
                hook(weight)
            dw = None
        dxv = dx.view(ctx.shape)
        return dxv, dw, None, (dxv if ctx.has_res else None)


def rms_norm(x: torch.Tensor, weight: torch.Tensor, eps: float = 1e-6,
             residual: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """Returns ``(norm(x + residual) * weight, x + residual)`` (pre-norm block helper)."""
    if x.is_cuda:
        if not x.is_contiguous():
            x = x.contiguous()
        if residual is not None and not residual.is_contiguous():
            residual = residual.contiguous()
        y, h = _RMSNorm.apply(x, weight, eps, residual)
        return y, (h if residual is not None else x)
    return rms_norm_ref(x, weight, eps, residual)


class _LayerNorm(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, eps, residual):
        H = x.shape[-1]
        x2 = x.reshape(-1, H)
        rows = x2.shape[0]
        y = torch.empty_like(x2)
        mean = torch.empty(rows, dtype=torch.float32, device=x.device)
        rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
        r2 = residual.reshape(-1, H) if residual is not None else None
        h = torch.empty_like(x2) if residual is not None else None
        _n.call("sxb_layernorm_fwd", x2.data_ptr(), _n.ptr(r2), weight.data_ptr(), _n.ptr(bias),
                y.data_ptr(), _n.ptr(h), mean.data_ptr(), rstd.data_ptr(), rows, H, float(eps),
                _n.dtype_code(x.dtype), _n.stream_ptr())
        src = h if h is not None else x2
        ctx.save_for_backward(src, weight, mean, rstd)
        ctx.has_res = residual is not None
        ctx.has_bias = bias is not None
        ctx.shape = x.shape
        ctx.set_materialize_grads(False)
        return y.view(x.shape), (h.view(x.shape) if h is not None else x.new_empty(0))

    @staticmethod
    def backward(ctx, dy, dh):
        src, weight, mean, rstd = ctx.saved_tensors
        H = src.shape[-1]
        rows = src.shape[0]
        if dy is None:
            dy = torch.zeros(ctx.shape, dtype=src.dtype, device=src.device)
        dy2 = dy.reshape(-1, H).contiguous()
        dres = dh.reshape(-1, H).contiguous() if (dh is not None and ctx.has_res) else None
        dx = torch.empty_like(src)
        dw = torch.empty_like(weight)
        db = torch.empty_like(weight)
        scratch = _scratch(src.device, _n.call("sxb_norm_bwd_scratch_elems", rows, H, 1))
        _n.call("sxb_layernorm_bwd", dy2.data_ptr(), src.data_ptr(), weight.data_ptr(), mean.data_ptr(),
                rstd.data_ptr(), _n.ptr(dres), dx.data_ptr(), dw.data_ptr(), db.data_ptr(),
                scratch.data_ptr(), rows, H, _n.dtype_code(src.dtype), 0, _n.stream_ptr())
        dxv = dx.view(ctx.shape)
        return dxv, dw, (db if ctx.has_bias else None), None, (dxv if ctx.has_res else None)


def layer_norm(x, weight, bias=None, eps: float = 1e-5, residual=None):
    """Returns ``(LN(x + residual), x + residual)``."""
    if x.is_cuda:
        y, h = _LayerNorm.apply(x.contiguous(), weight, bias, eps,
                                residual.contiguous() if residual is not None else None)
        return y, (h if residual is not None else x)
    h = x if residual is None else x + residual
    return F.layer_norm(h.float(), (h.shape[-1],), weight.float(),
                        None if bias is None else bias.float(), eps).to(x.dtype), h


# ----------------------------------------------------------------------------- RoPE
_rope_cache = {}


def rope_tables(max_pos: int, head_dim: int, theta: float, device) -> Tuple[torch.Tensor, torch.Tensor]:
    key = (max_pos, head_dim, float(theta), str(device))
    t = _rope_cache.get(key)
    if t is None:
        inv = 1.0 / (theta ** (torch.arange(0, head_dim, 2, dtype=torch.float64) / head_dim))
        ang = torch.outer(torch.arange(max_pos, dtype=torch.float64), inv)
        t = (ang.cos().float().to(device).contiguous(), ang.sin().float().to(device).contiguous())
        _rope_cache[key] = t
    return t


def rope_ref(x, positions, cos, sin, n_heads, head_dim, inverse=False):
    """x: [T, >= n_heads*head_dim]; rotate-half convention on the first n_heads heads."""
    T = x.shape[0]
    xr = x[:, : n_heads * head_dim].reshape(T, n_heads, head_dim).float()
    c = cos[positions].unsqueeze(1)
    s = sin[positions].unsqueeze(1)
    if inverse:
        s = -s
    a, b = xr[..., : head_dim // 2], xr[..., head_dim // 2:]
    out = torch.cat([a * c - b * s, b * c + a * s], -1).reshape(T, n_heads * head_dim).to(x.dtype)
    res = x.clone()
    res[:, : n_heads * head_dim] = out
    return res


class _Rope(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, positions, cos, sin, n_heads, head_dim):
        # x: [T, C] contiguous; rotates in place on a clone-free basis (x is marked dirty)
        ctx.mark_dirty(x)
        C = x.shape[-1]
        _n.call("sxb_rope", x.data_ptr(), _n.ptr(positions), cos.data_ptr(), sin.data_ptr(), x.numel() // C,
                n_heads, head_dim, C, 0, _n.dtype_code(x.dtype), _n.stream_ptr())
        ctx.save_for_backward(positions, cos, sin)
        ctx.nh, ctx.hd = n_heads, head_dim
        return x

    @staticmethod
    def backward(ctx, dx):
        positions, cos, sin = ctx.saved_tensors
        dx = dx.clone(memory_format=torch.contiguous_format)  # never rotate autograd's buffer in place
        C = dx.shape[-1]
        _n.call("sxb_rope", dx.data_ptr(), _n.ptr(positions), cos.data_ptr(), sin.data_ptr(), dx.numel() // C,
                ctx.nh, ctx.hd, C, 1, _n.dtype_code(dx.dtype), _n.stream_ptr())
        return dx, None, None, None, None, None


def rope(x: torch.Tensor, positions: Optional[torch.Tensor], cos: torch.Tensor, sin: torch.Tensor,
         n_heads: int, head_dim: int) -> torch.Tensor:
    """Rotate the first ``n_heads`` heads of every row of ``x`` ([..., C], contiguous); in place on CUDA."""
    if x.is_cuda:
        if not x.is_contiguous():
            x = x.contiguous()
        return _Rope.apply(x, positions, cos, sin, n_heads, head_dim)
    shape = x.shape
    x2 = x.reshape(-1, shape[-1])
    pos = positions if positions is not None else torch.arange(x2.shape[0])
    return rope_ref(x2, pos, cos, sin, n_heads, head_dim).view(shape)


# ------------------------------------------------------------------- gated activations
class _Gated(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, act):
        F_ = x.shape[-1] // 2
        x2 = x.reshape(-1, 2 * F_)
        out = torch.empty((x2.shape[0], F_), dtype=x.dtype, device=x.device)
        _n.call("sxb_gated_act_fwd", x2.data_ptr(), None, out.data_ptr(), x2.shape[0], F_, act,
                _n.dtype_code(x.dtype), _n.stream_ptr())
        ctx.save_for_backward(x2)
        ctx.act = act
        ctx.shape = x.shape
        return out.view(*x.shape[:-1], F_)

    @staticmethod
    def backward(ctx, dy):
        (x2,) = ctx.saved_tensors
        F_ = x2.shape[-1] // 2
        dy2 = dy.reshape(-1, F_)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dx = torch.empty_like(x2)
        _n.call("sxb_gated_act_bwd", dy2.data_ptr(), x2.data_ptr(), dx.data_ptr(), x2.shape[0], F_,
                ctx.act, _n.dtype_code(x2.dtype), _n.stream_ptr())
        return dx.view(ctx.shape), None


def gated_act(x: torch.Tensor, act="silu") -> torch.Tensor:
    """``act(x[..., :F]) * x[..., F:]`` (SwiGLU/GeGLU with concatenated gate|up projection)."""
    code = _act_code(act)
    if x.is_cuda:
        return _Gated.apply(x.contiguous(), code)
    F_ = x.shape[-1] // 2
    return (_torch_act(x[..., :F_].float(), code) * x[..., F_:].float()).to(x.dtype)


def swiglu(x):
    return gated_act(x, ACT_SILU)


class _BiasAct(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, bias, act):
        C = x.shape[-1]
        x2 = x.reshape(-1, C)
        y = torch.empty_like(x2)
        _n.call("sxb_bias_act", x2.data_ptr(), _n.ptr(bias), y.data_ptr(), x2.shape[0], C, act,
                _n.dtype_code(x.dtype), _n.stream_ptr())
        ctx.save_for_backward(x2, bias)
        ctx.act, ctx.shape = act, x.shape
        return y.view(x.shape)

    @staticmethod
    def backward(ctx, dy):
        x2, bias = ctx.saved_tensors
        C = x2.shape[-1]
        dy2 = dy.reshape(-1, C).contiguous()
        dx = torch.empty_like(x2)
        _n.call("sxb_bias_act_bwd", dy2.data_ptr(), x2.data_ptr(), _n.ptr(bias), dx.data_ptr(),
                x2.shape[0], C, ctx.act, _n.dtype_code(x2.dtype), _n.stream_ptr())
        db = dx.sum(0) if bias is not None else None
        return dx.view(ctx.shape), db, None


def bias_act(x, bias=None, act="gelu"):
    code = _act_code(act)
    if x.is_cuda:
        return _BiasAct.apply(x.contiguous(), bias, code)
    h = x if bias is None else x + bias
    return _torch_act(h.float(), code).to(x.dtype)


# ------------------------------------------------------------------------ cross entropy
class _CrossEntropy(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels, ignore_index):
        V = logits.shape[-1]
        l2 = logits.reshape(-1, V)
        lab = labels.reshape(-1)
        rows = l2.shape[0]
        loss_rows = torch.empty(rows, dtype=torch.float32, device=logits.device)
        nvalid = (lab != ignore_index).sum().clamp_(min=1).to(torch.float32)
        inv = (1.0 / nvalid).reshape(1)
        # gradient is written in place of the logits (saves a [T, V] buffer)
        _n.call("sxb_cross_entropy", l2.data_ptr(), lab.data_ptr(), loss_rows.data_ptr(), None, rows, V,
                l2.stride(0), int(ignore_index), inv.data_ptr(), 1, _n.dtype_code(l2.dtype), _n.stream_ptr())
        ctx.save_for_backward(l2)
        ctx.shape = logits.shape
        return loss_rows.sum() * inv[0]

    @staticmethod
    def backward(ctx, dloss):
        (g,) = ctx.saved_tensors
        g.mul_(dloss.to(g.dtype))  # in place: the saved buffer is the (consumed) logits tensor
        return g.view(ctx.shape), None, None


def cross_entropy(logits: torch.Tensor, labels: torch.Tensor, ignore_index: int = -100) -> torch.Tensor:
    """Mean token cross entropy.  On CUDA the logits buffer is consumed (overwritten by dlogits)."""
    if logits.is_cuda and logits.dtype in (torch.bfloat16, torch.float16) and logits.shape[-1] % 8 == 0:
        return _CrossEntropy.apply(logits.contiguous(), labels.contiguous(), ignore_index)
    if logits.is_cuda:
        raise _n.NativeError("sxb cross_entropy expects bf16/fp16 logits with V % 8 == 0")
    return F.cross_entropy(logits.reshape(-1, logits.shape[-1]).float(), labels.reshape(-1),
                           ignore_index=ignore_index)


# ---------------------------------------------------------------------------- embedding
class _Embedding(torch.autograd.Function):
    @staticmethod
    def forward(ctx, ids, table):
        flat = ids.reshape(-1).contiguous()
        out = torch.empty((flat.numel(), table.shape[1]), dtype=table.dtype, device=table.device)
        _n.call("sxb_embedding_fwd", flat.data_ptr(), table.data_ptr(), out.data_ptr(), flat.numel(),
                table.shape[1], _n.dtype_code(table.dtype), _n.stream_ptr())
        _note_use(table)
        ctx.save_for_backward(flat, table)
        return out.view(*ids.shape, table.shape[1])

    @staticmethod
    def backward(ctx, dout):
        flat, table = ctx.saved_tensors
        d2 = dout.reshape(-1, table.shape[1]).contiguous()
        mg = getattr(table, "_sxb_main_grad", None)
        if mg is not None:
            if not getattr(table, "_sxb_grad_written", False):
                mg.zero_()
            dt = mg
        else:
            dt = torch.zeros_like(table)
        _n.call("sxb_embedding_bwd", flat.data_ptr(), d2.data_ptr(), dt.data_ptr(), flat.numel(),
                table.shape[1], _n.dtype_code(table.dtype), _n.stream_ptr())
        if mg is not None:
            table._sxb_grad_written = True
            hook = getattr(table, "_sxb_grad_ready", None)
            if hook is not None:
                hook(table)
            return None, None
        return None, dt


def embedding(ids: torch.Tensor, table: torch.Tensor) -> torch.Tensor:
    if table.is_cuda and table.dtype in (torch.bfloat16, torch.float16):
        return _Embedding.apply(ids, table)
    return F.embedding(ids, table)


# ---------------------------------------------------------------------------- attention
def attention_ref(q, k, v, causal=True, scale=None):
    """q [B,S,Hq,D], k/v [B,S,Hk,D] -> [B,S,Hq,D]; fp32 math."""
    B, S, Hq, D = q.shape
    Hk = k.shape[2]
    scale = scale if scale is not None else 1.0 / math.sqrt(D)
    qf = q.float().permute(0, 2, 1, 3)
    kf = k.float().permute(0, 2, 1, 3).repeat_interleave(Hq // Hk, dim=1)
    vf = v.float().permute(0, 2, 1, 3).repeat_interleave(Hq // Hk, dim=1)
    s = qf @ kf.transpose(-1, -2) * scale
    if causal:
        mask = torch.ones(S, k.shape[1], dtype=torch.bool, device=q.device).tril()
        s = s.masked_fill(~mask, float("-inf"))
    p = s.softmax(-1)
    return (p @ vf).permute(0, 2, 1, 3).to(q.dtype)


class _FlashAttn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, k, v, causal, scale):
        B, S, Hq, D = q.shape
        Hk = k.shape[2]
        out = torch.empty((B, S, Hq, D), dtype=q.dtype, device=q.device)
        lse = torch.empty((B, Hq, S), dtype=torch.float32, device=q.device)
        if flop_sink is not None:
            flop_sink("attention", 4.0 * B * Hq * S * S * D * (0.5 if causal else 1.0))
        _n.call("sxb_attn_fwd", q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr(), lse.data_ptr(),
                B, S, Hq, Hk, D, q.stride(1), k.stride(1), v.stride(1), out.stride(1), int(causal),
                float(scale), _n.stream_ptr())
        ctx.save_for_backward(q, k, v, out, lse)
        ctx.causal, ctx.scale = causal, scale
        return out

    @staticmethod
    def backward(ctx, dout):
        q, k, v, out, lse = ctx.saved_tensors
        B, S, Hq, D = q.shape
        Hk = k.shape[2]
        if dout.stride(-1) != 1 or dout.stride(2) != D or dout.stride(0) != S * dout.stride(1):
            dout = dout.contiguous()
        # one fused buffer so dq|dk|dv is directly the gradient of a fused QKV projection output
        dqkv = torch.empty((B, S, (Hq + 2 * Hk) * D), dtype=q.dtype, device=q.device)
        dq = dqkv[..., : Hq * D].view(B, S, Hq, D)
        dk = dqkv[..., Hq * D: (Hq + Hk) * D].view(B, S, Hk, D)
        dv = dqkv[..., (Hq + Hk) * D:].view(B, S, Hk, D)
        delta = torch.empty((B, Hq, S), dtype=torch.float32, device=q.device)
        _n.call("sxb_attn_bwd", q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr(), dout.data_ptr(),
                lse.data_ptr(), delta.data_ptr(), dq.data_ptr(), dk.data_ptr(), dv.data_ptr(),
                B, S, Hq, Hk, D, q.stride(1), k.stride(1), v.stride(1), out.stride(1), dout.stride(1),
                dq.stride(1), dk.stride(1), dv.stride(1), int(ctx.causal), float(ctx.scale), _n.stream_ptr())
        return dq, dk, dv, None, None


class _FlashAttnQKV(torch.autograd.Function):
    """Attention over a fused QKV projection output ``[B, S, (Hq + 2 Hk) * D]``.  The backward kernels write dq|dk|dv
    into one buffer of the same layout, which IS the gradient of the fused tensor: no slice-backward zero fills,
    copies and adds (3 + 3 + 2 launches and ~1 GB of traffic per layer at 8192 tokens of Llama-3-8B)."""

    @staticmethod
    def forward(ctx, qkv, Hq, Hk, D, causal, scale):
        B, S, C = qkv.shape
        q, k, v = _split_qkv(qkv, Hq, Hk, D)
        out = torch.empty((B, S, Hq, D), dtype=qkv.dtype, device=qkv.device)
        lse = torch.empty((B, Hq, S), dtype=torch.float32, device=qkv.device)
        if flop_sink is not None:
            flop_sink("attention", 4.0 * B * Hq * S * S * D * (0.5 if causal else 1.0))
        _n.call("sxb_attn_fwd", q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr(), lse.data_ptr(),
                B, S, Hq, Hk, D, C, C, C, out.stride(1), int(causal), float(scale), _n.stream_ptr())
        ctx.save_for_backward(qkv, out, lse)
        ctx.meta = (Hq, Hk, D, causal, scale)
        return out

    @staticmethod
    def backward(ctx, dout):
        qkv, out, lse = ctx.saved_tensors
        Hq, Hk, D, causal, scale = ctx.meta
        B, S, C = qkv.shape
        if dout.stride(-1) != 1 or dout.stride(2) != D or dout.stride(0) != S * dout.stride(1):
            dout = dout.contiguous()
        q, k, v = _split_qkv(qkv, Hq, Hk, D)
        dqkv = torch.empty_like(qkv)
        dq, dk, dv = _split_qkv(dqkv, Hq, Hk, D)
        delta = torch.empty((B, Hq, S), dtype=torch.float32, device=qkv.device)
        _n.call("sxb_attn_bwd", q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr(), dout.data_ptr(),
                lse.data_ptr(), delta.data_ptr(), dq.data_ptr(), dk.data_ptr(), dv.data_ptr(),
                B, S, Hq, Hk, D, C, C, C, out.stride(1), dout.stride(1), C, C, C, int(causal), float(scale),
                _n.stream_ptr())
        return dqkv, None, None, None, None, None


def _split_qkv(qkv, Hq, Hk, D):
    B, S, _ = qkv.shape
    return (qkv[..., : Hq * D].view(B, S, Hq, D), qkv[..., Hq * D: (Hq + Hk) * D].view(B, S, Hk, D),
            qkv[..., (Hq + Hk) * D:].view(B, S, Hk, D))


def flash_attention_qkv(qkv: torch.Tensor, n_q: int, n_kv: int, head_dim: int, causal: bool = True,
                        scale: Optional[float] = None) -> torch.Tensor:
    """``qkv [B, S, (n_q + 2 n_kv) * head_dim]`` (q heads | k heads | v heads) -> ``[B, S, n_q, head_dim]``; the
    gradient comes back as one fused tensor."""
    scale = scale if scale is not None else 1.0 / math.sqrt(head_dim)
    if qkv.is_cuda:
        if qkv.dtype != torch.bfloat16 or head_dim not in (64, 128):
            # fp16 / fp32 or another head size: the general entry point pads or takes the library path (see there)
            q, k, v = _split_qkv(qkv, n_q, n_kv, head_dim)
            return flash_attention(q, k, v, causal, scale)
        if qkv.dim() != 3 or qkv.shape[-1] != (n_q + 2 * n_kv) * head_dim:
            raise ValueError(f"fused qkv shape {tuple(qkv.shape)} != [B, S, {(n_q + 2 * n_kv) * head_dim}]")
        return _FlashAttnQKV.apply(qkv.contiguous(), n_q, n_kv, head_dim, causal, scale)
    q, k, v = _split_qkv(qkv, n_q, n_kv, head_dim)
    return attention_ref(q, k, v, causal, scale)


def attention_fwd_lse(q, k, v, causal: bool, scale: float):
    """Raw (non-autograd) attention forward returning ``(out [B,S,Hq,D], lse [B,Hq,S] natural log)`` — the building
    block of chunked / ring schedules that merge partial results themselves (``sequence/fpdt_layer.py``)."""
    B, S, Hq, D = q.shape
    if q.is_cuda:
        q, k, v = (t if _attn_layout_ok(t) else t.contiguous() for t in (q, k, v))
        out = torch.empty((B, S, Hq, D), dtype=q.dtype, device=q.device)
        lse = torch.empty((B, Hq, S), dtype=torch.float32, device=q.device)
        _n.call("sxb_attn_fwd", q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr(), lse.data_ptr(),
                B, S, Hq, k.shape[2], D, q.stride(1), k.stride(1), v.stride(1), out.stride(1), int(causal),
                float(scale), _n.stream_ptr())
        return out, lse
    Hk = k.shape[2]
    qf = q.float().permute(0, 2, 1, 3)
    kf = k.float().permute(0, 2, 1, 3).repeat_interleave(Hq // Hk, dim=1)
    vf = v.float().permute(0, 2, 1, 3).repeat_interleave(Hq // Hk, dim=1)
    s = qf @ kf.transpose(-1, -2) * scale
    if causal:
        s = s.masked_fill(~torch.ones(S, k.shape[1], dtype=torch.bool).tril(), float("-inf"))
    lse = torch.logsumexp(s, dim=-1)
    out = (torch.softmax(s, -1) @ vf).permute(0, 2, 1, 3).to(q.dtype)
    return out, lse


def attention_bwd_raw(q, k, v, out, dout, lse, causal: bool, scale: float):
    """Raw attention backward for one (q-block, kv-block) pair given the FINAL ``out``/``lse`` of the q rows (which may
    have been merged over many kv blocks): returns this pair's ``(dq, dk, dv)`` contributions."""
    B, S, Hq, D = q.shape
    Hk = k.shape[2]
    if q.is_cuda:
        q, k, v, out, dout = (t if _attn_layout_ok(t) else t.contiguous() for t in (q, k, v, out, dout))
        dq = torch.empty_like(q)
        dk = torch.empty_like(k)
        dv = torch.empty_like(v)
        delta = torch.empty((B, Hq, S), dtype=torch.float32, device=q.device)
        _n.call("sxb_attn_bwd", q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr(), dout.data_ptr(),
                lse.data_ptr(), delta.data_ptr(), dq.data_ptr(), dk.data_ptr(), dv.data_ptr(),
                B, S, Hq, Hk, D, q.stride(1), k.stride(1), v.stride(1), out.stride(1), dout.stride(1),
                dq.stride(1), dk.stride(1), dv.stride(1), int(causal), float(scale), _n.stream_ptr())
        return dq, dk, dv
    g = Hq // Hk
    qf = q.float().permute(0, 2, 1, 3)
    kf = k.float().permute(0, 2, 1, 3).repeat_interleave(g, dim=1)
    vf = v.float().permute(0, 2, 1, 3).repeat_interleave(g, dim=1)
    of = out.float().permute(0, 2, 1, 3)
    dof = dout.float().permute(0, 2, 1, 3)
    s = qf @ kf.transpose(-1, -2) * scale
    if causal:
        s = s.masked_fill(~torch.ones(S, k.shape[1], dtype=torch.bool).tril(), float("-inf"))
    p = torch.exp(s - lse.float()[..., None])
    dv_ = p.transpose(-1, -2) @ dof
    dp = dof @ vf.transpose(-1, -2)
    delta = (dof * of).sum(-1, keepdim=True)
    ds = p * (dp - delta) * scale
    dq_ = ds @ kf
    dk_ = ds.transpose(-1, -2) @ qf
    dk_ = dk_.view(B, Hk, g, k.shape[1], D).sum(2)
    dv_ = dv_.view(B, Hk, g, k.shape[1], D).sum(2)
    return (dq_.permute(0, 2, 1, 3).to(q.dtype), dk_.permute(0, 2, 1, 3).to(k.dtype),
            dv_.permute(0, 2, 1, 3).to(v.dtype))


# ------------------------------------------------------------------- attention with an additive per-key bias
_n.register({
    "sxb_attn_fwd_kbias": "ppppp" "pi" "iiiii" "llll" "i" "f" "p",
    "sxb_attn_bwd_kbias": "pppppppppp" "pi" "iiiii" "llllllll" "i" "f" "p",
})


class _FlashAttnKeyBias(torch.autograd.Function):
    """Flash attention whose scores get ``key_bias[b, key]`` added before the softmax (key-padding masks).  The
    kernels take the bias in raw-score units (bias / scale), fp32; no gradient flows to it."""

    @staticmethod
    def forward(ctx, q, k, v, kb_raw, causal, scale):
        B, S, Hq, D = q.shape
        Hk = k.shape[2]
        out = torch.empty((B, S, Hq, D), dtype=q.dtype, device=q.device)
        lse = torch.empty((B, Hq, S), dtype=torch.float32, device=q.device)
        _n.call("sxb_attn_fwd_kbias", q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr(), lse.data_ptr(),
                kb_raw.data_ptr(), int(kb_raw.dim() == 3), B, S, Hq, Hk, D, q.stride(1), k.stride(1), v.stride(1),
                out.stride(1), int(causal), float(scale), _n.stream_ptr())
        ctx.save_for_backward(q, k, v, out, lse, kb_raw)
        ctx.causal, ctx.scale = causal, scale
        return out

    @staticmethod
    def backward(ctx, dout):
        q, k, v, out, lse, kb_raw = ctx.saved_tensors
        B, S, Hq, D = q.shape
        Hk = k.shape[2]
        if dout.stride(-1) != 1 or dout.stride(2) != D or dout.stride(0) != S * dout.stride(1):
            dout = dout.contiguous()
        dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
        delta = torch.empty((B, Hq, S), dtype=torch.float32, device=q.device)
        _n.call("sxb_attn_bwd_kbias", q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr(), dout.data_ptr(),
                lse.data_ptr(), delta.data_ptr(), dq.data_ptr(), dk.data_ptr(), dv.data_ptr(), kb_raw.data_ptr(),
                int(kb_raw.dim() == 3), B, S, Hq, Hk, D, q.stride(1), k.stride(1), v.stride(1), out.stride(1), dout.stride(1),
                dq.stride(1), dk.stride(1), dv.stride(1), int(ctx.causal), float(ctx.scale), _n.stream_ptr())
        return dq, dk, dv, None, None, None


def key_bias_supported(q: torch.Tensor, key_bias: torch.Tensor) -> bool:
    """Whether :func:`flash_attention_key_bias` runs this on the flash kernels (else: materialised scores)."""
    B, S, Hq = q.shape[0], q.shape[1], q.shape[2]
    return (q.is_cuda and q.dtype == torch.bfloat16 and q.shape[-1] in (64, 128) and S % 4 == 0
            and tuple(key_bias.shape) in ((B, S), (B, Hq, S)))


def flash_attention_key_bias(q, k, v, key_bias, causal: bool = False, scale: Optional[float] = None):
    """``softmax(q k^T * scale + key_bias[:, None, None, :]) v`` — q ``[B, S, Hq, D]``, k/v ``[B, S, Hk, D]``,
    ``key_bias`` ``[B, S]`` additive (``-inf`` / large negative on padded keys) or ``[B, Hq, S]`` (a bias per head and
    key: ALiBi in its BLOOM form, slope x key position).  BERT-style encoders and the v1 inference path pass their
    attention mask this way (reference: the additive ``attn_mask`` / ``alibi`` of the fused softmax)."""
    D = q.shape[-1]
    scale = scale if scale is not None else 1.0 / math.sqrt(D)
    if q.is_cuda and key_bias_supported(q, key_bias):
        q, k, v = (t if _attn_layout_ok(t) else t.contiguous() for t in (q, k, v))
        kb_raw = (key_bias.detach().to(torch.float32) / scale).contiguous()
        return _FlashAttnKeyBias.apply(q, k, v, kb_raw, causal, scale)
    B, S, Hq, _ = q.shape
    Hk = k.shape[2]
    qf = q.float().permute(0, 2, 1, 3)
    kf = k.float().permute(0, 2, 1, 3).repeat_interleave(Hq // Hk, dim=1)
    vf = v.float().permute(0, 2, 1, 3).repeat_interleave(Hq // Hk, dim=1)
    kb = key_bias.float()
    sc = qf @ kf.transpose(-1, -2) * scale + (kb[:, None, None, :] if kb.dim() == 2 else kb[:, :, None, :])
    if causal:
        sc = sc.masked_fill(~torch.ones(S, S, dtype=torch.bool, device=q.device).tril(), float("-inf"))
    return (sc.softmax(-1) @ vf).permute(0, 2, 1, 3).to(q.dtype)


# ------------------------------------------------------------------- packed-sequence ("varlen") attention
_n.register({
    "sxb_attn_fwd_varlen": "ppppp" "p" "iiiiii" "llll" "i" "f" "p",
    "sxb_attn_bwd_varlen": "pppppppppp" "p" "iiiiii" "llllllll" "i" "f" "p",
})


def attention_varlen_ref(q, k, v, cu_seqlens, causal=True, scale=None):
    """Oracle: every sequence ``[cu[i], cu[i+1])`` of the packed ``[T, H, D]`` tensors attends only to itself."""
    out = torch.empty_like(q)
    cu = cu_seqlens.tolist()
    for a, b in zip(cu[:-1], cu[1:]):
        if b > a:
            out[a:b] = attention_ref(q[None, a:b], k[None, a:b], v[None, a:b], causal, scale)[0]
    return out


class _FlashAttnVarlen(torch.autograd.Function):
    """q [T, Hq, D], k/v [T, Hk, D] packed; the same tcgen05 kernels as the dense path with a per-CTA row window
    (``csrc/attention.cu``: ``sxb_attn_fwd_varlen`` / ``sxb_attn_bwd_varlen``)."""

    @staticmethod
    def forward(ctx, q, k, v, cu_seqlens, max_seqlen, causal, scale):
        T, Hq, D = q.shape
        Hk = k.shape[1]
        nseq = cu_seqlens.numel() - 1
        out = torch.empty((T, Hq, D), dtype=q.dtype, device=q.device)
        lse = torch.empty((Hq, T), dtype=torch.float32, device=q.device)
        _n.call("sxb_attn_fwd_varlen", q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr(), lse.data_ptr(),
                cu_seqlens.data_ptr(), nseq, int(max_seqlen), T, Hq, Hk, D, q.stride(0), k.stride(0), v.stride(0),
                out.stride(0), int(causal), float(scale), _n.stream_ptr())
        ctx.save_for_backward(q, k, v, out, lse, cu_seqlens)
        ctx.cfg = (int(max_seqlen), bool(causal), float(scale))
        return out

    @staticmethod
    def backward(ctx, dout):
        q, k, v, out, lse, cu = ctx.saved_tensors
        max_seqlen, causal, scale = ctx.cfg
        T, Hq, D = q.shape
        Hk = k.shape[1]
        if dout.stride(-1) != 1 or dout.stride(1) != D:
            dout = dout.contiguous()
        dqkv = torch.empty((T, (Hq + 2 * Hk) * D), dtype=q.dtype, device=q.device)
        dq = dqkv[:, : Hq * D].view(T, Hq, D)
        dk = dqkv[:, Hq * D: (Hq + Hk) * D].view(T, Hk, D)
        dv = dqkv[:, (Hq + Hk) * D:].view(T, Hk, D)
        delta = torch.empty((Hq, T), dtype=torch.float32, device=q.device)
        _n.call("sxb_attn_bwd_varlen", q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr(), dout.data_ptr(),
                lse.data_ptr(), delta.data_ptr(), dq.data_ptr(), dk.data_ptr(), dv.data_ptr(), cu.data_ptr(),
                cu.numel() - 1, max_seqlen, T, Hq, Hk, D, q.stride(0), k.stride(0), v.stride(0), out.stride(0),
                dout.stride(0), dq.stride(0), dk.stride(0), dv.stride(0), int(causal), float(scale), _n.stream_ptr())
        return dq, dk, dv, None, None, None, None


def flash_attention_varlen(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, cu_seqlens: torch.Tensor,
                           max_seqlen: int, causal: bool = True, scale: Optional[float] = None) -> torch.Tensor:
    """Attention over packed sequences: q ``[T, Hq, D]``, k/v ``[T, Hk, D]`` (strided views of a fused QKV buffer are
    fine), ``cu_seqlens`` int32 ``[n + 1]`` on the same device, ``max_seqlen`` = the longest sequence (a host int: it
    sizes the grid)."""
    D = q.shape[-1]
    scale = scale if scale is not None else 1.0 / math.sqrt(D)
    if not q.is_cuda:
        return attention_varlen_ref(q, k, v, cu_seqlens, causal, scale)
    if q.dtype != torch.bfloat16 or D not in (64, 128):
        raise _n.NativeError(f"sxb flash_attention_varlen: bf16 and head_dim 64/128 only (got {q.dtype}, {D})")
    if cu_seqlens.dtype != torch.int32 or not cu_seqlens.is_cuda:
        raise _n.NativeError("sxb flash_attention_varlen: cu_seqlens must be an int32 CUDA tensor")
    ok = lambda t: t.stride(2) == 1 and t.stride(1) == D
    q, k, v = (t if ok(t) else t.contiguous() for t in (q, k, v))
    return _FlashAttnVarlen.apply(q, k, v, cu_seqlens.contiguous(), max_seqlen, causal, scale)


def _attn_layout_ok(t):
    B, S, H, D = t.shape
    return t.stride(3) == 1 and t.stride(2) == D and t.stride(0) == S * t.stride(1)


def flash_attention(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, causal: bool = True,
                    scale: Optional[float] = None) -> torch.Tensor:
    """q [B,S,Hq,D], k/v [B,S,Hk,D] (may be strided views of one fused QKV buffer)."""
    D = q.shape[-1]
    scale = scale if scale is not None else 1.0 / math.sqrt(D)
    if q.is_cuda:
        if q.dtype != torch.bfloat16 or D > 128 or D % 8:
            # fp16 / fp32 or head sizes beyond the kernels': PyTorch's attention, said once
            _warn_once("attn", f"sxb flash_attention: {q.dtype}, head_dim {D} runs on torch SDPA (the tcgen05 kernels "
                               "take bf16 with head_dim <= 128, multiple of 8)")
            o = F.scaled_dot_product_attention(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2), is_causal=causal,
                                               scale=scale, enable_gqa=q.shape[2] != k.shape[2])
            return o.transpose(1, 2).contiguous()
        if D not in (64, 128):
            # Other head sizes (32, 80, 96, ...) run on the 64 / 128 kernels with zero-padded channels: zeros add
            # nothing to q . k, the padded channels of v give zero output columns that are sliced off again, and the
            # softmax scale stays that of the true head size.  Exact; costs the padding copies.
            Dp = 64 if D < 64 else 128
            pad = (0, Dp - D)
            out = flash_attention(F.pad(q, pad), F.pad(k, pad), F.pad(v, pad), causal, scale)
            return out[..., :D]
        q = q if _attn_layout_ok(q) else q.contiguous()
        k = k if _attn_layout_ok(k) else k.contiguous()
        v = v if _attn_layout_ok(v) else v.contiguous()
        return _FlashAttn.apply(q, k, v, causal, scale)
    return attention_ref(q, k, v, causal, scale)
