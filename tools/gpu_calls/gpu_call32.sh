#!/bin/bash
timeout 120 python - <<'PY'
import torch, json
from shuffle_exchange_b200.ops import functional as F
x = torch.randn(8192, 28672, device="cuda", dtype=torch.bfloat16, requires_grad=True)
def t(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n
y = F.swiglu(x)
g = torch.randn_like(y)
fwd = t(lambda: F.swiglu(x.detach()))
def bwd():
    x.grad = None
    y.backward(g, retain_graph=True)
b = t(bwd)
print(json.dumps({"swiglu_fwd_ms": fwd, "fwd_TBps": 3 * 8192 * 14336 * 2 / fwd / 1e9, "swiglu_bwd_ms": b, "bwd_TBps": 5 * 8192 * 14336 * 2 / b / 1e9}))
PY
export SXB_LIB_PATH=build/libsxb200_oldew.so
timeout 120 python - <<'PY'
import torch, json
from shuffle_exchange_b200.ops import functional as F
x = torch.randn(8192, 28672, device="cuda", dtype=torch.bfloat16, requires_grad=True)
def t(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n
y = F.swiglu(x)
g = torch.randn_like(y)
fwd = t(lambda: F.swiglu(x.detach()))
def bwd():
    x.grad = None
    y.backward(g, retain_graph=True)
b = t(bwd)
print(json.dumps({"swiglu_fwd_ms": fwd, "fwd_TBps": 3 * 8192 * 14336 * 2 / fwd / 1e9, "swiglu_bwd_ms": b, "bwd_TBps": 5 * 8192 * 14336 * 2 / b / 1e9}))
PY
