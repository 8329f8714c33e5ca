"""Cycle accounting of the heaviest dQ tile (SXB_ATTN_DEBUG=3; diagnostic only).  Needs a library whose csrc/attention.cu was compiled with -DSXB_ATTN_PROF
(build it next to the default one and point SXB_LIB_PATH at it): where the MMA warp and a row warp of
one CTA spend their time.  S=8192, 32q/8kv, d=128, causal."""
import ctypes
import json
import os
import sys

os.environ["SXB_ATTN_DEBUG"] = "3"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from shuffle_exchange_b200.native import lib as n
from shuffle_exchange_b200.ops import functional as F

n.register({"sxb_attn_prof": "p"})
torch.manual_seed(0)
B, S, Hq, Hk, D = 1, 8192, 32, 8, 128
q = torch.randn(B, S, Hq, D, device="cuda", dtype=torch.bfloat16, requires_grad=True)
k = torch.randn(B, S, Hk, D, device="cuda", dtype=torch.bfloat16, requires_grad=True)
v = torch.randn(B, S, Hk, D, device="cuda", dtype=torch.bfloat16, requires_grad=True)
o = F.flash_attention(q, k, v, causal=True)
g = torch.randn_like(o)
for _ in range(3):
    o.backward(g, retain_graph=True)
torch.cuda.synchronize()
buf = (ctypes.c_ulonglong * 8)()
n.call("sxb_attn_prof", ctypes.cast(buf, ctypes.c_void_p))
names = ["mma_total", "mma_wait_kv", "mma_wait_ds", "iterations", "row_total", "row_wait_s", "row_ld_math", "row_st_fence"]
r = dict(zip(names, [int(x) for x in buf]))
it = max(1, r["iterations"])
r["per_iteration"] = {k: round(v / it, 1) for k, v in r.items() if k != "iterations"}
print(json.dumps(r))
