"""Time the attention backward kernels under the SXB_ATTN_DEBUG profiling modes (diagnostic only)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from shuffle_exchange_b200.ops import functional as F
torch.manual_seed(0)
B, S, Hq, Hk, D = 1, 8192, 32, 8, 128
q = torch.randn(B, S, Hq, D, device="cuda", dtype=torch.bfloat16, requires_grad=True)
k = torch.randn(B, S, Hk, D, device="cuda", dtype=torch.bfloat16, requires_grad=True)
v = torch.randn(B, S, Hk, D, device="cuda", dtype=torch.bfloat16, requires_grad=True)
o = F.flash_attention(q, k, v, causal=True)
g = torch.randn_like(o)
from torch.profiler import profile, ProfilerActivity
for _ in range(3):
    o.backward(g, retain_graph=True)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(5):
        with torch.no_grad():
            F.flash_attention(q, k, v, causal=True)
        o.backward(g, retain_graph=True)
    torch.cuda.synchronize()
out = {e.key[:40]: round(e.device_time_total / 5000.0, 3) for e in prof.key_averages() if "attn" in e.key}
print(json.dumps({"mode": os.environ.get("SXB_ATTN_DEBUG", "0"), "mc": os.environ.get("SXB_ATTN_MC", "1"), "ms": out}))
