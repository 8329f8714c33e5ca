"""Cycle accounting of the heaviest tile of the 128-key forward kernel (SXB_ATTN_DEBUG=3, SXB_ATTN_FWD128=1)."""
import ctypes, json, os, sys
os.environ["SXB_ATTN_DEBUG"] = "3"
os.environ.setdefault("SXB_ATTN_FWD128", "1")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from shuffle_exchange_b200.native import lib as n
from shuffle_exchange_b200.ops import functional as F
n.register({"sxb_attn_prof": "p"})
torch.manual_seed(0)
B, S, Hq, Hk, D = 1, 8192, 32, 8, 128
q = torch.randn(B, S, Hq, D, device="cuda", dtype=torch.bfloat16)
k = torch.randn(B, S, Hk, D, device="cuda", dtype=torch.bfloat16)
v = torch.randn(B, S, Hk, D, device="cuda", dtype=torch.bfloat16)
with torch.no_grad():
    for _ in range(3):
        F.flash_attention(q, k, v, causal=True)
torch.cuda.synchronize()
buf = (ctypes.c_ulonglong * 8)()
n.call("sxb_attn_prof", ctypes.cast(buf, ctypes.c_void_p))
names = ["mma_total", "mma_wait_kv_or_k", "mma_wait_p_or_p0", "blocks", "row_total", "row_wait_s", "row_work", "mma_wait_v_or_p1"]
r = dict(zip(names, [int(x) for x in buf]))
it = max(1, r["blocks"])
r["per_block"] = {k: round(v / it, 1) for k, v in r.items() if k != "blocks"}
print(json.dumps(r))
