"""Run a few cuBLAS bf16 GEMMs (torch.matmul) so that ncu can capture the library kernel as a reference point."""
import torch
M = N = K = 8192
a = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
b = torch.randn(N, K, device="cuda", dtype=torch.bfloat16)
out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
for _ in range(8):
    torch.matmul(a, b.t(), out=out)
torch.cuda.synchronize()
print("ok")
