"""ncu target: the same product as NT (K-major operands) and TN (MN-major operands), bf16 out."""
import sys, torch
sys.path.insert(0, "/root/repo")
from b200flow.ops import gemm as G
M, N, K = 11008, 4096, 4096
dy = torch.randn(K, M, device="cuda", dtype=torch.bfloat16)
x = torch.randn(K, N, device="cuda", dtype=torch.bfloat16)
dyt, xt = dy.t().contiguous(), x.t().contiguous()
out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
for _ in range(2):
    G.gemm(dyt, xt, out=out)
    G.gemm(dy, x, a_mn_major=True, b_mn_major=True, out=out)
torch.cuda.synchronize()
