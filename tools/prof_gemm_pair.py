"""ncu target: the same bf16 GEMM once through cuBLAS and once through the CTA-pair kernel."""
import sys, torch
sys.path.insert(0, "/root/repo")
from b200flow.ops import gemm as G
M, N, K = 4096, 22016, 4096
a = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
b = torch.randn(N, K, device="cuda", dtype=torch.bfloat16)
c = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
for _ in range(2):
    torch.matmul(a, b.t(), out=c)
    G.gemm(a, b, out=c)
torch.cuda.synchronize()
