"""ncu target for the final kernels: CTA-pair GEMM (bf16 TMA-store and fp32 reducing-store epilogues) and
attention forward / backward at the Llama-2 7B shapes."""
import sys
import torch
sys.path.insert(0, "/root/repo")
from b200flow.ops import attention as A
from b200flow.ops import gemm as G

M, N, K = 4096, 22016, 4096
a = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
b = torch.randn(N, K, device="cuda", dtype=torch.bfloat16)
c = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
dy = torch.randn(K, 11008, device="cuda", dtype=torch.bfloat16)
x = torch.randn(K, 4096, device="cuda", dtype=torch.bfloat16)
acc = torch.zeros(11008, 4096, device="cuda", dtype=torch.float32)
qkv = torch.randn(1, 4096, 3, 32, 128, device="cuda", dtype=torch.bfloat16)
q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]
for _ in range(2):
    G.gemm(a, b, out=c)
    G.gemm(dy, x, a_mn_major=True, b_mn_major=True, out=acc, accumulate=True)
    o, lse = A.attn_fwd_raw(q, k, v, True, 128 ** -0.5)
    A.attn_bwd_raw(q, k, v, o, torch.ones_like(o), lse, True, 128 ** -0.5)
torch.cuda.synchronize()
