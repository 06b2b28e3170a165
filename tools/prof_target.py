"""Tiny single-GPU programs for ncu captures: `python tools/prof_target.py gemm|attn|norm`."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from b200flow.ops import attention as A  # noqa: E402
from b200flow.ops import gemm as G  # noqa: E402
from b200flow.ops import norm as N  # noqa: E402

what = sys.argv[1] if len(sys.argv) > 1 else "gemm"
dev = "cuda"
if what == "gemm":
    a = torch.randn(4096, 4096, device=dev, dtype=torch.bfloat16)
    b = torch.randn(11008 * 2, 4096, device=dev, dtype=torch.bfloat16)
    c = torch.empty(4096, 11008 * 2, device=dev, dtype=torch.bfloat16)
    for _ in range(5):
        G.gemm(a, b, out=c)
elif what == "attn":
    qkv = torch.randn(1, 4096, 3, 32, 128, device=dev, dtype=torch.bfloat16)
    q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]
    for _ in range(3):
        o, lse = A.attn_fwd_raw(q, k, v, True, 128 ** -0.5)
        A.attn_bwd_raw(q, k, v, o, torch.randn_like(o), lse, True, 128 ** -0.5)
elif what == "norm":
    x = torch.randn(4096, 4096, device=dev, dtype=torch.bfloat16, requires_grad=True)
    g = torch.ones(4096, device=dev, dtype=torch.bfloat16, requires_grad=True)
    for _ in range(3):
        N.rmsnorm(x, g).sum().backward()
torch.cuda.synchronize()
print("ok", what)
