"""Phase timing of the attention backward kernels (build with B2F_NVCC_EXTRA=-DB2F_ATTN_DEBUG)."""
import ctypes, sys, torch
sys.path.insert(0, "/root/repo")
from b200flow.ops import attention as A, _lib
B, S, H, D = 1, 4096, 32, 128
qkv = torch.randn(B, S, 3, H, D, device="cuda", dtype=torch.bfloat16)
q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]
o, lse = A.attn_fwd_raw(q, k, v, True, D ** -0.5)
do = torch.randn_like(o)
for _ in range(3):
    A.attn_bwd_raw(q, k, v, o, do, lse, True, D ** -0.5)
torch.cuda.synchronize()
lib = _lib.cuda_lib()
out = (ctypes.c_ulonglong * 16)()
lib.b2f_attn_debug(out, 1)
A.attn_bwd_raw(q, k, v, o, do, lse, True, D ** -0.5)
torch.cuda.synchronize()
lib.b2f_attn_debug(out, 0)
v_ = list(out)
t = max(v_[4], 1)
print(f"dkv compute lane0: tiles {v_[4]}  per tile: wait {v_[1]/t:.0f}  load+bar {v_[2]/t:.0f}  math+store {v_[3]/t:.0f}")
t = max(v_[8], 1)
print(f"dq  compute lane0: tiles {v_[8]}  per tile: wait {v_[5]/t:.0f}  load+bar {v_[6]/t:.0f}  math+store {v_[7]/t:.0f}")
