"""Phase timing of the attention kernels (needs a build with B2F_NVCC_EXTRA=-DB2F_ATTN_DEBUG)."""
import ctypes, sys, torch
sys.path.insert(0, "/root/repo")
from b200flow.ops import attention as A, _lib
B, S, H, D = 1, 4096, 32, 128
qkv = torch.randn(B, S, 3, H, D, device="cuda", dtype=torch.bfloat16)
q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]
for _ in range(3):
    A.attn_fwd_raw(q, k, v, True, D ** -0.5)
torch.cuda.synchronize()
lib = _lib.cuda_lib()
out = (ctypes.c_ulonglong * 16)()
lib.b2f_attn_debug(out, 1)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
A.attn_fwd_raw(q, k, v, True, D ** -0.5)
e1.record()
torch.cuda.synchronize()
print("kernel ms", e0.elapsed_time(e1))
lib.b2f_attn_debug(out, 0)
v_ = list(out)
tiles = max(v_[4], 1)
print(f"softmax A lane0: tiles {v_[4]}  per tile: total {v_[0]/tiles:.0f}  wait_s {v_[1]/tiles:.0f}  ld+max {v_[2]/tiles:.0f}  exp+st {v_[3]/tiles:.0f}")
print(f"mma: loop total/tileA {v_[5]/tiles:.0f}  wait_p {v_[6]/tiles:.0f}  wait_kv {v_[7]/tiles:.0f}")
print(f"CTAs {v_[12]}  sum CTA lifetime {v_[8]} cyc = {v_[8]/148:.0f}/SM; prologue/CTA {v_[9]/max(v_[12],1):.0f}; softmax epilogue/CTA {v_[10]/max(v_[12],1):.0f}; "
      f"effective SM clock {v_[8]/max(v_[11],1):.3f} GHz; loop share {v_[5]/max(v_[8],1):.2f}")
