import ctypes, torch, sys
sys.path.insert(0, "/root/repo")
from b200flow.ops import gemm as G, _lib
a = torch.randn(4096, 4096, device="cuda", dtype=torch.bfloat16)
b = torch.randn(22016, 4096, device="cuda", dtype=torch.bfloat16)
for _ in range(3):
    G.gemm(a, b)
out = (ctypes.c_longlong * 8)()
_lib.cuda_lib().b2f_gemm2_debug(out)
tot, tfull, tacc, nkb = out[0], out[1], out[2], out[3]
print(f"total {tot} cyc, wait_full {tfull} ({100*tfull/tot:.1f}%), wait_acc {tacc} ({100*tacc/tot:.1f}%), kblocks {nkb}, cyc/kblock {tot/nkb:.1f}")
