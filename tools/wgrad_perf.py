"""wgrad-shaped GEMMs (TN, fp32 accumulate into main_grad) vs the same product with a bf16 output."""
import json, sys, torch
sys.path.insert(0, "/root/repo")
from b200flow.ops import gemm as G
from tools.kernel_check import timeit
for (M, N, K) in [(11008, 4096, 4096), (22016, 4096, 4096), (4096, 4096, 4096), (4096, 11008, 4096), (6144, 4096, 4096)]:
    dy = torch.randn(K, M, device="cuda", dtype=torch.bfloat16)   # [tokens, out]
    x = torch.randn(K, N, device="cuda", dtype=torch.bfloat16)    # [tokens, in]
    acc = torch.zeros(M, N, device="cuda", dtype=torch.float32)
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    fl = 2.0 * M * N * K
    t_acc = timeit(lambda: G.gemm(dy, x, a_mn_major=True, b_mn_major=True, out=acc, accumulate=True))
    t_f32 = timeit(lambda: G.gemm(dy, x, a_mn_major=True, b_mn_major=True, out=acc))
    t_bf = timeit(lambda: G.gemm(dy, x, a_mn_major=True, b_mn_major=True, out=out))
    dyt, xt = dy.t().contiguous(), x.t().contiguous()
    t_nt = timeit(lambda: G.gemm(dyt, xt, out=out))
    t_cb = timeit(lambda: torch.matmul(dy.t(), x, out=out))
    print(json.dumps({"shape": [M, N, K], "tn_f32_accum_tflops": round(fl / t_acc / 1e9), "tn_f32_tflops": round(fl / t_f32 / 1e9),
                      "tn_bf16_tflops": round(fl / t_bf / 1e9), "nt_bf16_tflops": round(fl / t_nt / 1e9), "cublas_tn_bf16": round(fl / t_cb / 1e9)}), flush=True)
