"""Sustained (power-capped) GEMM throughput: each kernel runs back-to-back for ~3 s while nvidia-smi
clocks and power are sampled.  Burst numbers hide this: the end-to-end step runs at the power cap."""
import json, subprocess, sys, threading, time
import torch
sys.path.insert(0, "/root/repo")
from b200flow.ops import gemm as G

def sample(stop, rows):
    while not stop.is_set():
        out = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,power.draw", "--format=csv,noheader,nounits", "-i", "0"],
                             capture_output=True, text=True).stdout.strip().split(",")
        try:
            rows.append((float(out[0]), float(out[1])))
        except Exception:
            pass
        stop.wait(0.1)

def run(name, fn, flops, secs=3.0):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); fn(); e1.record(); torch.cuda.synchronize()
    n = max(10, int(secs * 1e3 / e0.elapsed_time(e1)))
    stop, rows = threading.Event(), []
    th = threading.Thread(target=sample, args=(stop, rows)); th.start()
    e0.record()
    for _ in range(n):
        fn()
    e1.record(); torch.cuda.synchronize()
    stop.set(); th.join()
    ms = e0.elapsed_time(e1) / n
    tail = rows[len(rows) // 2:] or [(0, 0)]
    print(json.dumps({"name": name, "ms": round(ms, 4), "tflops": round(flops / ms / 1e9, 1),
                      "sm_mhz_late": sum(r[0] for r in tail) / len(tail), "power_w_late": sum(r[1] for r in tail) / len(tail)}), flush=True)
    time.sleep(2.0)

for (M, N, K) in [(4096, 22016, 4096), (4096, 4096, 11008), (8192, 8192, 8192)]:
    a = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
    b = torch.randn(N, K, device="cuda", dtype=torch.bfloat16)
    c = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    fl = 2.0 * M * N * K
    run(f"cublas_{M}x{N}x{K}", lambda: torch.matmul(a, b.t(), out=c), fl)
    run(f"b2f_2cta_{M}x{N}x{K}", lambda: G.gemm(a, b, out=c), fl)
    G.GEMM_IMPL = "1cta"
    run(f"b2f_1cta_{M}x{N}x{K}", lambda: G.gemm(a, b, out=c), fl)
    G.GEMM_IMPL = "2cta"
