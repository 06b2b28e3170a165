#!/bin/bash
# 1-GPU call: block-scaled fp8 GEMM bring-up, K5 on one GPU, re-check of the tolerance changes
mkdir -p gpurun_out
timeout 300 python tools/kernel_check.py mx > gpurun_out/c3_mx.jsonl 2>&1; echo "mx rc=$?"
timeout 300 python tools/kernel_check.py k5 > gpurun_out/c3_k5.jsonl 2>&1; echo "k5 rc=$?"
timeout 300 python tools/kernel_check.py attn_bwd fp8 gemm_tn gemm_tt > gpurun_out/c3_checks.jsonl 2>&1; echo "rechecks rc=$?"
grep -h '"ok": false' gpurun_out/c3_mx.jsonl gpurun_out/c3_k5.jsonl gpurun_out/c3_checks.jsonl | cut -c1-420
grep -h 'perf\|tflops\|"ms"' gpurun_out/c3_mx.jsonl gpurun_out/c3_k5.jsonl | cut -c1-300
tail -3 gpurun_out/c3_mx.jsonl | cut -c1-600; tail -3 gpurun_out/c3_k5.jsonl | cut -c1-600
