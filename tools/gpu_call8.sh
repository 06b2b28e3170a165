#!/bin/bash
mkdir -p gpurun_out
P="python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1"
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -15 > gpurun_out/c8_pytest.log; echo "pytest rc=$?"; tail -4 gpurun_out/c8_pytest.log
timeout 300 $P --master-port 29747 tools/multigpu_check.py arena allreduce zero1 > gpurun_out/c8_allreduce.log 2>&1; echo "allreduce rc=$?"
timeout 300 $P --master-port 29749 tools/multigpu_check.py zero1_perf > gpurun_out/c8_zero1_perf.log 2>&1; echo "zero1_perf rc=$?"
B2F_HALF=fp16 timeout 300 $P --master-port 29751 tools/multigpu_check.py fused_gemm tp > gpurun_out/c8_fp16_tp.log 2>&1; echo "fp16 fused tp rc=$?"
grep -h '"ok": false\|summary' gpurun_out/c8_allreduce.log gpurun_out/c8_zero1_perf.log gpurun_out/c8_fp16_tp.log | cut -c1-400
grep -h 'perf\|nccl_bf16' gpurun_out/c8_allreduce.log gpurun_out/c8_zero1_perf.log | grep '"rank": 0' | cut -c1-300
