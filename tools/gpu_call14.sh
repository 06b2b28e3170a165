#!/bin/bash
# 1-GPU call: attention forward with P published in quarters (P V starts on the first 32 keys) and 3-input max
mkdir -p gpurun_out
for P in 0 1; do
  B2F_ATTN_POLY=$P timeout 300 python tools/kernel_check.py attn_fwd attn_bwd attn_drop attn_perf > gpurun_out/c14_attn_poly$P.log 2>&1; echo "poly $P rc=$?"
  grep -h '"ok": false\|summary\|attn_perf' gpurun_out/c14_attn_poly$P.log | cut -c1-400
done
B2F_CUDA_LIB=libb200flow_cuda_attndbg.so timeout 120 python tools/dbg_attn.py 2>&1 | tail -4 | tee gpurun_out/c14_attn_dbg.txt
