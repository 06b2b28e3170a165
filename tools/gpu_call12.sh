#!/bin/bash
# 1-GPU call: attention forward with part of the exp2 work on the FMA pipe (B2F_ATTN_POLY = pairs of four)
mkdir -p gpurun_out
for P in 0 1 2; do
  B2F_ATTN_POLY=$P timeout 300 python tools/kernel_check.py attn_fwd attn_perf > gpurun_out/c12_attn_poly$P.log 2>&1; echo "poly $P rc=$?"
  grep -h '"ok": false\|summary\|attn_perf' gpurun_out/c12_attn_poly$P.log | cut -c1-400
done
