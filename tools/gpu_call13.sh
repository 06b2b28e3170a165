#!/bin/bash
# 1-GPU call: phase timing of the attention forward kernel (instrumented build), with parts of the work removed
mkdir -p gpurun_out
export B2F_CUDA_LIB=libb200flow_cuda_attndbg.so
for M in 0 1 2 3; do
  echo "## dbg_mode $M (1: softmax does no math, 2: also no P V, 3: also no Q K^T)"
  B2F_ATTN_DBG_MODE=$M timeout 120 python tools/dbg_attn.py 2>&1 | tail -5
done > gpurun_out/c13_attn_dbg.txt
cat gpurun_out/c13_attn_dbg.txt
