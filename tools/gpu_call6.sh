#!/bin/bash
mkdir -p gpurun_out
timeout 200 python tools/kernel_check.py attn_drop > gpurun_out/c6_attn_drop.jsonl 2>&1; echo "attn_drop rc=$?"
timeout 120 python tools/k4_loopback_perf.py > gpurun_out/c6_loop_default.jsonl 2>&1; echo "loop default rc=$?"
B2F_RS_INTERLEAVE=0 timeout 120 python tools/k4_loopback_perf.py > gpurun_out/c6_loop_nointer.jsonl 2>&1; echo "loop nointer rc=$?"
B2F_RS_DEBUG=1 timeout 120 python tools/k4_loopback_perf.py > gpurun_out/c6_loop_dbg1.jsonl 2>&1; echo "loop dbg1 rc=$?"
B2F_RS_DEBUG=2 timeout 120 python tools/k4_loopback_perf.py > gpurun_out/c6_loop_dbg2.jsonl 2>&1; echo "loop dbg2 rc=$?"
grep -h '"ok": false' gpurun_out/c6_attn_drop.jsonl | cut -c1-300
cat gpurun_out/c6_loop_*.jsonl | cut -c1-260
