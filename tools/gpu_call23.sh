#!/bin/bash
# 1-GPU call: the default bench line (Llama-2 7B with the twin inside it), attention against the libraries' backward,
# N=1 points of BERT-large and GPT-MoE
mkdir -p gpurun_out
timeout 600 python bench.py > gpurun_out/c23_bench_default.json 2> gpurun_out/c23_bench_default.err; echo "bench default rc=$?"
tail -n 1 gpurun_out/c23_bench_default.json | cut -c1-2500
timeout 200 python tools/kernel_check.py attn_perf > gpurun_out/c23_attn_perf.log 2>&1; echo "attn_perf rc=$?"; grep -h attn_perf gpurun_out/c23_attn_perf.log | cut -c1-900
timeout 300 python bench.py --workload bert-large --steps 12 --warmup 5 --baseline-steps 10 > gpurun_out/c23_bert1.json 2> gpurun_out/c23_bert1.err; echo "bert1 rc=$?"
tail -n 1 gpurun_out/c23_bert1.json | cut -c1-1500
timeout 300 python bench.py --workload gpt-moe-1.3b --steps 6 --warmup 8 --baseline none > gpurun_out/c23_moe1.json 2> gpurun_out/c23_moe1.err; echo "moe1 rc=$?"
tail -n 1 gpurun_out/c23_moe1.json | cut -c1-900
