#!/bin/bash
# 1-GPU call: column-sum / conv epilogue checks, Llama default arm (uniform tokens, constant rate), BERT-large with the
# column-sum kernel behind every bias gradient
mkdir -p gpurun_out
timeout 200 python tools/kernel_check.py detection > gpurun_out/c24_det.log 2>&1; echo "detection rc=$?"; grep -h '"ok": false\|summary\|colsum_perf' gpurun_out/c24_det.log | cut -c1-300
timeout 400 python bench.py --baseline none > gpurun_out/c24_llama1.json 2> gpurun_out/c24_llama1.err; echo "llama1 rc=$?"
tail -n 1 gpurun_out/c24_llama1.json | cut -c1-1600
timeout 300 python bench.py --workload bert-large --steps 12 --warmup 5 --baseline none > gpurun_out/c24_bert1.json 2> gpurun_out/c24_bert1.err; echo "bert1 rc=$?"
tail -n 1 gpurun_out/c24_bert1.json | cut -c1-900
