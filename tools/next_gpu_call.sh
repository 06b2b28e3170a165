#!/bin/bash
# First GPU call of a round, one command:  gpurun --gpus 2 --timeout 1500 -- 'bash tools/next_gpu_call.sh'
# Runs what could only be checked on CPU so far, cheapest first, and leaves every log under gpurun_out/.
mkdir -p gpurun_out
P="python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1"
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest -m gpu rc=$?"
timeout 200 $P --master-port 29741 tools/multigpu_check.py ulysses tp_replicated > gpurun_out/new_multigpu_groups.log 2>&1; echo "ulysses / tp_replicated rc=$?"
timeout 300 python bench.py --steps 4 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench n=1 rc=$?"
timeout 300 python bench.py --impl twin --workload maskrcnn --steps 4 --warmup 3 > gpurun_out/twin_maskrcnn.json 2> gpurun_out/twin_maskrcnn.err; echo "twin maskrcnn rc=$?"
timeout 300 python bench.py --workload maskrcnn --steps 4 --warmup 3 > gpurun_out/bench_maskrcnn.json 2> gpurun_out/bench_maskrcnn.err; echo "bench maskrcnn rc=$?"
timeout 400 $P --master-port 29742 bench.py --gpus 2 --impl twin --workload gpt-moe-1.3b --steps 3 --warmup 3 > gpurun_out/twin_moe_n2.json 2> gpurun_out/twin_moe_n2.err; echo "twin moe n=2 rc=$?"
timeout 400 $P --master-port 29743 bench.py --gpus 2 --workload gpt-moe-1.3b --steps 3 --warmup 3 > gpurun_out/bench_moe_n2.json 2> gpurun_out/bench_moe_n2.err; echo "bench moe n=2 rc=$?"
tail -n 3 gpurun_out/pytest_gpu.log; grep -h '"ok"' gpurun_out/new_multigpu_groups.log | cut -c1-300; cat gpurun_out/*.json | cut -c1-400
