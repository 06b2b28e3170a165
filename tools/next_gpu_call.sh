#!/bin/bash
# First GPU call of a round, one command:  gpurun --gpus 2 --timeout 1500 -- 'bash tools/next_gpu_call.sh'
# Re-establishes the measured state of the tree, cheapest first, and leaves every log under gpurun_out/.
mkdir -p gpurun_out
P="python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1"
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest -m gpu rc=$?"; tail -n 3 gpurun_out/pytest_gpu.log
bash tools/race_check.sh 2 1 > gpurun_out/race.log 2>&1; echo "race-delay build rc=$?"; tail -n 3 gpurun_out/race.log
timeout 400 python bench.py --steps 4 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench llama n=1 (+twin) rc=$?"
timeout 300 python bench.py --workload maskrcnn --steps 10 --warmup 5 > gpurun_out/bench_maskrcnn.json 2> gpurun_out/bench_maskrcnn.err; echo "bench maskrcnn (+twin) rc=$?"
timeout 500 $P --master-port 29742 bench.py --gpus 2 --steps 4 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench llama tp2 (+nccl arm) rc=$?"
timeout 500 $P --master-port 29743 bench.py --gpus 2 --workload gpt-moe-1.3b --steps 8 --warmup 8 > gpurun_out/bench_moe_n2.json 2> gpurun_out/bench_moe_n2.err; echo "bench moe n=2 (+twin) rc=$?"
for f in gpurun_out/bench_*.json; do tail -n 1 "$f" | cut -c1-500; done
