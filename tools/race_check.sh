#!/bin/bash
# Race-hunting run of the cross-GPU protocols: rebuild the CUDA library with
# -DB2F_RACE_DELAYS (pseudo-random pauses at every hand-rolled synchronisation point, see
# csrc/common/ptx.cuh::race_delay), run the multi-GPU protocol checks a few times, rebuild normally.
# usage (on a box with >= 2 GPUs):  tools/race_check.sh [nproc] [rounds]
set -u
NP=${1:-2}; ROUNDS=${2:-2}
cd "$(dirname "$0")/.."
touch csrc/comm/*.cu csrc/kernels/moe.cu
B2F_NVCC_EXTRA=-DB2F_RACE_DELAYS python -m b200flow.build > /dev/null || exit 2
rc=0
for i in $(seq 1 "$ROUNDS"); do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$NP" --master-addr 127.0.0.1 \
    --master-port $((29800 + i)) tools/multigpu_check.py arena allreduce zero1 fused_gemm moe_ep tp 2>&1 \
    | grep "summary\|\"ok\": false\|timed out" | head -8
  [ "${PIPESTATUS[0]}" -ne 0 ] && rc=1
done
touch csrc/comm/*.cu csrc/kernels/moe.cu
python -m b200flow.build > /dev/null
echo "race_check rc=$rc"
exit $rc
