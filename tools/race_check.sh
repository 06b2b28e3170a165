#!/bin/bash
# Race-hunting run of the cross-GPU protocols: a variant of the CUDA library built with
# -DB2F_RACE_DELAYS (pseudo-random pauses at every hand-rolled synchronisation point, see
# csrc/common/ptx.cuh::race_delay) next to the product libraries, then the multi-GPU protocol checks on it.
# usage (on a box with >= 2 GPUs):  tools/race_check.sh [nproc] [rounds]
set -u
NP=${1:-2}; ROUNDS=${2:-2}
cd "$(dirname "$0")/.."
python -c "from b200flow import build; build.build_cuda(variant='race', variant_flags=['-DB2F_RACE_DELAYS'])" > /dev/null || exit 2
rc=0
for i in $(seq 1 "$ROUNDS"); do
  B2F_CUDA_LIB=libb200flow_cuda_race.so timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$NP" \
    --master-addr 127.0.0.1 --master-port $((29800 + i)) tools/multigpu_check.py arena allreduce zero1 fused_gemm \
    tp_replicated moe_ep moe_k5 tp 2>&1 | grep "summary\|\"ok\": false\|timed out" | head -8
  [ "${PIPESTATUS[0]}" -ne 0 ] && rc=1
done
echo "race_check rc=$rc"
exit $rc
