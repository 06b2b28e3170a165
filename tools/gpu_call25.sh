#!/bin/bash
# 2-GPU call: the whole GPU test suite on the final tree
mkdir -p gpurun_out
timeout 1100 python -m pytest tests -m gpu -q -x 2>&1 | tail -25 > gpurun_out/c25_pytest.log; echo "pytest rc=${PIPESTATUS[0]}"; tail -8 gpurun_out/c25_pytest.log
