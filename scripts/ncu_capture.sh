#!/bin/bash
# ncu capture of the GEMM kernel (B200_PROFILING.md recipe): plain run first, then --set full on 3 launches.
# Usage (on the GPU box): bash scripts/ncu_capture.sh skinny|wgrad
set -u
cd "$(dirname "$0")/.."
which=${1:-skinny}
mkdir -p gpurun_out
python scripts/one_gemm.py $which > gpurun_out/plain_$which.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_bf16 -s 6 -c 3 -f \
    -o gpurun_out/gemm_$which python scripts/one_gemm.py $which > gpurun_out/ncu_$which.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_$which.log
