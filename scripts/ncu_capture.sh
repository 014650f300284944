#!/bin/bash
# ncu capture of the GEMM kernel (B200_PROFILING.md recipe): plain run first, then --set full on 3 launches.
# Usage (on the GPU box): bash scripts/ncu_capture.sh skinny|wgrad
set -u
cd "$(dirname "$0")/.."
which=${1:-skinny}
mkdir -p gpurun_out
python scripts/one_gemm.py $which > gpurun_out/plain_$which.log 2>&1 &&
skip=6; [ "$which" = "wgrad" ] && skip=2        # the wgrad workload launches 6 GEMMs, the skinny one 12
ncu --set full --clock-control none --import-source on -k regex:gemm_bf16 -s $skip -c 2 -f \
    -o gpurun_out/gemm_$which python scripts/one_gemm.py $which > gpurun_out/ncu_$which.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_$which.log
