#!/bin/bash
# One `ncu --set full` capture per kernel of the training step (B200_PROFILING.md recipe: the plain run first, one GPU,
# --clock-control none), then a roofline table from the reports (scripts/ncu_table.py runs on the CPU box).
#   bash scripts/ncu_all.sh            (on the GPU box; reports land in gpurun_out/ncu/)
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out/ncu
python scripts/one_step.py cfg3 2 > gpurun_out/ncu/plain_cfg3.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/ncu/plain_cfg3.log; exit 1; }
python scripts/one_step.py shipped 2 > gpurun_out/ncu/plain_shipped.log 2>&1 || { echo "plain run failed"; exit 1; }
cap() {   # name regex config skip
  timeout 240 ncu --set full --clock-control none --import-source on -k regex:$2 -s $4 -c 1 -f -o gpurun_out/ncu/$1 \
      python scripts/one_step.py $3 2 > gpurun_out/ncu/$1.log 2>&1
  echo "$1 rc=$?"
}
cap gemm_skinny gemm_bf16_tn_kernel cfg3 40
cap gemm_persistent gemm_bf16_persistent_kernel cfg3 6
cap attn_tile_fwd attn_tc_fwd_kernel cfg3 20
cap attn_tile_bwd attn_tc_bwd_kernel cfg3 20
cap add_ln_fwd add_ln_fwd_kernel cfg3 20
cap add_ln_bwd add_ln_bwd_kernel cfg3 20
cap embed_fwd embed_fwd_kernel cfg3 8
cap embed_bwd embed_bwd_kernel cfg3 8
cap cross_entropy ce_fwd_bwd_kernel cfg3 10
cap adam adam_tiles_kernel cfg3 1
cap transpose_multi transpose_multi_kernel cfg3 20
cap flash_fwd flash_fwd_kernel shipped 10
cap flash_bwd flash_bwd_kernel shipped 10
ls -la gpurun_out/ncu/*.ncu-rep | wc -l
