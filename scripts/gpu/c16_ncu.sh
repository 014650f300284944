#!/bin/bash
# refresh the ncu captures of the kernels that changed late in round 2 (plain run first, one GPU, no clock control)
mkdir -p gpurun_out/ncu
python scripts/one_step.py cfg3 2 > gpurun_out/ncu/plain_cfg3.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/ncu/plain_cfg3.log; exit 1; }
cap() {   # name regex config skip
  timeout 240 ncu --set full --clock-control none --import-source on -k regex:$2 -s $4 -c 1 -f -o gpurun_out/ncu/$1 \
      python scripts/one_step.py $3 2 > gpurun_out/ncu/$1.log 2>&1
  echo "$1 rc=$?"
}
cap gemm_skinny gemm_bf16_tn_kernel cfg3 40
cap attn_tile_fwd attn_tc_fwd_kernel cfg3 20
cap attn_tile_bwd attn_tc_bwd_kernel cfg3 20
UPTO=5 PB2_CHAIN_PREFETCH=0 python scripts/chain_probe.py > gpurun_out/ncu/plain_chain.log 2>&1 || { echo "plain chain run failed"; exit 1; }
UPTO=5 PB2_CHAIN_PREFETCH=0 timeout 240 ncu --set full --clock-control none --import-source on -k regex:chain_kernel -s 12 -c 1 -f -o gpurun_out/ncu/chain_kernel \
    python scripts/chain_probe.py > gpurun_out/ncu/chain_kernel.log 2>&1
echo "chain_kernel rc=$?"
ls -la gpurun_out/ncu/*.ncu-rep | wc -l
