#!/bin/bash
mkdir -p gpurun_out
timeout 120 python scripts/chain_debug.py > gpurun_out/chain_debug.txt 2>&1; cat gpurun_out/chain_debug.txt
PB2_CHAIN_PREFETCH=0 timeout 120 python scripts/chain_probe.py > gpurun_out/chain_probe.txt 2>&1
echo "rc=$?" >> gpurun_out/chain_probe.txt
grep -v "landed\|operands pre\|own chunks" gpurun_out/chain_probe.txt
