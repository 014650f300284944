#!/bin/bash
# evidence for profiles/chain_kernel_r2.txt: op-chain kernel timeline, stack forward with / without it, chain tests
mkdir -p gpurun_out
{
  echo "# Op-chain kernel (csrc/chain_sm100.cu), config-3 shapes: 75 tokens, E = FF = 4096, weights rotated through > L2"
  echo "# ---- scripts/chain_probe.py (bit comparison with the separate kernels, CUDA-event timing, in-kernel globaltimer stamps)"
  PB2_CHAIN_PREFETCH=0 timeout 120 python scripts/chain_probe.py 2>&1
  echo "# ---- the same with the L2 look-ahead of the weight producer on (PB2_CHAIN_PREFETCH=16 k-blocks)"
  PB2_CHAIN_PREFETCH=16 timeout 120 python scripts/chain_probe.py 2>&1 | grep "chain of 6\|kernel span"
  echo "# ---- scripts/stack_probe.py: forward of a 16-layer stack, per layer"
  timeout 300 python scripts/stack_probe.py 2>&1
} > gpurun_out/chain_kernel_r2.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/c11_pytest.txt 2>&1
tail -4 gpurun_out/c11_pytest.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/c11_smoke.txt 2>&1; tail -2 gpurun_out/c11_smoke.txt
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/c11_bench_n1.json 2> gpurun_out/c11_bench_n1.err
tail -c 1500 gpurun_out/c11_bench_n1.json
