#!/bin/bash
# Everything that needs more than one GPU, in one call on an 8-GPU box: the GPU test tier, the scaling bench at
# N = 2 / 4 / 8, the config-5 stress run and the hybrid (2 pipelines x 4 stages) bench.  Results land in gpurun_out/.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > gpurun_out/r8_pytest.txt
for n in 2 4 8; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500+n)) bench.py --gpus $n --steps 10 --warmup 3 2>&1 | tail -1 > gpurun_out/r8_bench_n$n.json
done
timeout 600 python scripts/stress_cfg5.py --steps 3 --out gpurun_out/r8_stress_cfg5.txt > gpurun_out/r8_stress.log 2>&1; echo "stress rc=$?" >> gpurun_out/r8_stress.log
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29611 scripts/bench_hybrid.py > gpurun_out/r8_hybrid.log 2>&1; echo "hybrid rc=$?" >> gpurun_out/r8_hybrid.log
cat gpurun_out/r8_pytest.txt; for n in 2 4 8; do cut -c1-260 gpurun_out/r8_bench_n$n.json; done; tail -5 gpurun_out/r8_stress.log; tail -6 gpurun_out/r8_hybrid.log
