#!/bin/bash
# op-chain kernel in the model: tests, then the single-GPU bench with and without it
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_chain_gpu.py tests/test_fused_model_gpu.py tests/test_gemm_gpu.py -m gpu -x -q > gpurun_out/c10_pytest.txt 2>&1
tail -5 gpurun_out/c10_pytest.txt
for c in 1 0; do
  PB2_CHAIN=$c PB2_BENCH_EXTRAS_BUDGET_S=1 timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/c10_bench_chain$c.json 2> gpurun_out/c10_bench_chain$c.err
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/c10_bench_chain$c.json").read().strip().splitlines()[-1])
    print("PB2_CHAIN=$c", d["ms_per_step"], d["value"], d.get("regions_ms"), d.get("gpu_launches"))
except Exception as e:
    print("PB2_CHAIN=$c failed", e); print(open("gpurun_out/c10_bench_chain$c.err").read()[-2000:])
PY
done
