mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gemm_gpu.py tests/test_chain_gpu.py tests/test_fused_model_gpu.py tests/test_kernels_gpu.py -m gpu -x -q 2>&1 | tail -3
timeout 300 python scripts/bench_gemm.py > gpurun_out/gemm_bench_r2_layout.txt 2>&1; cat gpurun_out/gemm_bench_r2_layout.txt | cut -c1-200
timeout 200 python scripts/stack_probe.py 2>&1 | tail -5
PB2_BENCH_EXTRAS_BUDGET_S=1 timeout 600 python bench.py --steps 10 --warmup 3 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('N=1', d['ms_per_step'], d['value'])"
