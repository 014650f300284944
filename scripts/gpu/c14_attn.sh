mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_fused_model_gpu.py tests/test_flash_attention_gpu.py -m gpu -x -q 2>&1 | tail -3
PB2_NO_PDL=1 timeout 200 python scripts/attn_probe.py 2>&1 | grep "^NB" | head -4
PB2_BENCH_EXTRAS_BUDGET_S=1 timeout 600 python bench.py --steps 10 --warmup 3 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('N=1', d['ms_per_step'], d['value'])"
