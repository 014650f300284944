mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/c15_pytest.txt 2>&1; tail -3 gpurun_out/c15_pytest.txt
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -1
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/c15_bench_n1.json 2> gpurun_out/c15_bench_n1.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/c15_bench_n1.json").read().strip().splitlines()[-1])
print({k: d[k] for k in ("ms_per_step", "value", "regions_ms", "gpu_launches", "step_ms_percentiles")})
print(d["e2e"], d.get("recompute"), d.get("eager_same_precision"))
PY
