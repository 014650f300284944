timeout 500 python -m pytest tests/test_allreduce_multi_gpu.py -m gpu -v 2>&1 | grep -E "PASS|FAIL|SKIP|passed|failed|Error" > gpurun_out/multi_gpu_tests_2gpu.txt; cat gpurun_out/multi_gpu_tests_2gpu.txt | tail -22
PB2_BENCH_NO_EAGER=1 PB2_BENCH_NO_PIPE=1 timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/c5_bench_n2.json 2> gpurun_out/c5_bench_n2.err
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/c5_bench_n2.json') if l.startswith('{')][-1])
print(d['ms_per_step'], {k:v for k,v in d.get('hybrid',{}).items() if k.endswith('_ms') or 'exposed' in k}, d.get('extras_error'))
PY
grep -v "^\[W\|Warning\|OMP_NUM\|^\*\*\*\|^$" gpurun_out/c5_bench_n2.err | tail -8
