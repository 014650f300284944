# final 2-GPU validation: multi-GPU tests, then the N = 2 bench line with all extras
mkdir -p gpurun_out
timeout 500 python -m pytest tests/test_allreduce_multi_gpu.py -m gpu -v 2>&1 | grep -E "PASS|FAIL|SKIP|passed|failed|Error" > gpurun_out/multi_gpu_tests_2gpu_final.txt; tail -24 gpurun_out/multi_gpu_tests_2gpu_final.txt
timeout 700 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29571 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/c12_bench_n2.json 2> gpurun_out/c12_bench_n2.err
cut -c1-5000 gpurun_out/c12_bench_n2.json
grep -v "^\[W\|Warning\|OMP_NUM\|^\*\*\*\|^$" gpurun_out/c12_bench_n2.err | tail -8
