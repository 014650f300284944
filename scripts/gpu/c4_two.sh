T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 500 $T --master-port 29541 scripts/sweep_cfg.py --config cfg2 --gpus 2 --chunks 8 2>&1 | grep -v "^\[W\|Warning\|OMP_NUM\|^\*\*\*\|^$" > gpurun_out/sweep_cfg2_n2.txt
timeout 700 $T --master-port 29542 scripts/sweep_cfg.py --config cfg3 --gpus 2 2>&1 | grep -v "^\[W\|Warning\|OMP_NUM\|^\*\*\*\|^$" > gpurun_out/sweep_cfg3_n2.txt
cat gpurun_out/sweep_cfg2_n2.txt gpurun_out/sweep_cfg3_n2.txt
timeout 600 python -m pytest tests/test_allreduce_multi_gpu.py -m gpu -v 2>&1 | grep -E "PASS|FAIL|SKIP|passed|failed|Error" > gpurun_out/multi_gpu_tests_2gpu.txt
cat gpurun_out/multi_gpu_tests_2gpu.txt
