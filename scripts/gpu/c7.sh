timeout 300 python -m pytest tests/test_allreduce_multi_gpu.py -m gpu -q -k "skip_and_changing" 2>&1 | tail -12
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 500 $T --master-port 29543 scripts/sweep_cfg.py --config shipped --gpus 2 --chunks 4 2>&1 | grep -v "^\[W\|Warning\|OMP_NUM\|^\*\*\*\|^$" > gpurun_out/sweep_shipped_n2.txt; cat gpurun_out/sweep_shipped_n2.txt
