timeout 700 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29621 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/c8_bench_n8.json 2> gpurun_out/c8_bench_n8.err
cat gpurun_out/c8_bench_n8.json | cut -c1-6000
echo unbatched-warnings: $(grep -c "unbatched P2P" gpurun_out/c8_bench_n8.err)
grep -v "^\[W\|Warning\|OMP_NUM\|^\*\*\*\|^$\|unbatched" gpurun_out/c8_bench_n8.err | tail -10
