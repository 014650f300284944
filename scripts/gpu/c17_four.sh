# N = 4 bench line with the final kernels (eager / in-process Pipe / hybrid extras off to keep the 4-GPU call short)
mkdir -p gpurun_out
PB2_BENCH_NO_EAGER=1 PB2_BENCH_NO_PIPE=1 PB2_BENCH_NO_HYBRID=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29581 bench.py --gpus 4 --steps 20 --warmup 5 > gpurun_out/c17_bench_n4.json 2> gpurun_out/c17_bench_n4.err
cut -c1-4000 gpurun_out/c17_bench_n4.json
grep -v "^\[W\|Warning\|OMP_NUM\|^\*\*\*\|^$" gpurun_out/c17_bench_n4.err | tail -5
