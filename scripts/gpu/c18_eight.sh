# N = 8 bench line with the final kernels (eager / in-process Pipe / hybrid extras off: 8-GPU time is nearly spent)
mkdir -p gpurun_out
PB2_BENCH_NO_EAGER=1 PB2_BENCH_NO_PIPE=1 PB2_BENCH_NO_HYBRID=1 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29591 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/c18_bench_n8.json 2> gpurun_out/c18_bench_n8.err
cut -c1-4500 gpurun_out/c18_bench_n8.json
grep -v "^\[W\|Warning\|OMP_NUM\|^\*\*\*\|^$" gpurun_out/c18_bench_n8.err | tail -5
