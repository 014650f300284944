"""Fused NVSwitch all-reduce vs NCCL: device-timed (CUDA events, max over ranks), bus bandwidth
= 2 (k-1)/k * bytes / time.  Launch with torchrun, one rank per GPU:
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 scripts/bench_allreduce.py
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from pipe_b200.parallel import SymmetricBuffer

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ.get("LOCAL_RANK", rank))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
dev = torch.device("cuda", local)
out_path = os.environ.get("PB2_OUT", "gpurun_out/allreduce_bench.txt")
lines = []


def timed(fn, iters):
    for _ in range(3):
        fn()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


for mib in (1, 16, 256, 1024, 1760):
    n = mib * 1024 * 1024 // 4
    buf = SymmetricBuffer(n, None, dev)
    ref = torch.randn(buf.numel, device=dev)
    iters = 20 if mib <= 256 else 5
    res = {}
    for ctas in (32, 64, 128):
        res[f"p2p/{ctas}"] = timed(lambda: buf.all_reduce_(scale=1.0 / world, ctas=ctas, use_multimem=False), iters)
        if buf.multicast_ptr:
            res[f"nvls/{ctas}"] = timed(lambda: buf.all_reduce_(scale=1.0 / world, ctas=ctas, use_multimem=True), iters)
    res["nccl"] = timed(lambda: dist.all_reduce(ref), iters)
    res["nccl+scale"] = timed(lambda: (dist.all_reduce(ref), ref.mul_(1.0 / world)), iters)
    if rank == 0:
        bus = lambda ms: 2 * (world - 1) / world * buf.numel * 4 / ms / 1e6
        line = f"world={world} mode={buf.mode} multicast={bool(buf.multicast_ptr)} {mib:5d} MiB: " + "  ".join(
            f"{k} {v * 1e3:8.1f} us ({bus(v):6.0f} GB/s)" for k, v in res.items())
        print(line, flush=True)
        lines.append(line)
    del buf
if rank == 0:
    os.makedirs(os.path.dirname(out_path) or ".", exist_ok=True)
    open(out_path, "a").write("\n".join(lines) + "\n")
dist.barrier()
dist.destroy_process_group()
