"""BASELINE config 4: Pipe-inside-DDP hybrid, k replicas x n stages (default 2 x 4 on 8 GPUs), the
fused NVSwitch gradient all-reduce vs NCCL.  torchrun, one rank per GPU:

    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 scripts/bench_hybrid.py --stages 4

Every replica trains on its own shard (global batch = replicas x 20 sequences, weak scaling in the
replica dimension).  Reports tokens/s (device-timed, max over ranks), the ideal bubble, and the
time of the reduction step alone.
"""
import argparse
import copy
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from pipe_b200.models import CONFIGS, build_stage
from pipe_b200.parallel import HybridPipe
from pipe_b200.runtime.schedule import bubble_fraction
from pipe_b200.utils.benchkit import timed_steps

ap = argparse.ArgumentParser()
ap.add_argument("--stages", type=int, default=4)
ap.add_argument("--config", default="cfg3")
ap.add_argument("--steps", type=int, default=10)
ap.add_argument("--warmup", type=int, default=3)
ap.add_argument("--checkpoint", default="never")
ap.add_argument("--out", default="gpurun_out/hybrid_bench.txt")
args = ap.parse_args()

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ.get("LOCAL_RANK", rank))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
dev = torch.device("cuda", local)
cfg = copy.copy(CONFIGS[args.config])
results = {}
for backend in ("nccl", "fused"):
    body = build_stage(cfg, rank % args.stages, args.stages, dev, impl="fused", seed=0)
    pipe = HybridPipe(body, stages=args.stages, chunks=cfg.chunks, checkpoint=args.checkpoint, device=dev)
    opt = pipe.make_optimizer(reduce_backend=backend, lr=1e-4, max_norm=0.5)
    g = torch.Generator().manual_seed(100 + pipe.replica)
    tok = torch.randint(0, cfg.ntokens, (cfg.batch, cfg.bptt), generator=g).to(dev)
    tgt = torch.randint(0, cfg.ntokens, (cfg.batch, cfg.bptt), generator=g).to(dev)

    def step(_k):
        opt.zero_grad()
        out = pipe(tok, tgt) if pipe.is_first else pipe()
        loss = out.local_value().sum() / (cfg.batch * cfg.bptt) if pipe.is_last else None
        pipe.backward(loss)
        opt.step()

    for k in range(args.warmup):
        step(k)
    ms = timed_steps(step, args.steps, dev) / args.steps
    # the reduction alone
    (st,) = opt.devs.values()
    red = timed_steps(lambda _k: opt._reduce_replicas(), 5, dev) / 5
    results[backend] = (ms, red, st.symm.numel * 4, st.symm.mode, bool(st.symm.multicast_ptr))
    del opt, pipe, body
    torch.cuda.empty_cache()
if rank == 0:
    k = world // args.stages
    m = len(torch.empty(cfg.batch).chunk(cfg.chunks))
    lines = []
    for backend, (ms, red, nbytes, mode, mc) in results.items():
        toks = k * cfg.batch * cfg.bptt / (ms / 1e3)
        bus = 2 * (k - 1) / k * nbytes / red / 1e6
        lines.append(f"hybrid {k}x{args.stages} ({args.config}, ckpt={args.checkpoint}) reduce={backend:5s}: {ms:7.2f} ms/step  "
                     f"{toks:9.0f} tokens/s  bubble_ideal {bubble_fraction(m, args.stages):.3f}  "
                     f"grad all-reduce alone {red * 1e3:8.1f} us over {nbytes / 2**20:.0f} MiB ({bus:.0f} GB/s bus, mode={mode}, multicast={mc})")
    print("\n".join(lines), flush=True)
    os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
    open(args.out, "a").write("\n".join(lines) + "\n")
dist.barrier()
dist.destroy_process_group()
