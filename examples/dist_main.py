#!/usr/bin/env python
"""The tutorial Transformer LM with one process per pipeline stage (and, optionally, several pipelines
side by side): what the reference leaves as future work (``pipe.py:295-302``) and what it only remarks on
for data parallelism (``pipe.py:290-293``).

    # 4 stages on 4 GPUs (NCCL) or 4 CPU processes (gloo)
    python -m torch.distributed.run --nproc-per-node 4 --master-addr 127.0.0.1 examples/dist_main.py except_last

    # 2 pipelines x 4 stages on 8 GPUs; gradients averaged by the fused NVLink all-reduce
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 examples/dist_main.py never --stages 4

    options: --impl eager|fused   --small   --steps N   --transport nccl|mailbox   --reduce fused|nccl

Rank layout: rank = replica * stages + stage.  Every rank builds only its own stage (same seed on every
replica), the first stage of each pipeline feeds its shard of the mini-batch, the last one computes the loss.
"""
import argparse
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.distributed as dist
import torch.nn as nn

from pipe_b200.models import TransformerConfig, build_stage
from pipe_b200.parallel import HybridPipe
from pipe_b200.utils.data import batchify, get_batch, synthetic_corpus


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("checkpoint", choices=["never", "except_last", "always"])
    ap.add_argument("--stages", type=int, default=None, help="stages per pipeline (default: world size = one pipeline)")
    ap.add_argument("--impl", default="eager", choices=["eager", "fused"])
    ap.add_argument("--small", action="store_true", help="emsize 64, 2 layers per stage pair (CPU-friendly)")
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--transport", default="nccl", choices=["nccl", "mailbox", "auto"])
    ap.add_argument("--reduce", default="fused", choices=["fused", "nccl"])
    args = ap.parse_args()

    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", rank))
    on_gpu = torch.cuda.is_available() and torch.cuda.device_count() > local
    if on_gpu:
        torch.cuda.set_device(local)
        device = torch.device("cuda", local)
        dist.init_process_group("nccl", device_id=device)
    else:
        if args.impl == "fused":
            raise SystemExit("--impl fused needs CUDA devices")
        device = torch.device("cpu")
        dist.init_process_group("gloo")
    stages = args.stages or world
    if world % stages:
        raise SystemExit(f"world size {world} is not a multiple of --stages {stages}")
    replicas = world // stages
    stage_idx, replica = rank % stages, rank // stages

    ntokens, batch_size, bptt, chunks = 28782, 32, 128, 4             # reference main.py:101-120,162
    cfg = TransformerConfig(ntokens=ntokens, emsize=2048, nhid=2048, nlayers=16, nhead=32, dropout=0.2,
                            bptt=bptt, batch=batch_size, chunks=chunks)
    if args.small or not on_gpu:
        cfg.ntokens = ntokens = 512
        cfg.emsize = cfg.nhid = 64
        cfg.nlayers, cfg.nhead, cfg.bptt, cfg.batch = 2 * stages, 2, 16, 8
        batch_size, bptt = cfg.batch, cfg.bptt

    body = build_stage(cfg, stage_idx, stages, device, impl=args.impl, seed=0)      # same weights on every replica
    pipe = HybridPipe(body, stages=stages, chunks=chunks, checkpoint=args.checkpoint, device=device,
                      transport=args.transport)
    if args.impl == "fused":
        opt = pipe.make_optimizer(reduce_backend=args.reduce, lr=1e-3, max_norm=0.5)
    else:
        opt = torch.optim.Adam(body.parameters(), lr=1e-3)
    criterion = nn.CrossEntropyLoss()

    # every replica reads its own shard of one synthetic token stream
    stream = synthetic_corpus(20_000 if args.small or not on_gpu else 200_000, ntokens, seed=0)
    data_all = batchify(stream, batch_size * replicas, device)
    shard = data_all[:, replica * batch_size:(replica + 1) * batch_size]

    losses = []
    t0 = time.perf_counter()
    for step in range(args.steps):
        data, targets = get_batch(shard, step * bptt, bptt)
        opt.zero_grad()
        if args.impl == "fused":
            # fused LM head: targets ride through the pipe, the last stage returns per-sequence loss sums
            out = pipe(data.contiguous(), targets.view(-1, data.size(0)).t().contiguous()) if pipe.is_first else pipe()
            loss = out.local_value().sum() / data.numel() if pipe.is_last else None
        else:
            out = pipe(data.contiguous()) if pipe.is_first else pipe()
            # eager bodies: ship the targets to the last stage of this pipeline, then the usual criterion
            tgt = targets
            if stages > 1:
                first_rank, last_rank = replica * stages, replica * stages + stages - 1
                if pipe.is_first:
                    dist.send(targets.contiguous(), last_rank)
                if pipe.is_last:
                    tgt = torch.empty_like(targets)
                    dist.recv(tgt, first_rank)
            loss = criterion(out.local_value().reshape(-1, ntokens).float(), tgt) if pipe.is_last else None
        pipe.backward(loss)
        if args.impl != "fused":
            if replicas > 1:
                for p in body.parameters():                       # plain DDP-style averaging for the eager arm
                    if p.grad is not None:
                        dist.all_reduce(p.grad, group=pipe.replica_group)
                        p.grad.div_(replicas)
            torch.nn.utils.clip_grad_norm_(body.parameters(), 0.5)
        opt.step()
        if pipe.is_last:
            losses.append(float(loss.detach()))
    if on_gpu:
        torch.cuda.synchronize()
    dist.barrier()
    if pipe.is_last and replica == 0:
        dt = (time.perf_counter() - t0) / args.steps
        print(f"{replicas} pipeline(s) x {stages} stages, checkpoint={args.checkpoint}, impl={args.impl}: "
              f"loss {losses[0]:.3f} -> {losses[-1]:.3f}, {dt * 1e3:.1f} ms/step (wall, incl. first-step set-up), "
              f"ideal bubble {pipe.last_bubble:.3f}")
    dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
