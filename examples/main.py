#!/usr/bin/env python
"""The reference's example app on pipe_b200: the tutorial Transformer LM in a 2-stage pipeline,
8 training steps inside a torch.profiler schedule (wait 1, warm-up 1, active 2), optional CUDA memory
snapshot — reference ``main.py:115-271`` — minus torchtext (synthetic Zipfian tokens, or ``--data``).

    python examples/main.py never|except_last|always [--impl eager|fused] [--gpus 2] [--small]
        [--logdir log/] [--memory-snapshot snapshot.pickle] [--data corpus.txt]

Without two GPUs it runs on two CPU partitions (the plumbing-only BASELINE config 1).
"""
import argparse
import math
import os
import sys
import time
from pickle import dump

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.nn as nn

from pipe_b200 import Pipe
from pipe_b200.models import TransformerConfig, build_stages, count_parameters
from pipe_b200.utils.data import batchify, build_corpus, get_batch, synthetic_corpus


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("checkpoint", choices=["never", "except_last", "always"])
    ap.add_argument("--impl", default="eager", choices=["eager", "fused"])
    ap.add_argument("--gpus", type=int, default=2)
    ap.add_argument("--small", action="store_true", help="emsize 64, 2 layers (CPU-friendly)")
    ap.add_argument("--logdir", default=None, help="write a TensorBoard trace here (torch.profiler)")
    ap.add_argument("--memory-snapshot", default=None)
    ap.add_argument("--data", default=None, help="text file to tokenise instead of synthetic tokens")
    ap.add_argument("--steps", type=int, default=8)
    args = ap.parse_args()

    on_gpu = torch.cuda.is_available() and torch.cuda.device_count() >= args.gpus
    devices = list(range(args.gpus)) if on_gpu else ["cpu"] * args.gpus
    if args.impl == "fused" and not on_gpu:
        print("no CUDA devices: the fused modules run their plain-PyTorch reference path")

    if args.data:
        stream, vocab = build_corpus(args.data)
        ntokens = len(vocab)
    else:
        ntokens = 28782                                          # WikiText-2 vocabulary size
        stream = synthetic_corpus(40_000 if args.small or not on_gpu else 200_000, ntokens)

    # reference main.py:101-120,162
    batch_size, bptt, chunks = 32, 128, 4
    cfg = TransformerConfig(ntokens=ntokens, emsize=2048, nhid=2048, nlayers=16, nhead=32, dropout=0.2,
                            bptt=bptt, batch=batch_size, chunks=chunks)
    if args.small or not on_gpu:
        cfg.emsize = cfg.nhid = 64
        cfg.nlayers, cfg.nhead = 2, 2
    first = torch.device("cuda", 0) if on_gpu else torch.device("cpu")
    last = torch.device("cuda", args.gpus - 1) if on_gpu else torch.device("cpu")
    train_data = batchify(stream, batch_size, first)

    model = Pipe(build_stages(cfg, devices, impl=args.impl), chunks=chunks, checkpoint=args.checkpoint)
    print("Total parameters in model: {:,}".format(count_parameters(model)))

    criterion = nn.CrossEntropyLoss()
    if args.impl == "fused":
        from pipe_b200.ops.optim import FusedClipAdam
        optimizer = FusedClipAdam(model, lr=1e-3, max_norm=0.5)
    else:
        optimizer = torch.optim.Adam(model.parameters(), lr=1e-3)   # the reference's lr = 5.0 diverges at once

    def one_step(i: int) -> float:
        data, targets = get_batch(train_data, i, bptt)
        optimizer.zero_grad()
        output = model(data.contiguous()).local_value()
        loss = criterion(output.reshape(-1, ntokens).float(), targets.to(last))
        loss.backward()
        if args.impl != "fused":
            torch.nn.utils.clip_grad_norm_(model.parameters(), 0.5)
        optimizer.step()
        return float(loss)

    model.train()
    nbatches = min(args.steps * bptt, train_data.size(0) - 1)
    if args.memory_snapshot and on_gpu:
        torch.cuda.memory._record_memory_history(enabled="all")
    t0 = time.time()
    if args.logdir:
        schedule = torch.profiler.schedule(wait=1, warmup=1, active=2, repeat=1)
        handler = torch.profiler.tensorboard_trace_handler(os.path.join(args.logdir, args.checkpoint))
        model.trace_cells(True)                                   # chunk{i}-part{j} ranges in the trace
        with torch.profiler.profile(schedule=schedule, on_trace_ready=handler, record_shapes=True,
                                    profile_memory=True, with_stack=True) as prof:
            for batch, i in enumerate(range(0, nbatches, bptt)):
                prof.step()
                loss = one_step(i)
    else:
        for batch, i in enumerate(range(0, nbatches, bptt)):
            loss = one_step(i)
            print(f"| batch {batch:3d} | loss {loss:6.3f} | ppl {math.exp(min(loss, 20)):10.2f} | "
                  f"{(time.time() - t0) * 1e3 / (batch + 1):7.1f} ms/batch")
    if args.memory_snapshot and on_gpu:
        with open(args.memory_snapshot, "wb") as f:
            dump(torch.cuda.memory._snapshot(), f)
        torch.cuda.memory._record_memory_history(enabled=None)
    if on_gpu:
        for d in range(args.gpus):
            print(f"cuda:{d} peak allocated {torch.cuda.max_memory_allocated(d) / 2**20:.1f} MiB")
    model.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
