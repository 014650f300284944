"""BASELINE config 3 sweep: checkpoint modes x chunks on an n-stage single-process Pipe —
tokens/s, recompute overhead vs 'never', ideal and *measured* bubble (cell timeline), peak memory.

    python scripts/bench_modes.py --stages 2 [--impl fused] [--chunks 8 16 32]
"""
import argparse
import copy
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from pipe_b200 import Timeline
from pipe_b200.models import CONFIGS
from pipe_b200.models.trainer import LMTrainer
from pipe_b200.runtime.schedule import bubble_fraction

ap = argparse.ArgumentParser()
ap.add_argument("--stages", type=int, default=2)
ap.add_argument("--impl", default="fused")
ap.add_argument("--config", default="cfg3")
ap.add_argument("--chunks", type=int, nargs="+", default=[8, 16, 32])
ap.add_argument("--steps", type=int, default=6)
ap.add_argument("--out", default="gpurun_out/modes_bench.txt")
args = ap.parse_args()

lines = []
for chunks in args.chunks:
    base = None
    for mode in ("never", "except_last", "always"):
        cfg = copy.copy(CONFIGS[args.config])
        cfg.chunks = chunks
        tr = LMTrainer(cfg, n_stages=args.stages, impl=args.impl, checkpoint=mode, chunks=chunks)
        tok = torch.randint(0, cfg.ntokens, (cfg.batch, cfg.bptt), device=tr.first_device)
        tgt_dev = tr.first_device if args.impl == "fused" else tr.last_device
        tgt = torch.randint(0, cfg.ntokens, (cfg.batch * cfg.bptt,), device=tgt_dev)
        for _ in range(3):
            tr.step(tok, tgt)
        torch.cuda.synchronize()
        for d in range(args.stages):
            torch.cuda.reset_peak_memory_stats(d)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            tr.step(tok, tgt)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.steps
        # one instrumented step for the measured bubble
        tl = Timeline()
        tr.pipe.timeline = tl
        for j, d in enumerate(tr.pipe.devices):
            with torch.cuda.device(d):
                tl.origin(j, d)
        tr.step(tok, tgt)
        summ = tl.summary(args.stages)
        tr.pipe.timeline = None
        m = len(torch.empty(cfg.batch).chunk(chunks))
        peak = [torch.cuda.max_memory_allocated(d) / 2**20 for d in range(args.stages)]
        if mode == "never":
            base = ms
        lines.append(f"{args.impl} n={args.stages} chunks={chunks} (m={m}) {mode:11s}: {ms:8.2f} ms/step  "
                     f"{cfg.tokens_per_step / ms * 1e3:9.0f} tokens/s  recompute overhead {100 * (ms / base - 1):+5.1f}%  "
                     f"bubble ideal {bubble_fraction(m, args.stages):.3f} measured fwd {summ.get('fwd_bubble', float('nan')):.3f} "
                     f"bwd {summ.get('bwd_bubble', float('nan')):.3f}  peak MiB/stage {[int(p) for p in peak]}")
        print(lines[-1], flush=True)
        tr.pipe.close()
        del tr
        torch.cuda.empty_cache()
os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
open(args.out, "a").write("\n".join(lines) + "\n")
