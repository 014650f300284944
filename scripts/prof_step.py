"""Where does a training step spend its time?  Host enqueue time vs device time, and the
per-kernel device-time table (torch.profiler / CUPTI).  Usage:
    python scripts/prof_step.py [--layers 16] [--impl fused] [--out gpurun_out/prof_step.txt]"""
import argparse
import copy
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch

from pipe_b200.models import CONFIGS
from pipe_b200.models.trainer import LMTrainer

ap = argparse.ArgumentParser()
ap.add_argument("--layers", type=int, default=16)
ap.add_argument("--impl", default="fused")
ap.add_argument("--checkpoint", default="never")
ap.add_argument("--out", default="gpurun_out/prof_step.txt")
args = ap.parse_args()

cfg = copy.copy(CONFIGS["cfg3"])
cfg.nlayers = args.layers
tr = LMTrainer(cfg, n_stages=1, impl=args.impl, checkpoint=args.checkpoint)
tok = torch.randint(0, cfg.ntokens, (cfg.batch, cfg.bptt), device="cuda")
tgt = torch.randint(0, cfg.ntokens, (cfg.batch * cfg.bptt,), device="cuda")
for _ in range(3):
    tr.step(tok, tgt)
torch.cuda.synchronize()

lines = []
t0 = time.perf_counter()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
N = 5
for _ in range(N):
    tr.step(tok, tgt)
host = (time.perf_counter() - t0) / N * 1e3
e1.record()
torch.cuda.synchronize()
lines.append(f"impl={args.impl} layers={args.layers} ckpt={args.checkpoint}: host enqueue {host:.2f} ms/step, "
             f"device span {e0.elapsed_time(e1) / N:.2f} ms/step")

from torch.profiler import ProfilerActivity, profile
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    tr.step(tok, tgt)
    torch.cuda.synchronize()
ka = prof.key_averages()
rows = [(k.key, k.device_time_total / 1e3, k.count) for k in ka if k.device_time_total > 0 and k.device_type.name == "CUDA"]
rows.sort(key=lambda r: -r[1])
total = sum(r[1] for r in rows)
lines.append(f"sum of kernel device time: {total:.2f} ms/step over {sum(r[2] for r in rows)} launches")
for name, ms, cnt in rows[:30]:
    lines.append(f"{ms:9.3f} ms  {cnt:6d}x  {ms / cnt * 1e3:8.1f} us  {name[:110]}")
os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
open(args.out, "w").write("\n".join(lines) + "\n")
print("\n".join(lines))
