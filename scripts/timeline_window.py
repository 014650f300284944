"""Kernel timeline of one training step: start offsets, durations and gaps for a window of kernels."""
import copy, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import ProfilerActivity, profile
from pipe_b200.models import CONFIGS
from pipe_b200.models.trainer import LMTrainer
cfg = copy.copy(CONFIGS["cfg3"]); cfg.nlayers = 4
tr = LMTrainer(cfg, n_stages=1, impl="fused", checkpoint="never")
tok = torch.randint(0, cfg.ntokens, (cfg.batch, cfg.bptt), device="cuda")
tgt = torch.randint(0, cfg.ntokens, (cfg.batch * cfg.bptt,), device="cuda")
for _ in range(3): tr.step(tok, tgt)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    tr.step(tok, tgt); torch.cuda.synchronize()
ev = sorted([e for e in prof.events() if e.device_time > 0], key=lambda e: e.time_range.start)
t0 = ev[0].time_range.start
start = int(sys.argv[1]) if len(sys.argv) > 1 else 0
prev_end = None
for e in ev[start:start + 70]:
    s, d = e.time_range.start - t0, e.device_time
    gap = (e.time_range.start - prev_end) if prev_end is not None else 0
    print(f"{s:10.1f} us  dur {d:7.1f}  gap {gap:7.1f}  {e.name[:60]}")
    prev_end = e.time_range.end
print("total span", ev[-1].time_range.end - t0, "us; kernels", len(ev))
# ---- idle-gap summary over the whole step ----
gaps = []
busy_end = ev[0].time_range.end
for a, b in zip(ev, ev[1:]):
    busy_end = max(busy_end, a.time_range.end)
    g = b.time_range.start - busy_end
    if g > 5:
        gaps.append((g, a.name[:38], b.name[:38], b.time_range.start - t0))
tot = sum(g for g, *_ in gaps)
print(f"idle gaps > 5 us: {len(gaps)} totalling {tot:.0f} us of {ev[-1].time_range.end - t0:.0f} us")
for g, a, b, at in sorted(gaps, reverse=True)[:25]:
    print(f"  gap {g:7.1f} us at {at:9.1f}: after {a}  before {b}")
