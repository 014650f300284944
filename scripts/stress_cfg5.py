"""BASELINE config 5 (deep stress): 32-layer emsize 8192, nhead 64 Transformer over 8 stages, chunks = 32
(-> 20 micro-batches), checkpoint='always', with an @skippable stash/pop pair that jumps from stage 0 to
stage 7.  Single-process Pipe (skip portals live in the in-process scheduler); fused kernels.

    python scripts/stress_cfg5.py [--stages 8] [--layers 32] [--steps 3]
"""
import argparse
import copy
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch import nn

from pipe_b200 import Pipe
from pipe_b200.models import CONFIGS, build_layer, count_parameters
from pipe_b200.models.transformer import stage_balance
from pipe_b200.ops.optim import FusedClipAdam
from pipe_b200.runtime.schedule import bubble_fraction
from pipe_b200.skip import pop, skippable, stash

ap = argparse.ArgumentParser()
ap.add_argument("--stages", type=int, default=8)
ap.add_argument("--layers", type=int, default=32)
ap.add_argument("--steps", type=int, default=3)
ap.add_argument("--emsize", type=int, default=8192)
ap.add_argument("--out", default="gpurun_out/stress_cfg5.txt")
args = ap.parse_args()


@skippable(stash=["emb"])
class StashEmb(nn.Module):
    def forward(self, x, tgt):
        yield stash("emb", x)
        return x, tgt


@skippable(pop=["emb"])
class PopAdd(nn.Module):
    def forward(self, x, tgt):
        emb = yield pop("emb")
        return x + 0.1 * emb, tgt


cfg = copy.copy(CONFIGS["cfg5"])
cfg.nlayers, cfg.emsize, cfg.nhid = args.layers, args.emsize, args.emsize
cfg.nhead = max(1, args.emsize // 128)
balance = stage_balance(cfg.nlayers, args.stages)
groups, at = [], 0
for j, count in enumerate(balance):
    layers = [build_layer(cfg, k, "fused", j) for k in range(at, at + count)]
    if j == 0:
        layers.insert(1, StashEmb().to(f"cuda:{j}"))
    if j == len(balance) - 1:
        layers.insert(len(layers) - 1, PopAdd().to(f"cuda:{j}"))
    groups.append(nn.Sequential(*layers))
    at += count
model = nn.Sequential(*groups)
pipe = Pipe(model, chunks=cfg.chunks, checkpoint="always")
opt = FusedClipAdam(pipe, lr=1e-4, max_norm=0.5)
params = count_parameters(pipe)
tok = torch.randint(0, cfg.ntokens, (cfg.batch, cfg.bptt), device="cuda:0")
tgt = torch.randint(0, cfg.ntokens, (cfg.batch, cfg.bptt), device="cuda:0")


def step():
    opt.zero_grad()
    per_seq = pipe(tok, tgt).local_value()
    loss = per_seq.sum() / (cfg.batch * cfg.bptt)
    loss.backward()
    opt.step()
    return loss


losses = [step().item() for _ in range(2)]
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
with torch.cuda.device(0):
    e0.record()
    for _ in range(args.steps):
        losses.append(step().item())
    e1.record()
for d in range(args.stages):
    torch.cuda.synchronize(d)
ms = e0.elapsed_time(e1) / args.steps
m = len(torch.empty(cfg.batch).chunk(cfg.chunks))
peak = [int(torch.cuda.max_memory_allocated(d) / 2**20) for d in range(args.stages)]
line = (f"cfg5 stress: {params / 1e9:.2f} G params, E={cfg.emsize} L={cfg.nlayers} heads={cfg.nhead}, {args.stages} stages, "
        f"chunks={cfg.chunks} (m={m}), checkpoint=always, cross-stage skip 0->{args.stages - 1}: {ms:.1f} ms/step, "
        f"{cfg.tokens_per_step / ms * 1e3:.0f} tokens/s, ideal bubble {bubble_fraction(m, args.stages):.3f}, "
        f"losses {['%.3f' % l for l in losses]}, peak MiB/stage {peak}")
print(line, flush=True)
os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
open(args.out, "a").write(line + "\n")
assert all(l == l for l in losses), "NaN loss"
pipe.close()
