"""Does the bf16 fused path train like the reference's fp32 arithmetic?  Same model, same initial weights, same data,
same optimizer (clip 0.5 + Adam), dropout off: the loss curves of impl='fused' (bf16 kernels, fp32 master weights) and
impl='eager' at fp32 / tf32 / bf16-autocast, side by side.  One GPU.

    python scripts/convergence_parity.py [steps]      -> table on stdout (profiles/convergence_parity_r2.txt)
"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pipe_b200.models.trainer import LMTrainer
from pipe_b200.models.transformer import TransformerConfig
from pipe_b200.utils.data import batchify, get_batch, synthetic_corpus

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 300
dev = torch.device("cuda:0")
cfg = TransformerConfig(ntokens=4096, emsize=512, nhid=1024, nlayers=4, nhead=8, dropout=0.0, bptt=25, batch=20, chunks=8)
# a stream with structure to learn: Zipf unigrams, and every other token is a fixed function of its predecessor
stream = synthetic_corpus(steps * cfg.bptt * cfg.batch + cfg.batch * 2, cfg.ntokens, seed=1)
stream[1::2] = (stream[0::2][: stream[1::2].numel()] * 7 + 3) % cfg.ntokens
data = batchify(stream, cfg.batch, dev)

def run(impl, precision):
    tr = LMTrainer(cfg, n_stages=1, impl=impl, precision=precision, checkpoint="never", lr=3e-4, seed=0)
    # the arms nest their stages differently (autocast wrapper, flattened Sequential) but enumerate the same tensors in
    # the same order: initial values are carried over by position, shapes checked
    sd = tr.pipe.state_dict()
    if run.init is None:
        run.init = [v.detach().clone() for v in sd.values()]
    else:
        assert len(sd) == len(run.init) and all(a.shape == b.shape for a, b in zip(sd.values(), run.init))
        tr.pipe.load_state_dict({k: v for k, v in zip(sd.keys(), run.init)})
    losses = []
    for s in range(steps):
        tokens, targets = get_batch(data, s * cfg.bptt, cfg.bptt)
        # batch-major targets (the model output is batch-major): the learnable pairing stays aligned
        targets = data[s * cfg.bptt + 1: s * cfg.bptt + 1 + cfg.bptt].t().reshape(-1)
        loss = tr.step(tokens.contiguous(), targets.contiguous())
        losses.append(float(loss.detach()))
    if hasattr(tr, "close"):
        tr.close()
    return losses
run.init = None

arms = [("eager", "fp32"), ("eager", "tf32"), ("eager", "bf16"), ("fused", "bf16")]
curves = {}
for impl, prec in arms:
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    curves[(impl, prec)] = run(impl, prec)
ref = curves[("eager", "fp32")]
print(f"# {cfg}")
print(f"# {steps} steps, lr 3e-4, clip 0.5, dropout 0; loss averaged over windows of 20 steps")
print("steps      " + "  ".join(f"{i}/{p:>5s}" for i, p in arms) + "   fused-bf16 vs eager-fp32")
for a in range(0, steps, 20):
    row = [sum(curves[k][a:a + 20]) / len(curves[k][a:a + 20]) for k in arms]
    print(f"{a:4d}-{a + 19:<4d}  " + "  ".join(f"{v:11.4f}" for v in row) + f"   {100 * (row[3] - row[0]) / row[0]:+.2f} %")
last = [sum(curves[k][-20:]) / 20 for k in arms]
print("final-window loss relative to eager fp32: " + ", ".join(f"{i}/{p} {100 * (v - last[0]) / last[0]:+.2f} %" for (i, p), v in zip(arms, last)))
