"""A few training steps of a 2-layer slice of config 3 at full width (for ncu captures of every kernel of the step):
python scripts/one_step.py [config] [layers]"""
import copy
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from pipe_b200.models import CONFIGS
from pipe_b200.models.trainer import LMTrainer

cfg = copy.copy(CONFIGS[sys.argv[1] if len(sys.argv) > 1 else "cfg3"])
cfg.nlayers = int(sys.argv[2]) if len(sys.argv) > 2 else 2
tr = LMTrainer(cfg, n_stages=1, impl="fused", checkpoint="never")
tok = torch.randint(0, cfg.ntokens, (cfg.batch, cfg.bptt), device="cuda")
tgt = torch.randint(0, cfg.ntokens, (cfg.batch * cfg.bptt,), device="cuda")
for _ in range(3):
    loss = tr.step(tok, tgt)
torch.cuda.synchronize()
print("ok", float(loss))
