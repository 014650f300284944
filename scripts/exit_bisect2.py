import copy, sys, torch
sys.path.insert(0, ".")
from pipe_b200.models import CONFIGS
from pipe_b200.models.trainer import LMTrainer
mode, steps, E, pin = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
cfg = copy.copy(CONFIGS["cfg3"]); cfg.nlayers = 2; cfg.emsize = cfg.nhid = E; cfg.nhead = max(4, E // 256)
tr = LMTrainer(cfg, n_stages=1, impl="fused", checkpoint=mode)
tok = torch.randint(0, cfg.ntokens, (cfg.batch, cfg.bptt)); tgt = torch.randint(0, cfg.ntokens, (cfg.batch * cfg.bptt,))
if pin: tok, tgt = tok.pin_memory(), tgt.pin_memory()
for _ in range(steps):
    loss = tr.step(tok.cuda(non_blocking=True), tgt.cuda(non_blocking=True))
torch.cuda.synchronize(); print("ok", mode, steps, E, pin, float(loss))
