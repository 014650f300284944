"""cProfile of the host side of a few training steps (where do the Python microseconds go?)."""
import cProfile, copy, os, pstats, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pipe_b200.models import CONFIGS
from pipe_b200.models.trainer import LMTrainer
cfg = copy.copy(CONFIGS["cfg3"])
if len(sys.argv) > 1 and sys.argv[1] == "tiny":
    # same layer / micro-batch structure, tiny width: the GPU never backs up, so the profile is pure host cost
    cfg.emsize = cfg.nhid = 256; cfg.nhead = 4; cfg.ntokens = 2048
tr = LMTrainer(cfg, n_stages=1, impl="fused", checkpoint="never")
tok = torch.randint(0, cfg.ntokens, (cfg.batch, cfg.bptt), device="cuda")
tgt = torch.randint(0, cfg.ntokens, (cfg.batch * cfg.bptt,), device="cuda")
for _ in range(3):
    tr.step(tok, tgt)
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(3):
    tr.step(tok, tgt)
pr.disable()
torch.cuda.synchronize()
st = pstats.Stats(pr)
st.sort_stats("cumulative").print_stats(45)
st.sort_stats("tottime").print_stats(25)
