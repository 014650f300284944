"""True host-side cost of a training step: same layer / micro-batch structure as config 3 but with a
tiny width, so the GPU is never the bottleneck and wall time per step = host time per step."""
import copy, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pipe_b200.models import CONFIGS
from pipe_b200.models.trainer import LMTrainer
for impl in ("fused", "eager"):
    cfg = copy.copy(CONFIGS["cfg3"]); cfg.emsize = cfg.nhid = 256; cfg.nhead = 4; cfg.ntokens = 2048
    tr = LMTrainer(cfg, n_stages=1, impl=impl, checkpoint="never")
    tok = torch.randint(0, cfg.ntokens, (cfg.batch, cfg.bptt), device="cuda")
    tgt = torch.randint(0, cfg.ntokens, (cfg.batch * cfg.bptt,), device="cuda")
    for _ in range(5): tr.step(tok, tgt)
    torch.cuda.synchronize()
    t0 = time.perf_counter(); c0 = time.process_time()
    N = 10
    for _ in range(N): tr.step(tok, tgt)
    torch.cuda.synchronize()
    print(f"{impl}: wall {1e3 * (time.perf_counter() - t0) / N:.2f} ms/step, cpu {1e3 * (time.process_time() - c0) / N:.2f} ms/step (16 layers x 7 micro-batches, tiny width)")
    tr.pipe.close()
