"""Independent baseline: the same eager stage bodies driven by torch's own successor of the
reference, ``torch.distributed.pipelining.ScheduleGPipe`` (one rank per stage, NCCL p2p), against this repo's
``DistPipe`` with eager bodies and with the fused kernels — same model, same micro-batch count, same step.

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 scripts/bench_schedule_gpipe.py

ScheduleGPipe splits a mini-batch evenly, so the comparison runs 10 micro-batches of 2 rows (chunks = 16 gives the
same 10 micro-batches in the reference's `Tensor.chunk` arithmetic for batch 20)."""
import copy
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import torch.nn.functional as F
from torch.distributed.pipelining import PipelineStage, ScheduleGPipe

from pipe_b200.models import CONFIGS, build_stage
from pipe_b200.models.trainer import AutocastStage, LMTrainer, clip_grad_norm_
from pipe_b200.utils.benchkit import timed_steps


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=dev)
    cfg = copy.copy(CONFIGS[os.environ.get("CFG", "cfg3")])
    m = 10
    cfg.chunks = 16
    V = cfg.ntokens
    g = torch.Generator().manual_seed(0)
    tok = torch.randint(0, V, (cfg.batch, cfg.bptt), generator=g).to(dev)
    tgt = torch.randint(0, V, (cfg.batch, cfg.bptt), generator=g).to(dev)
    steps = 8
    res = {}

    # ---- torch.distributed.pipelining.ScheduleGPipe, eager bodies under bf16 autocast -------------------
    body = AutocastStage(build_stage(cfg, rank, world, dev, "eager", 0))
    stage = PipelineStage(body, rank, world, dev)

    def loss_fn(logits, target):
        return F.cross_entropy(logits.reshape(-1, V).float(), target.reshape(-1))

    sched = ScheduleGPipe(stage, n_microbatches=m, loss_fn=loss_fn)
    params = list(body.parameters())
    opt = torch.optim.Adam(params, lr=1e-4, fused=True)

    def step_gpipe(k):
        opt.zero_grad(set_to_none=True)
        if rank == 0:
            sched.step(tok)
        elif rank == world - 1:
            sched.step(target=tgt, losses=[])
        else:
            sched.step()
        clip_grad_norm_(params, 0.5, distributed=True)
        opt.step()

    for k in range(3):
        step_gpipe(k)
    res["ScheduleGPipe + eager bf16"] = timed_steps(step_gpipe, steps, dev) / steps
    del sched, stage, body, opt, params
    torch.cuda.empty_cache()

    # ---- this repo's DistPipe: eager bodies, then fused kernels ---------------------------------------------
    for impl in ("eager", "fused"):
        tr = LMTrainer(cfg, n_stages=world, impl=impl, precision="bf16", checkpoint="never", chunks=16)
        feed_tgt = tgt.reshape(-1)
        a = tok if tr.is_first else None
        b = (feed_tgt if tr.is_first else None) if impl == "fused" else (feed_tgt if tr.is_last else None)
        for k in range(3):
            tr.step(a, b)
        res[f"DistPipe + {impl} bf16"] = timed_steps(lambda k: tr.step(a, b), steps, dev) / steps
        tr.close()
        del tr
        torch.cuda.empty_cache()
    if rank == 0:
        print(f"# {os.environ.get('CFG', 'cfg3')}, {world} stages, 10 micro-batches of 2 rows x 25 tokens, full step (clip 0.5 + Adam), ms/step (max over ranks)")
        for k, v in res.items():
            print(f"{k}: {v:.2f} ms/step, {cfg.tokens_per_step / v * 1e3:.0f} tokens/s")
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
