"""BASELINE.json config matrix on this box: chunks x checkpoint mode for one model / stage count.

    python scripts/sweep_cfg.py --config cfg3 --gpus 1
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 scripts/sweep_cfg.py --config cfg2 --gpus 2

Per (chunks, mode): ms/step (CUDA events, max over ranks, median of 3 regions), tokens/s, recompute overhead against
'never' at the same chunks, measured bubble per stage (N > 1, CUDA-event cell timelines) next to the ideal, peak MiB per
rank.  `strict` = except_last with strict_except_last=True (the real last micro-batch keeps its graph; the reference's
quirk checkpoints every micro-batch when batch / chunks leaves fewer micro-batches than chunks, README.md:398).
"""
import argparse
import copy
import os
import statistics
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from pipe_b200.models import CONFIGS
from pipe_b200.models.trainer import LMTrainer
from pipe_b200.utils.benchkit import timed_steps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="cfg3")
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--chunks", default="8,16,32")
    ap.add_argument("--steps", type=int, default=8)
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    if world > 1:
        torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
        dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
    cfg0 = CONFIGS[args.config]
    if rank == 0:
        print(f"# {args.config}: E {cfg0.emsize} layers {cfg0.nlayers} heads {cfg0.nhead} bptt {cfg0.bptt} batch {cfg0.batch}, "
              f"{args.gpus} stage(s), fused bf16, {args.steps} steps x 3 regions")
        print("chunks(m) | mode | ms/step | tokens/s | overhead vs never | ideal bubble | measured bubble per stage | peak MiB per rank")
    g = torch.Generator().manual_seed(1)
    for chunks in (int(c) for c in args.chunks.split(",")):
        cfg = copy.copy(cfg0)
        cfg.chunks = chunks
        m = len(torch.empty(cfg.batch).chunk(chunks))
        base = None
        for mode in ("never", "except_last", "strict", "always"):
            tr = LMTrainer(cfg, n_stages=args.gpus, impl="fused", checkpoint="except_last" if mode == "strict" else mode,
                           chunks=chunks)
            if mode == "strict":
                tr.pipe.strict_except_last = True
            tok = torch.randint(0, cfg.ntokens, (cfg.batch, cfg.bptt), generator=g).to(tr.first_device) if tr.is_first else None
            tgt = torch.randint(0, cfg.ntokens, (cfg.batch * cfg.bptt,), generator=g).to(tr.first_device) if tr.is_first else None
            for _ in range(3):
                tr.step(tok, tgt)
            torch.cuda.reset_peak_memory_stats()
            ms = statistics.median(timed_steps(lambda k: tr.step(tok, tgt), args.steps, tr.device) for _ in range(3)) / args.steps
            peak = torch.cuda.max_memory_allocated() / 2**20
            bubble = "-"
            if world > 1:
                tr.pipe.timeline = True
                for _ in range(2):
                    tr.step(tok, tgt)
                summ = tr.pipe.timeline_summary()
                bubble = " ".join(f"{s['bubble']:.3f}" for s in summ)
                peaks = [None] * world
                dist.all_gather_object(peaks, round(peak))
                peak_s = " / ".join(str(p) for p in peaks)
            else:
                peak_s = str(round(peak))
            if mode == "never":
                base = ms
            if rank == 0:
                print(f"{chunks} ({m}) | {mode} | {ms:.2f} | {cfg.tokens_per_step / ms * 1e3:.0f} | "
                      f"{(ms / base - 1) * 100:+.1f} % | {tr.bubble:.3f} | {bubble} | {peak_s}", flush=True)
            tr.close()
            del tr
            torch.cuda.empty_cache()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
