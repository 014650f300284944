#!/usr/bin/env python
"""Headline benchmark: tokens/s of one full training step of the tutorial
Transformer LM in a synchronous pipeline (BASELINE.json config 3: E=4096,
d_hid=4096, 16 layers, 16 heads, bptt 25, batch 20, vocab 28782, chunks=8,
split over N stages = N GPUs; strong scaling).

    python bench.py --gpus 1 --steps 20 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 \
        --master-port 29501 bench.py --gpus 4 --steps 20 --warmup 5

Arms (``--impl``):
  fused      (default) this repo's sm_100a kernels, bf16 compute / fp32 master weights
  eager      this repo's scheduler with stock torch.nn stage bodies (cuBLAS + ATen):
             semantically the reference; the on-box baseline of BASELINE.md §3
  reference  the unmodified reference from baseline/_ref (not installable: prints why)

Prints ONE JSON line on rank 0.  ``value`` is device-timed (CUDA events, barrier +
synchronize on both sides, max over ranks) with inputs resident on the device: the
K-step region is timed three times back to back and the median is reported (all three
are in ``regions_ms``; a single 0.2 s region on 8 ranks moved by 25 % run to run).
``e2e`` repeats the measurement through the public API with a pinned-host -> device
copy of every step's tokens/targets and a device -> host read of the loss inside the
timed region.

Extra keys (fused arm; ``PB2_BENCH_MINIMAL=1`` skips them), each measured after the
headline so that it cannot disturb it:
  bubble            measured idle share per stage from CUDA-event cell timelines, next to the ideal
  recompute         ms/step with checkpoint='always' on the same model -> overhead vs the headline mode
  eager_same_precision   this scheduler with stock torch.nn bodies under bf16 autocast, same step
  pipe_single_process    (N > 1) the in-process ``Pipe`` over the same N GPUs, driven by rank 0 alone
  hybrid            (N >= 2) 2 pipelines x N/2 stages: step time per gradient-reduction backend
"""
from __future__ import annotations

import argparse
import copy
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "tokens/sec (device-timed, max over ranks)"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="fused", choices=["fused", "eager", "reference"])
    ap.add_argument("--config", default="cfg3")
    ap.add_argument("--chunks", type=int, default=None)
    ap.add_argument("--checkpoint", default="never", choices=["never", "except_last", "always"])
    ap.add_argument("--precision", default="bf16", choices=["bf16", "tf32", "fp32"])
    ap.add_argument("--layers", type=int, default=None, help="override layer count (debug only; marks result invalid)")
    ap.add_argument("--timeline", action="store_true", help="(kept for compatibility; the bubble is always measured)")
    ap.add_argument("--minimal", action="store_true", help="headline and e2e only (same as PB2_BENCH_MINIMAL=1)")
    return ap.parse_args()


def reference_arm(args) -> int:
    """The reference cannot be installed: it is 2 of ~22 modules of a package torch
    2.11 no longer ships, has no build files, and imports torchtext."""
    why = ("reference tree has no setup.py/pyproject (pip install fails: 'neither setup.py nor pyproject.toml found'); "
           "pipe.py/pipeline.py import 11 sibling modules that are not shipped and torch 2.11 has no "
           "torch.distributed.pipeline; main.py needs torchtext + WikiText-2 (absent, no network)")
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    if os.path.isdir(ref_dir):
        sys.path.insert(0, ref_dir)
        try:
            import importlib
            importlib.import_module("pipe")     # would be the installed reference module
            why = "reference imported but has no runnable entry point without torchtext/WikiText-2"
        except Exception as e:  # noqa: BLE001
            why = f"{why}; import attempt: {type(e).__name__}: {str(e)[:120]}"
    if int(os.environ.get("RANK", "0")) == 0:
        print(json.dumps({"impl": "reference", "unavailable": why}))
    return 0


def _free(*objs) -> None:
    import gc

    import torch
    for o in objs:
        if o is not None and hasattr(o, "close"):
            try:
                o.close()
            except Exception:  # noqa: BLE001
                pass
    gc.collect()
    torch.cuda.empty_cache()


def _store_wait(key: str, rank: int, setter: bool) -> None:
    """Host-side rendezvous through the job's TCP store: no kernel sits on the waiting ranks' GPUs."""
    import torch.distributed as dist
    store = dist.distributed_c10d._get_default_store()
    if setter:
        store.set(key, "1")
    else:
        store.wait([key])


def main() -> int:
    args = parse_args()
    if args.impl == "reference":
        return reference_arm(args)

    import statistics

    import torch
    import torch.distributed as dist

    from pipe_b200.models import CONFIGS
    from pipe_b200.models.trainer import LMTrainer
    from pipe_b200.utils.benchkit import ClockSampler, timed_steps

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    minimal = args.minimal or bool(os.environ.get("PB2_BENCH_MINIMAL"))
    if world > 1:
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        if world != args.gpus:
            raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    else:
        torch.cuda.set_device(0)

    cfg = copy.copy(CONFIGS[args.config])
    if args.chunks:
        cfg.chunks = args.chunks
    if args.layers:
        cfg.nlayers = args.layers

    if args.impl == "fused":
        import pipe_b200.ops as ops
        ops.require_native()            # fail loudly rather than fall back

    trainer = LMTrainer(cfg, n_stages=args.gpus, impl=args.impl, precision=args.precision,
                        checkpoint=args.checkpoint, chunks=cfg.chunks, static_shapes=True)
    dev = trainer.device

    # ---- synthetic data: pinned host batches + device-resident copies ----------
    g = torch.Generator().manual_seed(1234)
    n_host = 4
    host_tok = [torch.randint(0, cfg.ntokens, (cfg.batch, cfg.bptt), generator=g).pin_memory() for _ in range(n_host)]
    host_tgt = [torch.randint(0, cfg.ntokens, (cfg.batch * cfg.bptt,), generator=g).pin_memory() for _ in range(n_host)]

    def feeds(tr, impl):
        # eager arm: targets go to the last stage (reference main.py:217); fused arm: they ride through
        # the pipe next to the tokens so the last stage can fuse the loss behind the vocab projection
        tgt_device = tr.first_device if impl == "fused" else tr.last_device
        feeds_tgt = tr.is_first if impl == "fused" else tr.is_last
        tok = [t.to(tr.first_device) for t in host_tok] if tr.is_first else [None] * n_host
        tgt = [t.to(tgt_device) for t in host_tgt] if feeds_tgt else [None] * n_host
        return tok, tgt, tgt_device, feeds_tgt

    dev_tok, dev_tgt, tgt_device, feeds_tgt = feeds(trainer, args.impl)

    def step_resident(k: int):
        trainer.step(dev_tok[k % n_host], dev_tgt[k % n_host])

    last_loss = [float("nan")]

    def step_e2e(k: int):
        tok = tgt = None
        if trainer.is_first:
            tok = host_tok[k % n_host].to(trainer.first_device, non_blocking=True)
        if feeds_tgt:
            tgt = host_tgt[k % n_host].to(tgt_device, non_blocking=True)
        loss = trainer.step(tok, tgt)
        if trainer.is_last:
            last_loss[0] = float(loss.item())           # device -> host read of the step's result

    for k in range(args.warmup):
        step_resident(k)
    torch.cuda.synchronize()

    launches0 = ops.launch_count() if args.impl == "fused" else 0
    sampler = ClockSampler(local_rank)
    if rank == 0 and not os.environ.get("PB2_BENCH_NO_SAMPLER"):
        sampler.start()
    regions = [timed_steps(step_resident, args.steps, dev) for _ in range(3)]
    clocks = sampler.stop() if rank == 0 else {}
    ms = statistics.median(regions)
    launches = (ops.launch_count() - launches0) // 3 if args.impl == "fused" else 0
    if world > 1:
        t = torch.tensor([launches], dtype=torch.int64, device=dev)
        dist.all_reduce(t)
        launches = int(t.item())

    if os.environ.get("PB2_BENCH_NO_E2E"):
        ms_e2e = ms
    else:
        for k in range(2):
            step_e2e(k)
        ms_e2e = timed_steps(step_e2e, args.steps, dev)
    h2d = host_tok[0].numel() * 8 + host_tgt[0].numel() * 8
    d2h = 4

    tokens = cfg.tokens_per_step
    value = tokens * args.steps / (ms / 1e3)
    e2e = tokens * args.steps / (ms_e2e / 1e3)
    m_actual = len(torch.empty(cfg.batch).chunk(cfg.chunks))
    dtype = "bf16" if args.precision == "bf16" else args.precision

    out = {
        "metric": METRIC, "value": round(value, 2), "unit": "tokens/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms / args.steps, 4),
        "higher_is_better": True, "scaling": "strong",
        # BASELINE.md holds no throughput number for this config (the reference publishes only memory
        # screenshots of another configuration, its traces were stripped): nothing to divide by
        "vs_baseline": None,
        "dtype": dtype, "data": "synthetic", "impl": args.impl,
        "config": {
            "model": f"tutorial TransformerModel emsize {cfg.emsize} d_hid {cfg.nhid} layers {cfg.nlayers} "
                     f"nhead {cfg.nhead} vocab {cfg.ntokens}",
            "global_batch": cfg.batch, "seq_len": cfg.bptt, "chunks": cfg.chunks, "micro_batches": m_actual,
            "checkpoint": args.checkpoint, "parallelism": f"pp{args.gpus}",
            "step": "zero_grad+fwd+CE+bwd+clip_grad_norm(0.5)+Adam",
            "l2": "no flush: per-step working set (weights+grads+Adam state, >3 GB/GPU) >> 126 MB L2",
            "launcher": "DistPipe (rank per stage, batched NCCL p2p, static_shapes)" if world > 1 else "Pipe (single process)",
            "timing": "median of 3 back-to-back regions of `steps` steps each",
        },
        "regions_ms": [round(r, 3) for r in regions],
        "bubble_ideal": round(trainer.bubble, 4),
        "e2e": {"value": round(e2e, 2), "unit": "tokens/s", "ms_per_step": round(ms_e2e / args.steps, 4),
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "last_loss": last_loss[0]},
        "gpu_launches": launches,
        "clocks": clocks,
    }
    if args.layers:
        out["invalid"] = "layer override"
    if world > 1:
        # the loss lives on the last rank; ship it to rank 0 for the record
        t = torch.tensor([last_loss[0] if trainer.is_last else 0.0], dtype=torch.float64, device=dev)
        dist.all_reduce(t)
        out["e2e"]["last_loss"] = float(t.item())

    if args.impl == "fused" and not minimal:
        try:
            extras(args, cfg, trainer, out, step_resident, dev, world, rank, local_rank, feeds, host_tok, host_tgt)
        except Exception as e:  # noqa: BLE001 - the headline must survive a failing extra
            import traceback
            out["extras_error"] = f"{type(e).__name__}: {e}"
            if rank == 0:
                traceback.print_exc()
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def extras(args, cfg, trainer, out, step_resident, dev, world, rank, local_rank, feeds, host_tok, host_tgt) -> None:
    """Measurements reported next to the headline (see the module docstring).  Every rank runs the same
    sequence; a watchdog prints the headline alone if anything here stalls."""
    import threading

    import torch
    import torch.distributed as dist

    from pipe_b200.models.trainer import LMTrainer
    from pipe_b200.utils.benchkit import timed_steps

    budget_s = float(os.environ.get("PB2_BENCH_EXTRAS_BUDGET_S", "420"))

    def bail():
        if rank == 0:
            out["extras_error"] = f"extras exceeded {budget_s:.0f} s; reporting what was measured"
            print(json.dumps(out), flush=True)
        os._exit(0)

    dog = threading.Timer(budget_s, bail)
    dog.daemon = True
    dog.start()
    steps = max(3, min(args.steps, 10))
    n_host = len(host_tok)
    tokens = cfg.tokens_per_step

    # ---- per-step spread: one CUDA-event pair per step on every rank, max over ranks per step -------------------
    n_spread = max(steps, 10)
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n_spread)]
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    for k, (a, b) in enumerate(evs):
        a.record(torch.cuda.current_stream(dev))
        step_resident(k)
        b.record(torch.cuda.current_stream(dev))
    torch.cuda.synchronize()
    per_step = torch.tensor([a.elapsed_time(b) for a, b in evs], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(per_step, op=dist.ReduceOp.MAX)
    q = torch.quantile(per_step.cpu(), torch.tensor([0.1, 0.5, 0.9], dtype=torch.float64))
    out["step_ms_percentiles"] = {"p10": round(float(q[0]), 4), "p50": round(float(q[1]), 4), "p90": round(float(q[2]), 4),
                                  "steps": n_spread, "how": "CUDA events around each step on every rank, max over ranks per step "
                                  "(a rank's events also see its wait for neighbours, so the median sits near ms_per_step)"}

    # ---- measured bubble ---------------------------------------------------------------------
    if world > 1:
        trainer.pipe.timeline = True
        for k in range(3):
            step_resident(k)
        summary = trainer.pipe.timeline_summary()
        trainer.pipe.timeline = False
        if summary:
            out["bubble"] = {
                "ideal": out["bubble_ideal"],
                "measured_per_stage": [round(s["bubble"], 4) for s in summary],
                "measured_mean": round(sum(s["bubble"] for s in summary) / len(summary), 4),
                "fwd_busy_ms": [round(s["fwd_busy_ms"], 3) for s in summary],
                "bwd_busy_ms": [round(s["bwd_busy_ms"], 3) for s in summary],
                "pipe_span_ms": round(summary[0]["pipe_span_ms"], 3),
                "how": "CUDA events around every cell on each rank's stream; bubble = 1 - busy / stage-0 span",
            }
    else:
        out["bubble"] = {"ideal": 0.0, "measured_mean": 0.0, "how": "one stage: no fill / drain"}

    # ---- recompute overhead: checkpoint='always' on the same model ----------------------------------
    base_ms = out["ms_per_step"]
    if args.checkpoint == "never":
        trainer.set_checkpoint("always")
        for k in range(2):
            step_resident(k)
        ms_always = timed_steps(step_resident, steps, dev) / steps
        trainer.set_checkpoint(args.checkpoint)
        peak = torch.cuda.max_memory_allocated(dev) / 2**20
        out["recompute"] = {"always_ms_per_step": round(ms_always, 4),
                            "overhead_vs_never": round(ms_always / base_ms - 1.0, 4),
                            "note": "with 20 rows / 8 chunks 'except_last' checkpoints every micro-batch too (reference quirk)"}
        out["peak_mib_rank0"] = round(peak, 1)

    # ---- the same step on stock torch.nn bodies, same precision ---------------------------------------
    _free(trainer)          # closes the pipe / optimizer and drops the model (the caller's reference stays, hollow)
    if not os.environ.get("PB2_BENCH_NO_EAGER"):
        eager = LMTrainer(cfg, n_stages=args.gpus, impl="eager", precision="bf16", checkpoint=args.checkpoint,
                          chunks=cfg.chunks, static_shapes=True)
        e_tok, e_tgt, _, _ = feeds(eager, "eager")

        def step_eager(k: int):
            eager.step(e_tok[k % n_host], e_tgt[k % n_host])

        for k in range(3):
            step_eager(k)
        ms_eager = timed_steps(step_eager, steps, dev) / steps
        out["eager_same_precision"] = {"ms_per_step": round(ms_eager, 4), "value": round(tokens / (ms_eager / 1e3), 2),
                                       "dtype": "bf16 autocast", "what": "this scheduler, stock torch.nn stage bodies (cuBLAS + ATen)"}
        _free(eager)
        del eager, e_tok, e_tgt
        _free()

    # ---- the in-process Pipe over the same GPUs, driven by rank 0 alone --------------------------------
    if world > 1 and not os.environ.get("PB2_BENCH_NO_PIPE"):
        dist.barrier()
        torch.cuda.synchronize()
        if rank == 0:
            try:
                single = _SingleProcess(cfg, args, host_tok, host_tgt)
                out["pipe_single_process"] = single.run(steps)
            except Exception as e:  # noqa: BLE001
                out["pipe_single_process"] = {"error": f"{type(e).__name__}: {e}"}
            torch.cuda.set_device(local_rank)
        _store_wait("pb2_pipe_single_done", rank, setter=rank == 0)

    # ---- Pipe-inside-DDP hybrid: 2 pipelines x world/2 stages ------------------------------------------
    if world >= 2 and world % 2 == 0 and not os.environ.get("PB2_BENCH_NO_HYBRID"):
        stages = world // 2
        hy = {"layout": f"2 pipelines x {stages} stages", "global_batch": 2 * cfg.batch, "steps": steps}
        tr = None
        for backend in ("nccl", "allreduce", "none", "gather"):
            if tr is None:
                tr = LMTrainer(cfg, n_stages=stages, impl="fused", checkpoint=args.checkpoint, chunks=cfg.chunks,
                               replicas=2, reduce_backend=backend, static_shapes=True)
                h_tok, h_tgt, _, _ = feeds(tr, "fused")
            else:
                tr.opt.close()
                tr.opt = tr.pipe.make_optimizer(reduce_backend=backend, max_tokens=cfg.batch * cfg.bptt, lr=1e-4,
                                                max_norm=0.5)

            def step_h(k: int):
                tr.step(h_tok[k % n_host], h_tgt[k % n_host])

            for k in range(3):
                step_h(k)
            ms_h = timed_steps(step_h, steps, dev) / steps
            hy[f"{backend}_ms"] = round(ms_h, 4)
        hy["tokens_per_s_gather"] = round(2 * tokens / (hy["gather_ms"] / 1e3), 2)
        for backend in ("nccl", "allreduce", "gather"):
            hy[f"exposed_reduce_us_{backend}"] = round((hy[f"{backend}_ms"] - hy["none_ms"]) * 1e3, 1)
        stash = getattr(tr.opt, "_stash", None)
        if stash is not None:
            hy["gather_nvlink_bytes_per_step_per_rank"] = stash.bytes_pushed_per_token * cfg.batch * cfg.bptt
        (st,) = tr.opt.devs.values()
        hy["dense_grad_bytes_per_rank"] = sum(p.numel() for p in st.params) * 4
        hy["what"] = ("nccl = bucketed async dist.all_reduce under the wgrad GEMMs (DDP-style baseline); allreduce = the "
                      "NVLink all-reduce kernel per bucket under the wgrad GEMMs; gather = wgrad operands stored into "
                      "every replica during backward, averaged inside the wgrad GEMM; none = no reduction (timing only)")
        out["hybrid"] = hy
        _free(tr)
    dog.cancel()


class _SingleProcess:
    """The reference-shaped launcher — one process, one worker thread per GPU — on the same config."""

    def __init__(self, cfg, args, host_tok, host_tgt) -> None:
        from pipe_b200.models.trainer import LMTrainer
        self.trainer = LMTrainer(cfg, n_stages=args.gpus, impl="fused", checkpoint=args.checkpoint, chunks=cfg.chunks,
                                 single_process=True)
        self.tok = [t.to(self.trainer.first_device) for t in host_tok]
        self.tgt = [t.to(self.trainer.first_device) for t in host_tgt]
        self.cfg = cfg

    def run(self, steps: int):
        import torch
        tr = self.trainer
        n = len(self.tok)
        devices = list(tr.pipe.devices)

        def sync():
            for d in devices:
                torch.cuda.synchronize(d)

        for k in range(3):
            tr.step(self.tok[k % n], self.tgt[k % n])
        sync()
        t0 = time.perf_counter()
        for k in range(steps):
            tr.step(self.tok[k % n], self.tgt[k % n])
        sync()
        ms = (time.perf_counter() - t0) * 1e3 / steps
        res = {"ms_per_step": round(ms, 4), "value": round(self.cfg.tokens_per_step / (ms / 1e3), 2),
               "launcher": "Pipe (one process, worker thread per GPU, peer copies on copy streams)",
               "timing": "host clock between full device synchronisations (the other ranks idle on the host)"}
        _free(tr)
        return res


if __name__ == "__main__":
    sys.exit(main())
