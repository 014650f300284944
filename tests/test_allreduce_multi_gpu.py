"""Fused NVSwitch all-reduce vs NCCL, and the hybrid (replicas x stages) step — needs >= 2 GPUs.
One process per GPU (never two ranks on one GPU: the kernels wait on each other)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _need(n):
    if torch.cuda.device_count() < n:
        pytest.skip(f"needs {n} GPUs")


def _spawn(fn, world, *args):
    port = _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    procs = [ctx.Process(target=_entry, args=(fn, r, world, port, q, args)) for r in range(world)]
    for p in procs:
        p.start()
    out = [q.get() for _ in range(world)]
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    bad = [o for o in out if o[1] is not None]
    assert not bad, bad
    return out


def _entry(fn, rank, world, port, q, args):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    err = None
    try:
        fn(rank, world, *args)
    except Exception as e:  # noqa: BLE001
        import traceback
        err = traceback.format_exc()
    q.put((rank, err))
    dist.destroy_process_group()


def _allreduce_case(rank, world):
    from pipe_b200.parallel import SymmetricBuffer
    dev = torch.device("cuda", rank)
    for numel in (1 << 10, 1000 * 1000 + 12, 64 * 1024 * 1024):
        buf = SymmetricBuffer(numel, None, dev)
        assert buf.mode in ("symm", "ipc"), getattr(buf, "_symm_error", "")
        modes = [False] + ([True] if buf.multicast_ptr else [])
        for use_mm in modes:
            for it in range(3):                       # repeated calls: epochs, no pad resets
                torch.manual_seed(1000 * it + rank)
                src = torch.randn(buf.numel, device=dev)
                buf.data.copy_(src)
                want = src.clone()
                dist.all_reduce(want)
                want.mul_(1.0 / world)
                ss = torch.zeros((), device=dev)
                cast = torch.empty(buf.numel, dtype=torch.bfloat16, device=dev)
                torch.cuda.synchronize()
                dist.barrier()
                buf.all_reduce_(scale=1.0 / world, sumsq=ss, cast_out=cast, use_multimem=use_mm)
                torch.cuda.synchronize()
                assert torch.allclose(buf.data, want, rtol=1e-5, atol=1e-5), (numel, use_mm, it)
                assert torch.allclose(cast.float(), want, rtol=1e-2, atol=1e-2)
                dist.all_reduce(ss)
                ref = (want.double() ** 2).sum().item()
                assert abs(ss.item() - ref) / ref < 1e-3
        dist.barrier()
    # buckets: successive calls over different, unevenly sized ranges of one buffer (what the bucketed
    # hybrid reduction does), with the bf16 cast-out whose epilogue must only read what has landed
    buf = SymmetricBuffer(3 * 1000 * 1000 + 96, None, dev)
    g = 4 * world
    edges = [0, 5 * g, 5 * g + 1000 * g, buf.numel // g // 2 * g, buf.numel]
    for it in range(3):
        torch.manual_seed(77 * it + rank)
        src = torch.randn(buf.numel, device=dev)
        buf.data.copy_(src)
        want = src.clone()
        dist.all_reduce(want)
        want.mul_(1.0 / world)
        cast = torch.zeros(buf.numel, dtype=torch.bfloat16, device=dev)
        torch.cuda.synchronize()
        dist.barrier()
        for lo, hi in zip(edges[:-1], edges[1:]):
            buf.all_reduce_(scale=1.0 / world, cast_out=cast[lo:hi], ctas=16 if hi - lo < 100000 else 48, lo=lo, hi=hi)
        torch.cuda.synchronize()
        assert torch.allclose(buf.data, want, rtol=1e-5, atol=1e-5), it
        assert torch.allclose(cast.float(), want, rtol=1e-2, atol=1e-2), it
    dist.barrier()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_fused_allreduce_matches_nccl(world):
    _need(world)
    _spawn(_allreduce_case, world)


def _hybrid_case(rank, world, stages):
    """k replicas x n stages: every reduction backend must produce the same averaged gradients, and the
    replicas must stay bit-identical after optimizer steps."""
    from pipe_b200.models import TransformerConfig, build_stage
    from pipe_b200.parallel import HybridPipe
    dev = torch.device("cuda", rank)
    cfg = TransformerConfig(ntokens=512, emsize=256, nhid=512, nlayers=4, nhead=4, dropout=0.0, bptt=16, batch=8, chunks=2)
    replicas = world // stages
    stage_idx = rank % stages
    g = torch.Generator().manual_seed(7)
    tokens = torch.randint(0, cfg.ntokens, (cfg.batch * replicas, cfg.bptt), generator=g)
    targets = torch.randint(0, cfg.ntokens, (cfg.batch * replicas, cfg.bptt), generator=g)
    grads = {}
    pipe = None
    for backend in ("nccl", "allreduce", "gather"):
        body = build_stage(cfg, stage_idx, stages, dev, impl="fused", seed=1)
        if pipe is None:
            pipe = HybridPipe(body, stages=stages, chunks=cfg.chunks, checkpoint="except_last", device=dev)
        else:
            pipe.stage = type(pipe.stage)(*body.children())         # same groups, fresh weights
            body = pipe.stage
        opt = pipe.make_optimizer(reduce_backend=backend, max_tokens=cfg.batch * cfg.bptt, lr=1e-2, max_norm=0.5)
        lo = pipe.replica * cfg.batch
        for step in range(3):
            opt.zero_grad()
            if pipe.is_first:
                out = pipe(tokens[lo:lo + cfg.batch].to(dev), targets[lo:lo + cfg.batch].to(dev))
            else:
                out = pipe()
            loss = out.local_value().sum() / (cfg.batch * cfg.bptt) if pipe.is_last else None
            pipe.backward(loss)
            if step == 0:
                opt._join_reduction()
                torch.cuda.synchronize()
                k = 1.0 / replicas if backend == "nccl" else 1.0
                grads[backend] = torch.cat([(p.grad * k).flatten() for p in body.parameters()])
            opt.step()
        torch.cuda.synchronize()
        # replicas must stay bit-identical after the updates (same averaged gradient everywhere)
        flat = torch.cat([p.detach().flatten() for p in body.parameters()])
        gathered = [torch.empty_like(flat) for _ in range(replicas)]
        dist.all_gather(gathered, flat, group=pipe.replica_group)
        for other in gathered[1:]:
            assert torch.equal(gathered[0], other), backend
        assert torch.isfinite(flat).all()
        opt.close()
        dist.barrier()
    ref = grads["nccl"]
    for backend in ("allreduce", "gather"):
        err = ((grads[backend] - ref).norm() / ref.norm()).item()
        assert err < 2e-3, (backend, err)      # same bf16 operands, fp32 sums in a different order


@pytest.mark.parametrize("world,stages", [(2, 1), (4, 2), (8, 4)])
def test_hybrid_backends_agree_and_replicas_stay_in_sync(world, stages):
    _need(world)
    _spawn(_hybrid_case, world, stages)


def _mailbox_case(rank, world, mode):
    """The NVLink mailbox transport must train like the NCCL transport (same losses, same weights)."""
    from pipe_b200.models import TransformerConfig, build_stage
    from pipe_b200.ops.optim import FusedClipAdam
    from pipe_b200.parallel import DistPipe
    dev = torch.device("cuda", rank)
    cfg = TransformerConfig(ntokens=512, emsize=256, nhid=512, nlayers=4, nhead=4, dropout=0.1, bptt=16, batch=10, chunks=4)
    g = torch.Generator().manual_seed(11)
    tokens = torch.randint(0, cfg.ntokens, (cfg.batch, cfg.bptt), generator=g).to(dev)
    targets = torch.randint(0, cfg.ntokens, (cfg.batch, cfg.bptt), generator=g).to(dev)
    results = {}
    for transport in ("nccl", "mailbox"):
        torch.manual_seed(5)
        body = build_stage(cfg, rank, world, dev, impl="fused", seed=1)
        pipe = DistPipe(body, chunks=cfg.chunks, checkpoint=mode, device=dev, transport=transport)
        opt = FusedClipAdam(body, lr=1e-2, max_norm=0.5, distributed=True)
        losses, grads = [], []
        for step in range(4):
            opt.zero_grad()
            out = pipe(tokens, targets) if pipe.is_first else pipe()
            loss = out.local_value().sum() / (cfg.batch * cfg.bptt) if pipe.is_last else None
            pipe.backward(loss)
            if step in (0, 1):       # step 0 runs over the process group (handshake), step 1 over the mailboxes
                torch.cuda.synchronize()
                grads.append(torch.cat([p.grad.flatten() for p in body.parameters()]).clone())
                if step == 1:
                    break            # weights after an Adam step differ by sign flips of near-zero gradients
            opt.step()
            if pipe.is_last:
                losses.append(loss.item())
        torch.cuda.synchronize()
        if transport == "mailbox":
            assert pipe._mb is not None, getattr(pipe, "_mb_error", "mailboxes were not built")
        results[transport] = (losses, grads)
        opt.close()
        dist.barrier()
    # same seeds, same masks: only fp32 atomics (LayerNorm / embedding gradients) reorder sums
    for a, b in zip(results["nccl"][0], results["mailbox"][0]):
        assert abs(a - b) <= 1e-4 * abs(a), (results["nccl"][0], results["mailbox"][0])
    g_nccl, g_mb = results["nccl"][1][0], results["mailbox"][1][0]
    err = ((g_nccl - g_mb).norm() / g_nccl.norm()).item()
    assert err < 1e-4, err            # step 0: identical transport
    for g in results["mailbox"][1]:
        assert torch.isfinite(g).all()
    # step 1 starts from weights that went through one Adam step each; compare direction, not bits
    a, b = results["nccl"][1][1], results["mailbox"][1][1]
    cos = torch.nn.functional.cosine_similarity(a, b, dim=0).item()
    assert cos > 0.999, cos


@pytest.mark.parametrize("world,mode", [(2, "never"), (2, "except_last"), (4, "always")])
def test_mailbox_transport_matches_nccl(world, mode):
    _need(world)
    _spawn(_mailbox_case, world, mode)


# ---------------------------------------------------------------------------------------------------
# stress of the cross-GPU flag protocols (stands in for compute-sanitizer's racecheck, which is closed on
# the pool): many back-to-back collectives of random sizes under concurrent compute, every result checked
# ---------------------------------------------------------------------------------------------------
def _allreduce_stress(rank, world):
    import random

    from pipe_b200.parallel import SymmetricBuffer
    dev = torch.device("cuda", rank)
    g = 4 * world
    buf = SymmetricBuffer(600 * 1000 // g * g, None, dev)
    rnd = random.Random(1234)                         # same sequence on every rank
    noise = torch.cuda.Stream(dev)
    a = torch.randn(2048, 2048, device=dev, dtype=torch.bfloat16)
    errors = torch.zeros((), dtype=torch.int64, device=dev)
    base = float(sum(range(1, world + 1)))
    iters = 1200
    for it in range(iters):
        if it % 8 == 0:
            with torch.cuda.stream(noise):            # tensor-core work competing for SMs and HBM
                a = (a @ a).clamp_(-1, 1)
        lo = rnd.randrange(0, buf.numel // g // 2) * g
        hi = lo + rnd.randrange(1, (buf.numel - lo) // g + 1) * g
        ctas = rnd.choice([1, 3, 16, 48, 64])
        use_mm = bool(buf.multicast_ptr) and rnd.random() < 0.5
        bias = float(it % 7)
        buf.data[lo:hi].fill_(rank + 1 + bias)
        buf.all_reduce_(scale=0.5, ctas=ctas, use_multimem=use_mm, lo=lo, hi=hi)
        errors += (buf.data[lo:hi] != 0.5 * (base + world * bias)).sum()
    torch.cuda.synchronize()
    assert int(errors.item()) == 0, f"{int(errors.item())} wrong elements over {iters} collectives"
    dist.barrier()


@pytest.mark.parametrize("world", [2, 8])
def test_allreduce_flag_protocol_stress(world):
    _need(world)
    _spawn(_allreduce_stress, world)


def _mailbox_stress(rank, world):
    from pipe_b200.parallel.mailbox import PeerMailbox
    dev = torch.device("cuda", rank)
    pair = dist.new_group([0, 1])
    sizes = [16, 4096, 65536 * 3 + 16, 1 << 20]
    layout = {slot: [((n // 4,), torch.float32)] for slot, n in enumerate(sizes)}
    box = PeerMailbox(pair, dev, sender_rank=0, layout=layout)
    noise = torch.cuda.Stream(dev)
    a = torch.randn(2048, 2048, device=dev, dtype=torch.bfloat16)
    errors = torch.zeros((), dtype=torch.int64, device=dev)
    rounds = 300
    for seq in range(1, rounds + 1):
        if seq % 8 == 0:
            with torch.cuda.stream(noise):
                a = (a @ a).clamp_(-1, 1)
        for slot, n in enumerate(sizes):                   # 4 x 300 = 1200 messages
            if rank == 0:
                src = torch.full((n // 4,), float(seq * 10 + slot), device=dev)
                box.send(slot, 0, src, seq)
            else:
                dst = torch.empty(n // 4, device=dev)
                box.recv(slot, 0, dst, seq)
                errors += (dst != float(seq * 10 + slot)).sum()
    torch.cuda.synchronize()
    assert int(errors.item()) == 0
    dist.barrier()


def test_mailbox_flag_protocol_stress():
    _need(2)
    _spawn(_mailbox_stress, 2)


def _peer_store_barrier_case(rank, world):
    """The ordering the replicated stashes rely on: kernel A stores into every peer's memory, a LATER kernel (the
    group barrier) exchanges flags, a kernel after that reads what the peers stored — no host synchronisation in
    between, many rounds back to back."""
    from pipe_b200.ops import functional as F2
    from pipe_b200.parallel import SymmetricBuffer
    dev = torch.device("cuda", rank)
    R, C = 64, 256                                  # each rank contributes a [C, R] block (transposed [R, C] source)
    buf = SymmetricBuffer(world * C * R // 2, None, dev)     # bf16 [C, world * R]
    slab = buf.data.view(torch.bfloat16)[: C * world * R].view(C, world * R)
    me = buf.peer_data[rank]
    deltas = [p - me for q, p in enumerate(buf.peer_data) if q != rank]
    errors = torch.zeros((), dtype=torch.int64, device=dev)
    for it in range(300):
        src = torch.full((R, C), float((it % 50) + rank * 64), device=dev).bfloat16()
        F2.transpose_multi([(src, slab, rank * R, None, 1.0)], deltas)
        buf.barrier(phase=0)
        for q in range(world):
            errors += (slab[:, q * R:(q + 1) * R].float() != float((it % 50) + q * 64)).sum()
        buf.barrier(phase=1)                        # nobody overwrites before everybody has read
    torch.cuda.synchronize()
    assert int(errors.item()) == 0, int(errors.item())
    dist.barrier()


@pytest.mark.parametrize("world", [2, 8])
def test_peer_stores_are_visible_after_group_barrier(world):
    _need(world)
    _spawn(_peer_store_barrier_case, world)


def _embed_stash_case(rank, world):
    """EmbedStash alone: every replica packs its own rows, after the group barrier every replica must hold the same,
    complete gradient (= dense scatter-add of all replicas' rows / world), round after round without host syncs."""
    from pipe_b200.ops import functional as F2
    from pipe_b200.parallel import SymmetricBuffer
    from pipe_b200.parallel.replica_stash import EmbedStash
    dev = torch.device("cuda", rank)
    V, E, N, S = 300, 256, 4, 16
    weight = torch.nn.Parameter(torch.zeros(V, E, device=dev))
    weight.grad = torch.zeros(V, E, device=dev)
    sync = SymmetricBuffer(8, None, dev)                    # its pad carries the barriers
    stash = EmbedStash(weight, dist.group.WORLD, dev, max_tokens=2 * N * S)
    bad = torch.zeros((), dtype=torch.int64, device=dev)
    for it in range(40):
        g = torch.Generator().manual_seed(1000 + it)
        all_tok = torch.randint(0, V, (world, 2, N, S), generator=g).to(dev)
        all_dout = torch.randn(world, 2, S, N, E, generator=g).to(dev).bfloat16()
        weight.grad.zero_()
        for mb in range(2):
            stash.pack(all_tok[rank, mb].contiguous(), all_dout[rank, mb].contiguous(), 2.0, 0.0, 0, 0)
        sync.barrier(phase=0)
        stash.scatter(torch.cuda.current_stream(dev).cuda_stream)
        sync.barrier(phase=1)
        want = torch.zeros(V, E, device=dev)
        for q in range(world):
            for mb in range(2):
                F2.embed_bwd(all_tok[q, mb].contiguous(), all_dout[q, mb].contiguous(), want, 2.0, 0.0, 0, 0)
        bad += (~torch.isclose(weight.grad, want / world, rtol=1e-4, atol=1e-4)).sum()
    torch.cuda.synchronize()
    assert int(bad.item()) == 0, int(bad.item())
    dist.barrier()


@pytest.mark.parametrize("world", [2, 4])
def test_embedding_rows_exchange_gives_every_replica_the_full_gradient(world):
    _need(world)
    _spawn(_embed_stash_case, world)


def _nccl_skip_case(rank, world):
    """DistPipe over NCCL: a @skippable stash on the first rank popped on the last one (direct route over the
    pipeline group), in all three checkpoint modes, a short last batch, and an eval pass after training."""
    from torch import nn

    from pipe_b200.parallel import DistPipe
    from pipe_b200.skip import Namespace, pop, skippable, stash
    dev = torch.device("cuda", rank)

    @skippable(stash=["res"])
    class Stash(nn.Module):
        def __init__(self):
            super().__init__()
            self.lin = nn.Linear(16, 16)

        def forward(self, x):
            y = self.lin(x)
            yield stash("res", y)
            return torch.tanh(y)

    @skippable(pop=["res"])
    class Pop(nn.Module):
        def __init__(self):
            super().__init__()
            self.lin = nn.Linear(16, 16)

        def forward(self, x):
            res = yield pop("res")
            return self.lin(x) + 2.0 * res

    def build():
        torch.manual_seed(1)
        ns = Namespace(key="blk")
        layers = [Stash().isolate(ns)] + [nn.Sequential(nn.Linear(16, 16), nn.Tanh()) for _ in range(world - 2)] + \
                 [nn.Sequential(Pop().isolate(ns), nn.Linear(16, 4))]
        if world == 2:
            layers = [nn.Sequential(layers[0]), layers[-1]]
        else:
            layers = [nn.Sequential(layers[0])] + layers[1:]
        return layers

    for mode in ("never", "always", "except_last"):
        stages = build()
        mine = stages[rank].to(dev)
        pipe = DistPipe(mine, chunks=3, checkpoint=mode, device=dev)
        full = nn.Sequential(*[l for st in build() for l in st]).to(dev)
        torch.manual_seed(7)
        for rows in (6, 6, 4):                               # the last batch is shorter
            x = torch.rand(rows, 16, device=dev)
            for p in mine.parameters():
                p.grad = None
            out = pipe(x) if pipe.is_first else pipe()
            pipe.backward(out.local_value().pow(2).sum() if pipe.is_last else None)
        full.zero_grad()
        ref = full(x)
        ref.pow(2).sum().backward()
        if pipe.is_last:
            assert torch.allclose(out.local_value(), ref, atol=1e-5), mode
        ref_params = dict(full.named_parameters())
        n_before = sum(len(list(s.parameters())) for s in stages[:rank])
        for a, b in zip(mine.parameters(), list(full.parameters())[n_before:]):
            assert torch.allclose(a.grad, b.grad, atol=1e-4), (mode, rank)
        pipe.eval()
        with torch.no_grad():
            out = pipe(x) if pipe.is_first else pipe()
        if pipe.is_last:
            assert torch.allclose(out.local_value(), full(x), atol=1e-5)
        dist.barrier()


@pytest.mark.parametrize("world", [2, 4])
def test_dist_pipe_skip_and_changing_batches_over_nccl(world):
    _need(world)
    _spawn(_nccl_skip_case, world)
