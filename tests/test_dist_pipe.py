"""DistPipe over gloo, world_size 2 and 3, on CPU: transparency vs the plain Sequential."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
from torch import nn


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _build(n_stages):
    torch.manual_seed(0)
    layers = [nn.Linear(4, 8), nn.Tanh(), nn.Dropout(0.3), nn.Linear(8, 8), nn.Tanh(), nn.Linear(8, 2)]
    per = len(layers) // n_stages
    return [nn.Sequential(*layers[k * per:(k + 1) * per if k < n_stages - 1 else None]) for k in range(n_stages)]


def _worker(rank, world, port, mode, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from pipe_b200.parallel import DistPipe
        stages = _build(world)
        pipe = DistPipe(stages[rank], chunks=4, checkpoint=mode, device=torch.device("cpu"))
        torch.manual_seed(5)
        x = torch.rand(10, 4)
        for step in range(2):                      # second step reuses cached boundary shapes
            torch.manual_seed(100 + rank)          # per-rank RNG: dropout must replay under recompute
            out = pipe(x) if pipe.is_first else pipe()
            loss = out.local_value().pow(2).sum() if pipe.is_last else None
            pipe.backward(loss)
        assert pipe._mb is None and pipe.transport == "nccl"      # the process group is the default transport
        with pytest.raises(ValueError):
            DistPipe(stages[rank], chunks=4, checkpoint=mode, device=torch.device("cpu"), transport="carrier-pigeon")
        grads = {f"{rank}.{k}": p.grad.clone() for k, p in stages[rank].named_parameters()}
        q.put((rank, grads, out.local_value().detach() if pipe.is_last else None))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,mode", [(2, "never"), (2, "always"), (3, "except_last")])
def test_dist_pipe_matches_itself_across_modes(world, mode):
    """With dropout, 'always' must give the same gradients as 'never' given the same
    seeds (bit-identical recompute), and without it both must match plain autograd."""
    ctx = mp.get_context("spawn")
    results = {}
    for m in ("never", mode):
        q = ctx.SimpleQueue()
        port = _free_port()
        procs = [ctx.Process(target=_worker, args=(r, world, port, m, q)) for r in range(world)]
        for p in procs:
            p.start()
        got = [q.get() for _ in range(world)]
        for p in procs:
            p.join(60)
            assert p.exitcode == 0
        merged = {}
        for _, g, _o in got:
            merged.update(g)
        results[m] = merged
    for k in results["never"]:
        assert torch.allclose(results["never"][k], results[mode][k], atol=1e-6), k
