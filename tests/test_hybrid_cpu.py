"""Host-side logic of the hybrid layout on CPU / gloo: group construction and the torch-fallback
reduction of SymmetricBuffer (world_size 4 = 2 replicas x 2 stages)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from pipe_b200.parallel import SymmetricBuffer, hybrid_groups
        pipe_g, rep_g, replica, stage, replicas = hybrid_groups(stages=2)
        assert (replica, stage, replicas) == (rank // 2, rank % 2, 2)
        assert dist.get_process_group_ranks(pipe_g) == [replica * 2, replica * 2 + 1]
        assert dist.get_process_group_ranks(rep_g) == [stage, stage + 2]
        buf = SymmetricBuffer(10, rep_g, torch.device("cpu"))
        assert buf.mode == "torch" and buf.numel % (4 * 2) == 0
        buf.data.fill_(float(rank + 1))
        ss = torch.zeros(())
        buf.all_reduce_(scale=0.5, sumsq=ss)
        expect = (stage + 1 + stage + 3) * 0.5
        assert torch.allclose(buf.data, torch.full_like(buf.data, expect))
        dist.all_reduce(ss, group=rep_g)
        assert abs(ss.item() - expect * expect * buf.numel) < 1e-4
        q.put((rank, None))
    except Exception:
        import traceback
        q.put((rank, traceback.format_exc()))
    finally:
        dist.destroy_process_group()


def test_hybrid_groups_and_fallback_reduce():
    world = 4
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get() for _ in range(world)]
    for p in procs:
        p.join(60)
    assert all(e is None for _, e in res), res
