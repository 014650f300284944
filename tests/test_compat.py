"""`import pipe_b200.compat` keeps code written against the reference's import paths working
(torch >= 2.4 removed torch.distributed.pipeline).  Runs in a subprocess so the aliases do not leak
into the other tests' sys.modules.  Reference names: pipe.py:24 (__all__), main.py:12, pipeline.py:21."""
import subprocess
import sys
import textwrap


def test_reference_import_paths_resolve():
    code = textwrap.dedent("""
        import torch
        from torch import nn
        import pipe_b200.compat
        from torch.distributed.pipeline.sync import Pipe, NoChunk, WithDevice, is_checkpointing, is_recomputing
        from torch.distributed.pipeline.sync.pipe import Pipe as P2, BalanceError, PipeSequential
        from torch.distributed.pipeline.sync.skip import skippable, stash, pop, verify_skippables
        from torch.distributed.pipeline.sync.skip.tracker import SkipTrackerThroughPotals
        from torch.distributed.pipeline.sync.pipeline import Pipeline, _clock_cycles
        from torch.distributed.pipeline.sync.microbatch import Batch, scatter, gather
        from torch.distributed.pipeline.sync.checkpoint import Checkpointing
        from torch.distributed.pipeline.sync.batchnorm import DeferredBatchNorm
        from torch.distributed.pipeline.sync.utils import partition_model
        import pipe_b200
        assert Pipe is pipe_b200.Pipe and P2 is Pipe
        assert list(_clock_cycles(2, 2)) == [[(0, 0)], [(1, 0), (0, 1)], [(1, 1)]]
        model = Pipe(nn.Sequential(nn.Linear(4, 4), nn.ReLU(), nn.Linear(4, 2)), chunks=2, checkpoint="always")
        out = model(torch.rand(6, 4)).local_value()
        out.sum().backward()
        assert out.shape == (6, 2)
        print("ok")
    """)
    res = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert res.returncode == 0 and res.stdout.strip().endswith("ok"), res.stderr[-2000:]
