"""pipe_b200 — synchronous pipeline parallelism, built for B200.

Public surface (same names as ``torch.distributed.pipeline.sync``):

    Pipe, NoChunk, WithDevice, is_checkpointing, is_recomputing
    skip.skippable / stash / pop / Namespace / verify_skippables
    balance.balance_by_time / balance_by_size,  utils.partition_model

plus what this framework adds: ``balance=/devices=`` partitioning, the cell
timeline (bubble %), state-dict converters, fused sm_100a layers under
``pipe_b200.ops`` / ``pipe_b200.models``, the Pipe-inside-DDP hybrid with a fused
NVSwitch all-reduce under ``pipe_b200.parallel``.

``import pipe_b200.compat`` additionally registers the package under the
reference's module paths, so ``from torch.distributed.pipeline.sync import Pipe``
keeps working on torch builds that no longer ship it.
"""
from .api import Pipe, MOVING_DENIED  # noqa: F401
from .autograd.checkpoint import is_checkpointing, is_recomputing  # noqa: F401
from .microbatch import NoChunk  # noqa: F401
from .partition import BalanceError, PipeSequential, WithDevice  # noqa: F401
from .rref import LocalRRef  # noqa: F401
from .runtime.timeline import Timeline  # noqa: F401

__version__ = "0.1.0"

__all__ = ["Pipe", "NoChunk", "WithDevice", "PipeSequential", "BalanceError", "LocalRRef",
           "Timeline", "is_checkpointing", "is_recomputing", "__version__"]
