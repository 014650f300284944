"""Multi-process / multi-replica parallelism: DistPipe (rank-per-stage pipeline),
the Pipe-inside-DDP hybrid and its fused NVSwitch gradient all-reduce."""
from .dist_pipe import DistPipe  # noqa: F401
from .hybrid import HybridPipe, hybrid_groups  # noqa: F401
from .symmetric import SymmetricBuffer  # noqa: F401
