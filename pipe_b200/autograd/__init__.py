"""Autograd plumbing: streams, phonies, fork/join, Copy/Wait, checkpointing."""
from .checkpoint import (Checkpointing, checkpoint, is_checkpointing,  # noqa: F401
                         is_recomputing)
from .copy import Copy, Wait  # noqa: F401
from .dependency import fork, join  # noqa: F401
from .phony import get_phony  # noqa: F401
from .stream import (CPUStream, AbstractStream, current_stream, default_stream,  # noqa: F401
                     get_device, new_stream, record_stream, use_device, use_stream,
                     wait_stream)
