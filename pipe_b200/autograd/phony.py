"""Zero-element "phony" tensors used as pure autograd dependency carriers.

A phony holds no data; it exists so that one autograd node can be made to
depend on another without moving a single byte.  One phony is cached per
``(device, requires_grad)``.  It is created under the device's *default*
stream, so no side stream is ever recorded against its (empty) allocation.

Parity: reference ``README.md:140-160`` (excerpt of the absent ``phony``
module); SURVEY.md Appendix A "phony".
"""
from __future__ import annotations

import threading
from typing import Dict, Tuple

import torch

from .stream import default_stream, use_stream

__all__ = ["get_phony"]

_cache: Dict[Tuple[torch.device, bool], torch.Tensor] = {}
_lock = threading.Lock()


def get_phony(device: torch.device, *, requires_grad: bool) -> torch.Tensor:
    """Return the cached empty tensor for ``(device, requires_grad)``.

    Callers that feed the phony into an autograd ``Function`` should pass it
    as-is; ones that *return* it from a Function should ``detach()`` it first so
    the cached leaf never acquires a ``grad_fn``.
    """
    device = torch.device(device)
    key = (device, bool(requires_grad))
    phony = _cache.get(key)
    if phony is None:
        with _lock:
            phony = _cache.get(key)
            if phony is None:
                with use_stream(default_stream(device)):
                    phony = torch.empty(0, device=device, requires_grad=requires_grad)
                _cache[key] = phony
    return phony
