"""Drop-in module aliases for code written against the reference.

    import pipe_b200.compat            # once, before the old imports
    from torch.distributed.pipeline.sync import Pipe
    from torch.distributed.pipeline.sync.skip import skippable, stash, pop

torch >= 2.4 removed ``torch.distributed.pipeline``; importing this module
registers pipe_b200's modules under those names (it never shadows a real one).
"""
from __future__ import annotations

import importlib
import sys
import types

_BASE = "torch.distributed.pipeline"

_ALIASES = {
    f"{_BASE}.sync": "pipe_b200",
    f"{_BASE}.sync.pipe": "pipe_b200.api",
    f"{_BASE}.sync.pipeline": "pipe_b200.runtime.executor",
    f"{_BASE}.sync.microbatch": "pipe_b200.microbatch",
    f"{_BASE}.sync.stream": "pipe_b200.autograd.stream",
    f"{_BASE}.sync.phony": "pipe_b200.autograd.phony",
    f"{_BASE}.sync.dependency": "pipe_b200.autograd.dependency",
    f"{_BASE}.sync.copy": "pipe_b200.autograd.copy",
    f"{_BASE}.sync.checkpoint": "pipe_b200.autograd.checkpoint",
    f"{_BASE}.sync.worker": "pipe_b200.runtime.workers",
    f"{_BASE}.sync.batchnorm": "pipe_b200.batchnorm",
    f"{_BASE}.sync.utils": "pipe_b200.utils",
    f"{_BASE}.sync._balance": "pipe_b200.balance",
    f"{_BASE}.sync.balance": "pipe_b200.balance",
    f"{_BASE}.sync.skip": "pipe_b200.skip",
    f"{_BASE}.sync.skip.skippable": "pipe_b200.skip.skippable",
    f"{_BASE}.sync.skip.layout": "pipe_b200.skip.layout",
    f"{_BASE}.sync.skip.tracker": "pipe_b200.skip.tracker",
    f"{_BASE}.sync.skip.portal": "pipe_b200.skip.portal",
    f"{_BASE}.sync.skip.namespace": "pipe_b200.skip.namespace",
}


def install() -> None:
    import torch.distributed  # noqa: F401
    try:
        importlib.import_module(f"{_BASE}.sync")
        return  # a genuine implementation exists; leave it alone
    except Exception:
        pass
    if _BASE not in sys.modules:
        pkg = types.ModuleType(_BASE)
        pkg.__path__ = []  # type: ignore[attr-defined]
        sys.modules[_BASE] = pkg
        setattr(sys.modules["torch.distributed"], "pipeline", pkg)
    for alias, target in _ALIASES.items():
        mod = importlib.import_module(target)
        sys.modules[alias] = mod
        parent, _, leaf = alias.rpartition(".")
        if parent in sys.modules:
            try:
                setattr(sys.modules[parent], leaf, mod)
            except Exception:
                pass


install()
