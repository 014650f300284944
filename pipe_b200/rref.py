"""What ``Pipe.forward`` returns: an RRef when RPC is up, else a local stand-in.

The reference's contract is ``forward(*inputs) -> RRef`` (``pipe.py:431-435,
491-494``), kept "for future inter-node pipelining".  A real
``torch.distributed.rpc.RRef`` can only be constructed after ``rpc.init_rpc``
(otherwise: ``RuntimeError: Current RPC agent is not set!``) — which is why the
reference's author patched the RRef out (``pipe.py:491-494``, ``main.py:210-214``).
Here the same call works both ways: with an RPC agent the result is a genuine
``RRef``; without one it is a :class:`LocalRRef` that answers the same three
questions (``local_value``, ``to_here``, ``is_owner``) and otherwise behaves
like the value it wraps, so ``model(x).local_value()`` and the patched
``model(x)`` style both run unchanged.

Parity: reference ``pipe.py:13,295-302,317-323,431-435,481,491-494``;
SURVEY.md A10.
"""
from __future__ import annotations

from typing import Any

__all__ = ["LocalRRef", "make_rref", "rpc_is_initialized"]


def rpc_is_initialized() -> bool:
    try:
        from torch.distributed import rpc
        return bool(rpc._is_current_rpc_agent_set())
    except Exception:
        return False


class LocalRRef:
    """Owner-side remote reference to a value that never left this process."""

    __slots__ = ("_value",)

    def __init__(self, value: Any) -> None:
        self._value = value

    # -- the RRef surface used with Pipe -------------------------------------
    def local_value(self) -> Any:
        return self._value

    def to_here(self, timeout: float = -1.0) -> Any:
        return self._value

    def is_owner(self) -> bool:
        return True

    def confirmed_by_owner(self) -> bool:
        return True

    def owner_name(self) -> str:
        return "local"

    # -- transparent use as the wrapped value (the reference author's style) --
    def __getattr__(self, name: str) -> Any:
        return getattr(object.__getattribute__(self, "_value"), name)

    def __iter__(self):
        return iter(self._value)

    def __len__(self) -> int:
        return len(self._value)

    def __getitem__(self, item):
        return self._value[item]

    def __repr__(self) -> str:
        return f"LocalRRef({self._value!r})"


def make_rref(value: Any):
    if rpc_is_initialized():
        from torch.distributed.rpc import RRef
        return RRef(value)
    return LocalRRef(value)
