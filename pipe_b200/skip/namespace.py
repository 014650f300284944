"""Namespaces that keep skip names of re-used ``@skippable`` classes apart.

Parity: SURVEY.md A8 (``skip.namespace`` is absent from the reference tree; its
use is implied by ``pipe.py:20-21``).
"""
from __future__ import annotations

import abc
import uuid
from functools import total_ordering
from typing import Any

__all__ = ["Namespace"]


@total_ordering
class Namespace(metaclass=abc.ABCMeta):
    """A unique, hashable, orderable token.  ``isolate(ns)`` on a skippable
    module scopes its stash/pop names to ``ns`` so two instances of the same
    class do not collide.

    ``Namespace()`` is unique to the process that created it.  ``Namespace(key="block3")`` is the
    same token in every process that uses the same key: the multi-process pipeline
    (:class:`~pipe_b200.parallel.DistPipe`) needs that to pair a stash on one rank with its pop on
    another, because each rank builds its own module objects."""

    __slots__ = ("_id", "key")

    def __init__(self, key: "str | None" = None) -> None:
        self.key = None if key is None else str(key)
        self._id = uuid.uuid4() if key is None else uuid.uuid5(uuid.NAMESPACE_OID, "pipe_b200.skip/" + self.key)

    def __repr__(self) -> str:
        return f"<Namespace '{self._id}'>"

    def __hash__(self) -> int:
        return hash(self._id)

    def __eq__(self, other: Any) -> bool:
        return isinstance(other, Namespace) and self._id == other._id

    def __lt__(self, other: Any) -> bool:
        # ``None`` (the default, un-isolated scope) sorts before every Namespace.
        if isinstance(other, Namespace):
            return self._id < other._id
        return False


# The default scope is ``None``; registering it makes isinstance checks uniform.
Namespace.register(type(None))
