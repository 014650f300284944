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
    class do not collide."""

    __slots__ = ("_id",)

    def __init__(self) -> None:
        self._id = uuid.uuid4()

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
