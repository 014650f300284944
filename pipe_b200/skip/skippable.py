"""``@skippable``: skip connections that may cross pipeline stages.

A module class decorated with ``@skippable(stash=[...], pop=[...])`` writes its
``forward`` as a generator::

    @skippable(stash=['skip'])
    class Stash(nn.Module):
        def forward(self, x):
            yield stash('skip', x)
            return self.body(x)

    @skippable(pop=['skip'])
    class Pop(nn.Module):
        def forward(self, x):
            skip = yield pop('skip')
            return x + skip

The wrapper drives the generator, hands stashed tensors to the micro-batch's
skip tracker and feeds popped tensors back in.  ``verify_skippables`` checks a
whole ``nn.Sequential`` statically: every stash is popped exactly once, later,
and no name is declared twice in one namespace.

Parity: reference ``pipe.py:21,334-336`` (call sites; module absent);
SURVEY.md A7 and Appendix A "skip".
"""
from __future__ import annotations

import inspect
from typing import (Any, Callable, ClassVar, Dict, FrozenSet, Generator, Iterable, List,
                    Optional, Set, Tuple, Type, TypeVar, Union)

from torch import Tensor, nn

from ..microbatch import Batch
from .namespace import Namespace
from .tracker import current_skip_tracker

__all__ = ["skippable", "stash", "pop", "Skippable", "verify_skippables"]

T = TypeVar("T", bound="Skippable")


class stash:
    """Command: ``yield stash(name, tensor)`` files ``tensor`` under ``name``."""

    __slots__ = ("name", "tensor")

    def __init__(self, name: str, tensor: Optional[Tensor]) -> None:
        self.name = name
        self.tensor = tensor


class pop:
    """Command: ``tensor = yield pop(name)`` retrieves what was stashed."""

    __slots__ = ("name",)

    def __init__(self, name: str) -> None:
        self.name = name


def _quoted(names: Iterable[str]) -> str:
    return ", ".join(f"'{n}'" for n in sorted(names))


class Skippable(nn.Module):
    """Base of the classes produced by :func:`skippable`.  Holds the user module
    as ``.module`` and the declared names as class attributes."""

    module_cls: ClassVar[Type[nn.Module]]
    stashable_names: ClassVar[FrozenSet[str]]
    poppable_names: ClassVar[FrozenSet[str]]

    def __init__(self, *args: Any, **kwargs: Any) -> None:
        super().__init__()
        self.module = self.module_cls(*args, **kwargs)  # type: ignore[call-arg]
        self.namespaces: Dict[str, Namespace] = {}

    def __repr__(self) -> str:
        return f"@skippable({self.module})"

    # -- naming ------------------------------------------------------------
    def namespaced(self, name: str) -> Tuple[Namespace, str]:
        return (self.namespaces.get(name), name)  # type: ignore[return-value]

    def stashable(self) -> Iterable[Tuple[Namespace, str]]:
        for name in self.stashable_names:
            yield self.namespaced(name)

    def poppable(self) -> Iterable[Tuple[Namespace, str]]:
        for name in self.poppable_names:
            yield self.namespaced(name)

    def isolate(self: T, ns: Namespace, *, only: Optional[Iterable[str]] = None) -> T:
        """Scope this instance's skip names (all, or ``only`` some) to ``ns``."""
        names = set(self.stashable_names | self.poppable_names) if only is None else set(only)
        for name in names:
            self.namespaces[name] = ns
        return self

    # -- execution ---------------------------------------------------------
    def _drive(self, inputs, popped: Dict[str, Optional[Tensor]],
               stashed: Dict[str, Optional[Tensor]]):
        """Run the user's forward; service stash/pop commands it yields."""
        result = self.module(*inputs) if isinstance(inputs, tuple) else self.module(inputs)
        if not inspect.isgenerator(result):
            return result  # forward had no yield at all

        gen: Generator = result
        try:
            command = next(gen)
            while True:
                if isinstance(command, stash):
                    if command.name not in self.stashable_names:
                        raise RuntimeError(f"'{command.name}' has not been declared as stashable")
                    stashed[command.name] = command.tensor
                    command = next(gen)
                elif isinstance(command, pop):
                    if command.name not in self.poppable_names:
                        raise RuntimeError(f"'{command.name}' has not been declared as poppable")
                    command = gen.send(popped.pop(command.name))
                else:
                    raise TypeError(f"{command!r} is not a command from @skippable")
        except StopIteration as done:
            return done.value

    def forward(self, *inputs: Any):  # type: ignore[override]
        tracker = current_skip_tracker()
        values = inputs[0] if len(inputs) == 1 else inputs

        # Pull in everything this layer may pop (this also wires the portal's
        # autograd edge into the incoming batch).
        batch = Batch(values)
        popped: Dict[str, Optional[Tensor]] = {}
        for ns, name in self.poppable():
            try:
                popped[name] = tracker.load(batch, ns, name)
            except KeyError:
                raise RuntimeError(f"'{name}' has not been stashed") from None
        values = batch.values

        stashed: Dict[str, Optional[Tensor]] = {}
        output = self._drive(values, popped, stashed)

        missing = self.stashable_names - stashed.keys()
        if missing:
            raise RuntimeError(f"{_quoted(missing)} must be stashed but have not")
        if popped:
            raise RuntimeError(f"{_quoted(popped.keys())} must be popped but have not")

        # File the stashes (joins the portal's phony into the outgoing batch).
        batch = Batch(output)
        for ns, name in self.stashable():
            tracker.save(batch, ns, name, stashed[name])
        return batch.values


def skippable(stash: Iterable[str] = (), pop: Iterable[str] = ()) -> Callable[[Type[nn.Module]], Type[Skippable]]:
    """Class decorator declaring which names a module stashes and pops."""
    stash_names = frozenset(stash)
    pop_names = frozenset(pop)

    def wrap(module_cls: Type[nn.Module]) -> Type[Skippable]:
        return type(module_cls.__name__, (Skippable,), {
            "module_cls": module_cls,
            "stashable_names": stash_names,
            "poppable_names": pop_names,
            "__doc__": module_cls.__doc__,
            "__module__": module_cls.__module__,
        })

    return wrap


def _walk(module: nn.Module):
    """Children in execution order, opening nested ``nn.Sequential`` children one
    level — exactly what ``Pipe`` flattens into a stage."""
    for child_name, child in module.named_children():
        if isinstance(child, nn.Sequential):
            for sub_name, sub in child.named_children():
                yield f"{child_name}.{sub_name}", sub
        else:
            yield child_name, child


def verify_skippables(module: nn.Sequential) -> None:
    """Statically check that stashes and pops pair up.

    Raises one ``TypeError`` listing every problem found: a name declared both
    stashable and poppable on one layer, a pop without an earlier stash, a name
    re-declared without a separate namespace, a stash that is never popped.
    """
    stashed: Set[Tuple[Namespace, str]] = set()
    popped: Set[Tuple[Namespace, str]] = set()
    problems: List[str] = []

    for path, layer in _walk(module):
        if not isinstance(layer, Skippable):
            continue

        for name in sorted(layer.stashable_names & layer.poppable_names):
            problems.append(f"'{path}' declared '{name}' both as stashable and as poppable")

        for ns, name in layer.stashable():
            if name in layer.poppable_names:
                continue
            if (ns, name) in stashed:
                problems.append(f"'{path}' redeclared '{name}' as stashable "
                                "but not isolated by namespace")
                continue
            stashed.add((ns, name))

        for ns, name in layer.poppable():
            if name in layer.stashable_names:
                continue
            if (ns, name) in popped:
                problems.append(f"'{path}' redeclared '{name}' as poppable "
                                "but not isolated by namespace")
                continue
            if (ns, name) not in stashed:
                problems.append(f"'{path}' declared '{name}' as poppable but it was not stashed")
                continue
            popped.add((ns, name))

    for ns, name in sorted(stashed - popped, key=lambda k: k[1]):
        problems.append(f"no module declared '{name}' as poppable but stashed")

    if problems:
        raise TypeError("one or more pairs of stash and pop do not match:\n\n"
                        + "\n".join(f"* {p}" for p in problems))
