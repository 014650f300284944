"""Skip connections across pipeline stages (``@skippable`` / ``stash`` / ``pop``)."""
from .layout import SkipLayout, inspect_skip_layout  # noqa: F401
from .namespace import Namespace  # noqa: F401
from .portal import Portal  # noqa: F401
from .skippable import Skippable, pop, skippable, stash, verify_skippables  # noqa: F401
from .tracker import (SkipTracker, SkipTrackerThroughPotals, current_skip_tracker,  # noqa: F401
                      use_skip_tracker)

__all__ = ["skippable", "stash", "pop", "verify_skippables", "Namespace", "Skippable",
           "SkipLayout", "inspect_skip_layout", "SkipTracker", "SkipTrackerThroughPotals",
           "use_skip_tracker", "current_skip_tracker", "Portal"]
