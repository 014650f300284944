"""Utilities: ``partition_model`` (re-nest a flat Sequential per device),
benchmark configs, clock sampling, roofline arithmetic."""
from .partition_model import partition_model  # noqa: F401
from . import data  # noqa: F401

__all__ = ["partition_model"]
