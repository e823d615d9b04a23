"""Stub `matplotlib` (oracle shim): the reference imports pyplot at module import time only."""
from . import pyplot  # noqa: F401
