"""Failure detection, fault injection and debug aids."""
from . import debug, watchdog  # noqa: F401
