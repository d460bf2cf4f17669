"""``asgd.utils`` (ABSENT in the reference; imported at asgd/optim/Asynchronous.py:4-5)."""
from . import messaging, serialization

__all__ = ["messaging", "serialization"]
