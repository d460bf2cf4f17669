"""The reference's asgd/optim/__init__.py imports ``.downpour_sgd.DownpourSGD`` but ships Asynchronous.py
(the rename was left half-done).  This shim re-exports the shipped class under both names
without touching the shipped files: it also defines the global ``DownpourSGD`` that
``super(DownpourSGD, self).__init__`` looks up at asgd/optim/Asynchronous.py:40."""
import importlib
import sys

_mod = importlib.import_module("asgd.optim.Asynchronous")
DownpourSGD = _mod.Asynchronous
_mod.DownpourSGD = _mod.Asynchronous
# `from asgd.optim import Asynchronous` (example/main.py:20) must yield the class, not the submodule
sys.modules["asgd.optim"].Asynchronous = _mod.Asynchronous
