"""Reference asgd/optim/__init__.py:1 exports ``DownpourSGD`` (from a file that does not exist); example/main.py:20
imports ``Asynchronous``.  Both names are exported here and are the same class."""
from distributed_ml_pytorch_b200.parallel.async_optim import Asynchronous, DownpourSGD, Listener
from distributed_ml_pytorch_b200.parallel.fused_sgd import FusedSGD
from distributed_ml_pytorch_b200.parallel.sync_optim import Synchronous

__all__ = ["Asynchronous", "DownpourSGD", "Listener", "Synchronous", "FusedSGD"]

# The reference keeps the optimizer in asgd/optim/Asynchronous.py, i.e. ``from asgd.optim.Asynchronous import Asynchronous,
# Listener`` is a valid import there.  A real submodule of that name would shadow the class exported above, so the module path is
# provided as a sys.modules entry instead.
import sys as _sys
import types as _types

_mod = _types.ModuleType(__name__ + ".Asynchronous", "module path of the reference's asgd/optim/Asynchronous.py")
_mod.Asynchronous, _mod.Listener, _mod.DownpourSGD = Asynchronous, Listener, DownpourSGD
_sys.modules[_mod.__name__] = _mod
