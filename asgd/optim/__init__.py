"""Reference asgd/optim/__init__.py:1 exports ``DownpourSGD`` (from a file that does not exist); example/main.py:20
imports ``Asynchronous``.  Both names are exported here and are the same class."""
from distributed_ml_pytorch_b200.parallel.async_optim import Asynchronous, DownpourSGD, Listener
from distributed_ml_pytorch_b200.parallel.fused_sgd import FusedSGD
from distributed_ml_pytorch_b200.parallel.sync_optim import Synchronous

__all__ = ["Asynchronous", "DownpourSGD", "Listener", "Synchronous", "FusedSGD"]
