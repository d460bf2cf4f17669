"""The module the reference's asgd/optim/__init__.py:1 imports (``from .downpour_sgd import DownpourSGD``) but does not ship."""
from distributed_ml_pytorch_b200.parallel.async_optim import DownpourSGD

__all__ = ["DownpourSGD"]
