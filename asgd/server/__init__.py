"""``asgd.server.ParameterServer`` (imported at reference example/main.py:21; the module is ABSENT there)."""
from distributed_ml_pytorch_b200.parallel.server import ParameterServer

__all__ = ["ParameterServer"]
