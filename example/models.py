"""``from models import LeNet, AlexNet`` — the import the reference's example/main.py:15 uses.  The classes are the
kernel-backed, state_dict-compatible models of distributed_ml_pytorch_b200 (reference example/models.py:5-49), plus
the NORTH-STAR ResNets."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from distributed_ml_pytorch_b200.models import AlexNet, LeNet, build_model  # noqa: E402,F401

try:  # ResNets are optional at import time
    from distributed_ml_pytorch_b200.models.resnet import resnet18, resnet50  # noqa: F401
except ImportError:  # pragma: no cover
    pass
