"""Functional ops: sm_100a kernels on CUDA, ``torch.nn.functional`` on CPU (see :mod:`.functional`)."""
from .functional import (conv_block, cross_entropy, dropout, flatten, linear, max_pool2d, relu, set_unit_loss_grad)

__all__ = ["conv_block", "linear", "max_pool2d", "relu", "flatten", "dropout", "cross_entropy", "set_unit_loss_grad"]
