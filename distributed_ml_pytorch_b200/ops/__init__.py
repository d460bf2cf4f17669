"""Functional ops: sm_100a kernels on CUDA, ``torch.nn.functional`` on CPU (see :mod:`.functional`)."""
from .functional import (batch_norm, conv_block, cross_entropy, dropout, flatten, global_avg_pool, linear, max_pool2d, relu,
                         set_unit_loss_grad)

__all__ = ["conv_block", "linear", "max_pool2d", "relu", "flatten", "dropout", "cross_entropy", "batch_norm", "global_avg_pool",
           "set_unit_loss_grad"]
