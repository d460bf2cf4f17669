from distributed_ml_pytorch_b200.utils.serialization import (numel, ravel_model_params, unravel_model_grads, unravel_model_params)

__all__ = ["ravel_model_params", "unravel_model_params", "unravel_model_grads", "numel"]
