"""distributed_ml_pytorch_b200 — a Blackwell-native asynchronous/synchronous SGD framework with the
capabilities of bkpcoding/distributed_ML_pytorch (see SURVEY.md / DESIGN.md)."""
__version__ = "0.1.0"
