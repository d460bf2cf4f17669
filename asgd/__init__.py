"""``asgd`` — the reference's package name (bkpcoding/distributed_ML_pytorch), kept as a thin namespace over
:mod:`distributed_ml_pytorch_b200` so reference user code keeps working unchanged:

    from asgd.optim import Asynchronous          # example/main.py:20
    from asgd.server import ParameterServer      # example/main.py:21
    from asgd.utils.serialization import ravel_model_params, unravel_model_params   # Asynchronous.py:4
    from asgd.utils.messaging import MessageCode, MessageListener, send_message     # Asynchronous.py:5
"""
