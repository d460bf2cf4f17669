"""Minimal stand-in for the reference's ABSENT asgd.server.ParameterServer (contract: SURVEY.md §0.2 A3).

ParameterUpdate sets the global parameters, GradientUpdate adds the (already -lr-scaled) payload,
ParameterRequest is answered with a ParameterUpdate to the sender.  run() is the blocking serve loop the
reference calls on rank 0 (example/main.py:135-138).
"""
import torch

from asgd.utils.messaging import MessageCode, MessageListener, send_message
from asgd.utils.serialization import ravel_model_params


class ParameterServer(MessageListener):
    def __init__(self, model):
        super().__init__(model)
        self.parameter_shard = ravel_model_params(model).clone()
        self.done = 0

    def receive(self, sender, message_code, parameter):
        if message_code == MessageCode.ParameterUpdate:
            self.parameter_shard = parameter.clone()
        elif message_code == MessageCode.GradientUpdate:
            self.parameter_shard.add_(parameter)
        elif message_code == MessageCode.ParameterRequest:
            send_message(MessageCode.ParameterUpdate, self.parameter_shard, dst=sender)
