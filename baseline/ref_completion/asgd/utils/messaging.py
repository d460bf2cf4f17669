"""Minimal stand-in for the reference's ABSENT asgd.utils.messaging (contract: SURVEY.md §0.2 A2, §3.7).

Wire format: one flat fp32 tensor [sender_rank, code, payload...] of fixed length N+2, sent with
dist.isend and received by a thread blocked in dist.recv — the transport the reference hard-codes
(gloo, CPU tensors: example/main.py:165).
"""
import threading
from enum import Enum

import torch
import torch.distributed as dist


class MessageCode(Enum):
    ParameterRequest = 0
    GradientUpdate = 1
    ParameterUpdate = 2
    Shutdown = 3  # harness-only: lets the benchmark stop the (otherwise endless) server loop


_keep = []


def send_message(message_code, payload, dst=0):
    m = torch.cat([torch.tensor([float(dist.get_rank()), float(message_code.value)]), payload.detach().reshape(-1).cpu().float()])
    w = dist.isend(tensor=m, dst=dst)
    _keep.append((w, m))
    if len(_keep) > 8:
        w0, _ = _keep.pop(0)
        w0.wait()


class MessageListener(threading.Thread):
    def __init__(self, model):
        super().__init__(daemon=True)
        self.model = model
        n = sum(p.numel() for p in model.parameters())
        self.m_parameter = torch.zeros(n + 2)
        self.running = True

    def receive(self, sender, message_code, parameter):
        raise NotImplementedError

    def run(self):
        while self.running:
            try:
                dist.recv(tensor=self.m_parameter)
            except Exception:
                return
            code = MessageCode(int(self.m_parameter[1].item()))
            if code == MessageCode.Shutdown:
                self.running = False
                return
            self.receive(int(self.m_parameter[0].item()), code, self.m_parameter[2:])
