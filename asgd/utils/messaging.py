from distributed_ml_pytorch_b200.utils.messaging import (MessageCode, MessageListener, flush_sends, recv_message, send_message)

__all__ = ["MessageCode", "MessageListener", "send_message", "recv_message", "flush_sends"]
